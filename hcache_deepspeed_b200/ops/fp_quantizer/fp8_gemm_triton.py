"""The reference implements the weight-only FP8 matmul in Triton here. Triton is not used by this framework; the name
resolves to the same entry point as :mod:`.fp8_gemm`."""
from .fp8_gemm import matmul_fp8_fallback as matmul_fp8  # noqa: F401
