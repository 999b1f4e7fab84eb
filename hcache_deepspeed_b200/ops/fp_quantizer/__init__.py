from .quantize import FP_Quantize, Quantizer, fp_dequantize, fp_quantize  # noqa: F401
