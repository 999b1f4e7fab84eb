"""Op-builder compatibility surface. The reference JIT-compiles ~25 pybind extensions through ``OpBuilder.load()``
(``op_builder/builder.py``); here every native op lives in two pre-built in-tree libraries, so a "builder" reports
whether its ops are present in them and ``load()`` returns the ctypes handle."""
from __future__ import annotations

from typing import Dict


class OpBuilder:
    NAME = "op"
    LIB = "cuda"            # which in-tree library provides the op: "cuda" (libhds_b200.so) or "host" (libhds_host.so)
    SYMBOLS = ()

    def __init__(self, name=None):
        self.name = name or self.NAME

    def absolute_name(self):
        return f"hcache_deepspeed_b200.ops.{self.name}"

    def sources(self):
        return []

    def is_compatible(self, verbose: bool = False) -> bool:
        import shutil
        return shutil.which("nvcc") is not None or self.LIB == "host" or self.installed()

    def installed(self) -> bool:
        from .. import native
        path = native.LIB_DIR / ("libhds_b200.so" if self.LIB == "cuda" else "libhds_host.so")
        return path.exists()

    def load(self, verbose: bool = True):
        from .. import native
        lib = native.cuda_lib() if self.LIB == "cuda" else native.host_lib()
        missing = [s for s in self.SYMBOLS if not hasattr(lib, s)]
        if missing:
            raise RuntimeError(f"{self.name}: symbols {missing} missing from the in-tree library; rebuild with "
                               "`python -m hcache_deepspeed_b200.ops.build`")
        return lib

    builder = load


def _mk(name, lib, symbols, doc=""):
    return type(name, (OpBuilder, ), {"NAME": name, "LIB": lib, "SYMBOLS": tuple(symbols), "__doc__": doc})


InferenceCoreBuilder = _mk("inference_core_ops", "cuda", ["hds_gemm", "hds_gemm_2sm", "hds_gemm_fp8", "hds_quantize_fp8_rows", "hds_norm"])
RaggedOpsBuilder = _mk("ragged_device_ops", "cuda", ["hds_paged_attention_splitkv", "hds_paged_attention_prefill", "hds_embed",
                                                       "hds_rope_kv_append", "hds_gather_rows", "hds_moe_topk_gating",
                                                       "hds_moe_scatter", "hds_moe_gather"])
InferenceCutlassBuilder = _mk("cutlass_ops", "cuda", ["hds_moe_gemm"])
RaggedUtilsBuilder = _mk("ragged_ops", "host", ["hds_alloc_create", "hds_build_batch"])
InferenceBuilder = _mk("transformer_inference", "cuda", ["hds_gemm", "hds_norm", "hds_rope_kv_append"])
AsyncIOBuilder = _mk("async_io", "host", ["hds_aio_create", "hds_aio_pread", "hds_aio_pwrite"])
GDSBuilder = _mk("gds", "host", ["hds_aio_create"])
QuantizerBuilder = _mk("quantizer", "cuda", ["hds_quantize_int", "hds_dequantize_int"])
FPQuantizerBuilder = _mk("fp_quantizer", "cuda", ["hds_fp_quantize", "hds_fp_dequantize"])
FusedAdamBuilder = _mk("fused_adam", "cuda", ["hds_multi_tensor_adam"])
FusedLionBuilder = _mk("fused_lion", "cuda", ["hds_multi_tensor_lion"])
FusedLambBuilder = _mk("fused_lamb", "cuda", [])
CPUAdamBuilder = _mk("cpu_adam", "host", ["hds_cpu_adam_step"])
CPUAdagradBuilder = _mk("cpu_adagrad", "host", ["hds_cpu_adagrad_step"])
CPULionBuilder = _mk("cpu_lion", "host", ["hds_cpu_lion_step"])
TransformerBuilder = _mk("transformer", "cuda", [])
StochasticTransformerBuilder = _mk("stochastic_transformer", "cuda", [])
SparseAttnBuilder = _mk("sparse_attn", "cuda", [])
RandomLTDBuilder = _mk("random_ltd", "cuda", [])
SpatialInferenceBuilder = _mk("spatial_inference", "cuda", [])
EvoformerAttnBuilder = _mk("evoformer_attn", "cuda", [])
ShareMemCommBuilder = _mk("deepspeed_shm_comm", "host", ["hds_shm_open", "hds_shm_allreduce", "hds_shm_close"])


class UtilsBuilder(OpBuilder):
    """``UtilsBuilder().load()`` exposes ``flatten`` / ``unflatten`` (``csrc/utils/flatten_unflatten.cpp``: thin
    wrappers over torch's dense-tensor flattening, which is what they are here too)."""
    NAME, LIB = "utils", "host"

    def load(self, verbose: bool = True):
        from types import SimpleNamespace
        from torch._utils import _flatten_dense_tensors, _unflatten_dense_tensors
        return SimpleNamespace(flatten=_flatten_dense_tensors, unflatten=_unflatten_dense_tensors)

ALL_OPS: Dict[str, OpBuilder] = {b.NAME: b() for b in (
    InferenceCoreBuilder, RaggedOpsBuilder, InferenceCutlassBuilder, RaggedUtilsBuilder, InferenceBuilder, AsyncIOBuilder,
    GDSBuilder, QuantizerBuilder, FPQuantizerBuilder, FusedAdamBuilder, FusedLionBuilder, FusedLambBuilder, CPUAdamBuilder,
    CPUAdagradBuilder, CPULionBuilder, TransformerBuilder, StochasticTransformerBuilder, SparseAttnBuilder, RandomLTDBuilder,
    SpatialInferenceBuilder, EvoformerAttnBuilder, UtilsBuilder, ShareMemCommBuilder)}


def get_builder_class(class_name: str):
    return globals().get(class_name)


def get_builder(class_name: str):
    cls = get_builder_class(class_name)
    return cls() if cls is not None else None
