"""Op builders: one per native op of the reference's ``op_builder/`` table. Each knows its translation units, probes
whether it can be built here, and ``load()`` hands back a library that provides it — the pre-built in-tree one, or its
own JIT-compiled object (see :mod:`.builder`). Ops the reference implements natively but this build formulates in
torch (training transformer, sparse attention, ...) keep a builder with no sources: compatible, nothing to build."""
from __future__ import annotations

from typing import Dict


from .builder import (OpBuildError, OpBuilder, find_library_file, has_function, has_header, nvcc_knows_sm100a,  # noqa: F401
                      nvcc_version, object_dependencies)

_GEMM_TUS = ("gemm.cu", "gemm_2sm.cu", "gemm_fp8.cu")
_TP = ("tp_comm.cu", )          # every GEMM / norm launcher consumes the armed flag-exchange state defined there


def _mk(name, lib, symbols, sources=(), depends=(), doc=""):
    return type(name, (OpBuilder, ), {"NAME": name, "LIB": lib, "SYMBOLS": tuple(symbols), "SOURCES": tuple(sources),
                                      "DEPENDS": tuple(depends), "__doc__": doc})


InferenceCoreBuilder = _mk("inference_core_ops", "cuda", ["hds_gemm", "hds_gemm_2sm", "hds_gemm_fp8", "hds_quantize_fp8_rows", "hds_norm"],
                           sources=_GEMM_TUS + ("elementwise.cu", ), depends=_TP)
RaggedOpsBuilder = _mk("ragged_device_ops", "cuda", ["hds_paged_attention_splitkv", "hds_paged_attention_prefill", "hds_embed",
                                                       "hds_rope_kv_append", "hds_gather_rows", "hds_moe_topk_gating",
                                                       "hds_moe_scatter", "hds_moe_gather"],
                       sources=("attention_splitkv.cu", "attention_prefill.cu", "elementwise.cu", "moe.cu", "host_buffer.cu"),
                       depends=_TP)
InferenceCutlassBuilder = _mk("cutlass_ops", "cuda", ["hds_moe_gemm"], sources=("moe.cu", ))
RaggedUtilsBuilder = _mk("ragged_ops", "host", ["hds_alloc_create", "hds_build_batch"], sources=("host_runtime.cpp", ))
InferenceBuilder = _mk("transformer_inference", "cuda", ["hds_gemm", "hds_norm", "hds_rope_kv_append"],
                       sources=_GEMM_TUS + ("elementwise.cu", ), depends=_TP)


class AsyncIOBuilder(OpBuilder):
    """Kernel AIO file engine. Needs the kernel's AIO ABI header (the engine issues io_submit through raw syscalls, so
    no libaio is linked — the reference probes for ``libaio`` here, ``op_builder/async_io.py``)."""
    NAME, LIB = "async_io", "host"
    SOURCES = ("aio.cpp", )
    SYMBOLS = ("hds_aio_create", "hds_aio_pread", "hds_aio_pwrite")

    def extra_problems(self):
        return [] if has_header("linux/aio_abi.h") else ["<linux/aio_abi.h> not found (kernel headers)"]


class GDSBuilder(AsyncIOBuilder):
    """GPUDirect Storage goes through libcufile at run time (``io/gds.py``, ctypes); the bounce path is the AIO engine."""
    NAME = "gds"
    SYMBOLS = ("hds_aio_create", )

    def extra_problems(self):
        why = super().extra_problems()
        if find_library_file("libcufile.so") is None:
            why.append("libcufile.so not found: GDS transfers fall back to the pinned bounce buffer")
        return why


QuantizerBuilder = _mk("quantizer", "cuda", ["hds_quantize_int", "hds_dequantize_int"], sources=("quantize.cu", ))
FPQuantizerBuilder = _mk("fp_quantizer", "cuda", ["hds_fp_quantize", "hds_fp_dequantize"], sources=("quantize.cu", ))
FusedAdamBuilder = _mk("fused_adam", "cuda", ["hds_multi_tensor_adam"], sources=("optimizers.cu", ))
FusedLionBuilder = _mk("fused_lion", "cuda", ["hds_multi_tensor_lion"], sources=("optimizers.cu", ))
FusedLambBuilder = _mk("fused_lamb", "cuda", [])
CPUAdamBuilder = _mk("cpu_adam", "host", ["hds_cpu_adam_step"], sources=("cpu_optimizers.cpp", ))
CPUAdagradBuilder = _mk("cpu_adagrad", "host", ["hds_cpu_adagrad_step"], sources=("cpu_optimizers.cpp", ))
CPULionBuilder = _mk("cpu_lion", "host", ["hds_cpu_lion_step"], sources=("cpu_optimizers.cpp", ))
TransformerBuilder = _mk("transformer", "cuda", [])
StochasticTransformerBuilder = _mk("stochastic_transformer", "cuda", [])
SparseAttnBuilder = _mk("sparse_attn", "cuda", [])
RandomLTDBuilder = _mk("random_ltd", "cuda", [])
SpatialInferenceBuilder = _mk("spatial_inference", "cuda", [])
EvoformerAttnBuilder = _mk("evoformer_attn", "cuda", [])
ShareMemCommBuilder = _mk("deepspeed_shm_comm", "host", ["hds_shm_open", "hds_shm_allreduce", "hds_shm_broadcast", "hds_shm_allgather",
                                                          "hds_shm_reduce_scatter", "hds_shm_barrier", "hds_shm_close"],
                          sources=("shm_comm.cpp", ))
ShareMemCommBuilder.extra_problems = lambda self: [] if has_header("immintrin.h") else ["<immintrin.h> not found (x86-64 only)"]


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
