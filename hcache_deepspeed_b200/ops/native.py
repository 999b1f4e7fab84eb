"""ctypes bindings of the two in-tree native libraries.

``libhds_b200.so``  — sm_100a CUDA kernels (``include/hds_b200.h``)
``libhds_host.so``  — host runtime: block allocator, ragged batch builder, async file IO
                      (``include/hds_host.h``)

Loading is lazy. On a machine with a GPU a missing kernel library is a hard error (the framework
never silently falls back to eager PyTorch on the device); on a CPU-only machine the CUDA library is
simply not needed.
"""
from __future__ import annotations

import ctypes
import os
from ctypes import c_char_p, c_float, c_int, c_int32, c_int64, c_void_p
from pathlib import Path
from typing import Optional

LIB_DIR = Path(__file__).resolve().parent / "lib"

_cuda_lib: Optional[ctypes.CDLL] = None
_host_lib: Optional[ctypes.CDLL] = None


class NativeLibraryError(RuntimeError):
    pass


class HdsKernelError(RuntimeError):
    pass


def _declare_cuda(lib: ctypes.CDLL) -> None:
    P, I, L, F = c_void_p, c_int, c_int64, c_float
    lib.hds_last_error.restype = c_char_p
    lib.hds_last_error.argtypes = []
    sigs = {
        "hds_gemm": [P, L, P, L, I, I, I, I, I, I, P, P, L, I, I, P, L, P],
        "hds_gemm_accum": [P, L, P, L, I, I, I, I, P, P, I, I, L, I, I, P],
        "hds_restore_kv_proj": [P, L, P, L, I, I, I, P, P, P, P, P, I, I, I, I, P],
        "hds_qkv_rope_proj": [P, L, P, L, I, I, I, P, P, P, P, P, P, I, I, I, I, I, P],
        "hds_restore_kv_proj_peers": [P, I, I, L, P, L, I, I, I, P, P, P, P, P, I, I, I, I, P],
        "hds_norm": [P, L, P, I, I, P, P, L, P, P, P, I, I, F, I, I, I, P],
        "hds_norm_tp": [I, I, I, P, L, P, P, I, P, P, P, L, P, P, I, I, I, P, P, I, F, I, I, P],
        "hds_fp8_next_weight_format": [I],
        "hds_attention_next_alibi": [P],
        "hds_attention_next_window": [I],
        "hds_host_buffer_alloc": [L, I, I, P],
        "hds_host_buffer_device_ptr": [P, P],
        "hds_host_buffer_free": [P],
        "hds_tp_ctx_create": [I, I, P],
        "hds_tp_ctx_destroy": [I],
        "hds_tp_arm": [I, I, I],
        "hds_tp_disarm": [],
        "hds_tp_signal": [I, P],
        "hds_tp_wait": [I, P],
        "hds_tp_barrier": [I, P],
        "hds_tp_gather_cols": [I, I, P, L, P, L, P, P, I, I, P],
        "hds_rope_kv_append": [P, L, P, P, P, P, P, I, I, I, I, I, I, I, I, P],
        "hds_embed": [P, P, P, P, I, I, P, I, I, I, I, P],
        "hds_gather_rows": [P, P, P, I, I, P],
        "hds_argmax": [P, L, I, I, P, P, I, P],
        "hds_kv_blocks_copy": [P, P, P, I, I, L, L, I, P],
        "hds_cvt_f32": [P, P, L, I, I, P],
        "hds_p2p_gather_rows": [P, P, I, I, I, L, I, P],
        "hds_paged_attention_splitkv": [P, P, P, L, P, I, P, I, I, I, I, I, I, I, F, P, L, I, P],
        "hds_paged_attention_prefill": [P, P, P, L, P, I, P, I, I, I, I, I, I, F, I, P, L, I, P],
        "hds_paged_attention_packed": [P, P, P, L, P, I, P, I, I, I, I, I, I, F, I, P, L, I, P],
        "hds_moe_topk_gating": [P, L, I, I, I, I, I, P, P, P, P, P],
        "hds_moe_scatter": [P, L, P, L, P, P, P, P, P, I, I, I, I, P],
        "hds_moe_gemm": [P, L, P, L, P, I, I, I, I, I, I, I, P, L, P],
        "hds_moe_gather": [P, L, P, P, P, L, I, I, I, I, P],
        "hds_gemm_2sm": [P, L, P, L, I, I, I, I, I, I, P, P, L, I, P],
        "hds_quantize_fp8_rows": [P, L, I, I, I, P, L, P, P],
        "hds_gemm_fp8_skinny": [P, L, P, P, L, P, I, I, I, I, I, I, P, P, L, P, L, P, L, I, P],
        "hds_gemm_fp8": [P, L, P, P, L, P, I, I, I, I, I, I, P, P, L, I, P],
        "hds_quantize_int": [P, I, L, I, I, I, P, P, P],
        "hds_dequantize_int": [P, P, L, I, I, P, I, P],
        "hds_fp_quantize": [P, I, L, I, I, P, P, P],
        "hds_fp_dequantize": [P, P, P, L, I, I, P, I, P],
        "hds_multi_tensor_adam": [I, P, P, P, P, P, I, I, F, F, F, F, F, I, I, I, F, P],
        "hds_multi_tensor_lion": [I, P, P, P, P, I, I, F, F, F, F, F, P],
    }
    for name, argtypes in sigs.items():
        fn = getattr(lib, name)
        fn.argtypes = argtypes
        fn.restype = c_int
    lib.hds_paged_attention_splitkv_workspace_bytes.argtypes = [I, I, I, I, I, I]
    lib.hds_paged_attention_splitkv_workspace_bytes.restype = c_int64
    lib.hds_paged_attention_prefill_workspace_bytes.argtypes = [I, I, I, I]
    lib.hds_paged_attention_prefill_workspace_bytes.restype = c_int64
    lib.hds_gemm_splitk_workspace_bytes.argtypes = [I, I]
    lib.hds_gemm_splitk_workspace_bytes.restype = c_int64
    lib.hds_moe_padded_rows.argtypes = [I, I, I]
    lib.hds_moe_padded_rows.restype = c_int64
    # optional entry points (declared when present so older builds keep loading)
    optional = {
        "hds_memset_u32": ([P, c_int32, L, P], c_int),
    }
    for name, (argtypes, restype) in optional.items():
        if hasattr(lib, name):
            fn = getattr(lib, name)
            fn.argtypes = argtypes
            fn.restype = restype


def cuda_lib(build_if_missing: bool = True) -> ctypes.CDLL:
    """Return the loaded sm_100a kernel library (building it in-tree first if it is absent)."""
    global _cuda_lib
    if _cuda_lib is not None:
        return _cuda_lib
    path = LIB_DIR / "libhds_b200.so"
    if not path.exists() and build_if_missing and os.environ.get("HDS_NO_AUTOBUILD") != "1":
        from . import build
        build.build_cuda(verbose=True)
    if not path.exists():
        raise NativeLibraryError(
            f"{path} is missing: run `python -m hcache_deepspeed_b200.ops.build` (needs nvcc). "
            "The GPU path never falls back to eager PyTorch.")
    lib = ctypes.CDLL(str(path), mode=ctypes.RTLD_GLOBAL)
    _declare_cuda(lib)
    _cuda_lib = lib
    return lib


def cuda_lib_loaded() -> bool:
    return _cuda_lib is not None


def check(rc: int, what: str) -> None:
    """Raise with the library's message when an entry point reports failure."""
    if rc != 0:
        msg = cuda_lib().hds_last_error()
        raise HdsKernelError(f"{what} failed (code {rc}): {msg.decode() if msg else 'no message'}")


def host_lib(build_if_missing: bool = True) -> ctypes.CDLL:
    """Return the loaded host runtime library (g++-built, no CUDA dependency)."""
    global _host_lib
    if _host_lib is not None:
        return _host_lib
    path = LIB_DIR / "libhds_host.so"
    if not path.exists() and build_if_missing and os.environ.get("HDS_NO_AUTOBUILD") != "1":
        from . import build
        build.build_host(verbose=True)
    if not path.exists():
        raise NativeLibraryError(f"{path} is missing: run `python -m hcache_deepspeed_b200.ops.build` (needs g++)")
    lib = ctypes.CDLL(str(path))
    from . import host_bindings
    host_bindings.declare(lib)
    _host_lib = lib
    return lib
