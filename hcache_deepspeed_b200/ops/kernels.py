"""torch-facing wrappers over the sm_100a kernel library (``include/hds_b200.h``).

Every function takes CUDA tensors, launches on the current stream and returns immediately; there
is no hidden allocation unless an ``out=`` argument is omitted. These wrappers are the ONLY way the
framework runs compute on the device — there is no eager-PyTorch fallback behind them (the fp32
PyTorch restatements used as test oracles live in :mod:`hcache_deepspeed_b200.ops.reference`).

Reference counterparts: the ``DSKernelBase`` callables under
``deepspeed/inference/v2/kernels/{core_ops,ragged_ops}/`` (e.g. ``core_ops/blas_kernels/blas.py``,
``ragged_ops/linear_blocked_kv_rotary/blocked_kv_rotary.py``).
"""
from __future__ import annotations

import ctypes
from typing import Optional, Sequence, Tuple

import torch

from . import native

ACT_IDENTITY, ACT_GELU, ACT_RELU, ACT_SILU, ACT_GELU_TANH = 0, 1, 2, 3, 4
ACT_CODES = {"identity": 0, None: 0, "gelu": 1, "relu": 2, "silu": 3, "gelu_tanh": 4, "gelu_new": 4,
             "gelu_pytorch_tanh": 4, "swish": 3}

EPI_STORE, EPI_GATED = 0, 1

#: counts launches of our own kernels (bench.py reports it as ``gpu_launches``)
launch_count = 0


def _dt(t: torch.Tensor) -> int:
    if t.dtype == torch.bfloat16:
        return 0
    if t.dtype == torch.float16:
        return 1
    raise TypeError(f"unsupported activation dtype {t.dtype}: the sm_100a kernels take bf16 or fp16")


def _p(t: Optional[torch.Tensor]):
    return None if t is None else t.data_ptr()


def _stream() -> int:
    return torch.cuda.current_stream().cuda_stream


def _count(n: int = 1) -> None:
    global launch_count
    launch_count += n


def _check_2d(t: torch.Tensor, name: str) -> None:
    if t.dim() != 2 or t.stride(1) != 1:
        raise ValueError(f"{name} must be 2-D with a contiguous last dimension, got {tuple(t.shape)} / {t.stride()}")


class TiledWeight:
    """A weight matrix ``[N, K]`` stored pre-tiled for TMA: ``data[ceil(N/32), ceil(K/64), 32, 64]`` so that the
    32x64 block a GEMM tile needs per k-step is one contiguous 4 KB chunk (zero padded at the edges). Streaming a
    row-major weight fetches 128-byte segments 2*K bytes apart — poor DRAM page locality for decode-size GEMMs
    that are pure weight streaming; the tiled layout makes every fetch sequential."""

    def __init__(self, data: torch.Tensor, n: int, k: int):
        self.data, self.n, self.k = data, n, k

    @staticmethod
    def from_dense(w: torch.Tensor) -> "TiledWeight":
        n, k = w.shape
        np_, kp = -(-n // 32) * 32, -(-k // 64) * 64
        if (np_, kp) != (n, k):
            wp = w.new_zeros((np_, kp))
            wp[:n, :k] = w
            w = wp
        data = w.view(np_ // 32, 32, kp // 64, 64).permute(0, 2, 1, 3).contiguous()
        return TiledWeight(data, n, k)

    def to_dense(self) -> torch.Tensor:
        s, kb = self.data.shape[0], self.data.shape[1]
        return self.data.permute(0, 2, 1, 3).reshape(s * 32, kb * 64)[:self.n, :self.k]

    def rows_from(self, start: int) -> "TiledWeight":
        """Zero-copy view of rows ``[start, n)`` (start must be a multiple of 32)."""
        assert start % 32 == 0
        return TiledWeight(self.data[start // 32:], self.n - start, self.k)

    @property
    def shape(self):
        return (self.n, self.k)

    @property
    def dtype(self):
        return self.data.dtype

    @property
    def device(self):
        return self.data.device

    def data_ptr(self) -> int:
        return self.data.data_ptr()


def _w(w):
    """(pointer, ldw) of a weight operand: ldw = -1 marks the tiled layout for the C ABI."""
    if isinstance(w, TiledWeight):
        return w.data.data_ptr(), -1
    return w.data_ptr(), w.stride(0)


def gemm_splitk_workspace_bytes(M: int, N: int) -> int:
    return int(native.cuda_lib().hds_gemm_splitk_workspace_bytes(M, N))


def gemm(a: torch.Tensor, w: torch.Tensor, bias: Optional[torch.Tensor] = None, act: int = ACT_IDENTITY,
         gated: bool = False, out: Optional[torch.Tensor] = None, bn: int = 0, splits: int = 0,
         workspace: Optional[torch.Tensor] = None) -> torch.Tensor:
    """``out = epilogue(a @ w.T + bias)`` on the tcgen05 GEMM. ``gated``: interleaved (gate, up) columns.
    ``workspace`` (zero-initialised uint8 scratch) enables split-K for decode-size M."""
    _check_2d(a, "a")
    M, K = a.shape
    N = w.shape[0]
    assert w.shape[1] == K and w.dtype == a.dtype
    wp, ldw = _w(w)
    n_out = N // 2 if gated else N
    if out is None:
        out = torch.empty((M, n_out), dtype=a.dtype, device=a.device)
    _check_2d(out, "out")
    lib = native.cuda_lib()
    rc = lib.hds_gemm(_p(a), a.stride(0), wp, ldw, M, N, K, _dt(a), EPI_GATED if gated else EPI_STORE,
                      act, _p(bias), _p(out), out.stride(0), bn, splits, _p(workspace),
                      0 if workspace is None else workspace.numel() * workspace.element_size(), _stream())
    native.check(rc, "hds_gemm")
    _count()
    return out


def gemm_accum(a: torch.Tensor, w: torch.Tensor, accum: torch.Tensor, bias: Optional[torch.Tensor] = None,
               peer_ptrs: Optional[Sequence[int]] = None, multicast_ptr: Optional[int] = None, bn: int = 0,
               splits: int = 0) -> None:
    """``accum (fp32) += a @ w.T``. With ``peer_ptrs`` / ``multicast_ptr`` the epilogue adds its tile into every
    tensor-parallel rank's buffer (``red.sys`` / ``multimem.red``: the push form of GEMM -> all-reduce; the serving path
    uses the pull form — local accumulation, flag signal, reduction in :func:`norm_tp`)."""
    _check_2d(a, "a")
    M, K = a.shape
    N = w.shape[0]
    wp, ldw = _w(w)
    assert accum.dtype == torch.float32 and accum.stride(-1) == 1
    if multicast_ptr is not None:
        ptrs, mode = [multicast_ptr], 2
    elif peer_ptrs is not None:
        ptrs, mode = list(peer_ptrs), 1
    else:
        ptrs, mode = [accum.data_ptr()], 0
    arr = (ctypes.c_void_p * len(ptrs))(*ptrs)
    lib = native.cuda_lib()
    rc = lib.hds_gemm_accum(_p(a), a.stride(0), wp, ldw, M, N, K, _dt(a), _p(bias), arr, len(ptrs), mode,
                            accum.stride(0) if accum.dim() == 2 else N, bn, splits, _stream())
    native.check(rc, "hds_gemm_accum")
    _count()


def restore_kv_proj(latents: torch.Tensor, w_kv: torch.Tensor, kv_cache_layer: torch.Tensor, tok_pos: torch.Tensor,
                    tok_slot: torch.Tensor, inv_freq: Optional[torch.Tensor], n_kv_heads: int, head_dim: int,
                    block_size: int, bias_kv: Optional[torch.Tensor] = None) -> None:
    """HCache restore of one layer: K/V projection + RoPE + paged KV write in one tcgen05 kernel."""
    _check_2d(latents, "latents")
    T, H = latents.shape
    lib = native.cuda_lib()
    wp, ldw = _w(w_kv)
    rc = lib.hds_restore_kv_proj(_p(latents), latents.stride(0), wp, ldw, T, H, _dt(latents),
                                 _p(bias_kv), _p(kv_cache_layer), _p(tok_pos), _p(tok_slot), _p(inv_freq), n_kv_heads,
                                 head_dim, block_size, 0 if inv_freq is None else 1, _stream())
    native.check(rc, "hds_restore_kv_proj")
    _count()


def restore_kv_proj_peers(peer_ptrs: Sequence[int], rows_per_peer: int, T: int, H: int, dtype: torch.dtype, w_kv,
                          kv_cache_layer: torch.Tensor, tok_pos: torch.Tensor, tok_slot: torch.Tensor,
                          inv_freq: Optional[torch.Tensor], n_kv_heads: int, head_dim: int, block_size: int,
                          bias_kv: Optional[torch.Tensor] = None) -> None:
    """TP restore: latents row-sharded over peers' symmetric buffers, gathered by the kernel's TMA loads."""
    arr = (ctypes.c_void_p * len(peer_ptrs))(*peer_ptrs)
    wp, ldw = _w(w_kv)
    lib = native.cuda_lib()
    rc = lib.hds_restore_kv_proj_peers(arr, len(peer_ptrs), rows_per_peer, H, wp, ldw, T, H,
                                       0 if dtype == torch.bfloat16 else 1, _p(bias_kv), _p(kv_cache_layer), _p(tok_pos),
                                       _p(tok_slot), _p(inv_freq), n_kv_heads, head_dim, block_size,
                                       0 if inv_freq is None else 1, _stream())
    native.check(rc, "hds_restore_kv_proj_peers")
    _count()


def qkv_rope_proj(x: torch.Tensor, w_qkv: torch.Tensor, q_out: torch.Tensor, kv_cache_layer: torch.Tensor,
                  tok_pos: torch.Tensor, tok_slot: torch.Tensor, inv_freq: Optional[torch.Tensor], n_q_heads: int,
                  n_kv_heads: int, head_dim: int, block_size: int, bias_qkv: Optional[torch.Tensor] = None) -> None:
    """Forward QKV projection with RoPE + KV-append fused into the GEMM epilogue."""
    _check_2d(x, "x")
    T, H = x.shape
    lib = native.cuda_lib()
    wp, ldw = _w(w_qkv)
    rc = lib.hds_qkv_rope_proj(_p(x), x.stride(0), wp, ldw, T, H, _dt(x), _p(bias_qkv), _p(q_out),
                               _p(kv_cache_layer), _p(tok_pos), _p(tok_slot), _p(inv_freq), n_q_heads, n_kv_heads,
                               head_dim, block_size, 0 if inv_freq is None else 1, _stream())
    native.check(rc, "hds_qkv_rope_proj")
    _count()


def norm(x: torch.Tensor, gamma: torch.Tensor, out: torch.Tensor, *, beta: Optional[torch.Tensor] = None,
         delta: Optional[torch.Tensor] = None, zero_delta: bool = False, res_out: Optional[torch.Tensor] = None,
         row_idx: Optional[torch.Tensor] = None, rows: Optional[int] = None, eps: float = 1e-5,
         layer_norm: bool = False, post_norm: bool = False) -> torch.Tensor:
    """RMSNorm / LayerNorm with fused residual add (``delta`` may be the fp32 partial-sum buffer of
    :func:`gemm_accum`, which is consumed and optionally re-zeroed) and optional row gather."""
    cols = x.shape[-1]
    n = rows if rows is not None else (row_idx.numel() if row_idx is not None else x.numel() // cols)
    lib = native.cuda_lib()
    rc = lib.hds_norm(_p(x), x.stride(-2) if x.dim() > 1 else cols, _p(delta),
                      1 if (delta is not None and delta.dtype == torch.float32) else 0, 1 if zero_delta else 0,
                      _p(res_out), _p(out), out.stride(-2) if out.dim() > 1 else cols, _p(gamma), _p(beta), _p(row_idx),
                      n, cols, eps, 1 if layer_norm else 0, 1 if post_norm else 0, _dt(x), _stream())
    native.check(rc, "hds_norm")
    _count()
    return out


def _ptr_array(ptrs: Optional[Sequence[int]]):
    if ptrs is None:
        return None
    return (ctypes.c_void_p * len(ptrs))(*[int(p) for p in ptrs])


def norm_tp(ctx: int, x: torch.Tensor, gamma: torch.Tensor, out: torch.Tensor, src_ptrs: Sequence[int], *,
            src_mc: Optional[int] = None, partial_16bit: bool = False, zero_buf: Optional[torch.Tensor] = None,
            beta: Optional[torch.Tensor] = None, res_out: Optional[torch.Tensor] = None, out_ptrs: Optional[Sequence[int]] = None,
            out_mc: Optional[int] = None, row0: int = 0, n_rows: Optional[int] = None, wait: bool = True, signal: bool = False,
            eps: float = 1e-5, layer_norm: bool = False) -> torch.Tensor:
    """The collective after a row-parallel GEMM fused with residual add + norm (``hds_norm_tp``): all-reduce shape
    (``out_ptrs`` / ``out_mc`` unset: every rank reduces all rows) or reduce-scatter + all-gather shape (this rank reduces
    rows ``[row0, row0 + n_rows)`` and writes their normalised values into every rank's ``out``)."""
    cols = x.shape[-1]
    total = x.numel() // cols
    n = total if n_rows is None else n_rows
    lib = native.cuda_lib()
    rc = lib.hds_norm_tp(ctx, int(wait), int(signal), _p(x), x.stride(-2) if x.dim() > 1 else cols, _ptr_array(src_ptrs),
                         src_mc, int(partial_16bit), _p(zero_buf), _p(res_out), _p(out), out.stride(-2) if out.dim() > 1 else cols,
                         _ptr_array(out_ptrs), out_mc, row0, n, total, _p(gamma), _p(beta), cols, eps, 1 if layer_norm else 0,
                         _dt(x), _stream())
    native.check(rc, "hds_norm_tp")
    _count()
    return out


# ---- tensor-parallel flag protocol (tp_comm.cu) ------------------------------------------------------------------------
def tp_ctx_create(rank: int, n_ranks: int, pad_ptrs: Sequence[int]) -> int:
    ctx = native.cuda_lib().hds_tp_ctx_create(rank, n_ranks, _ptr_array(pad_ptrs))
    if ctx < 0:
        native.check(ctx, "hds_tp_ctx_create")
    return ctx


def tp_ctx_destroy(ctx: int) -> None:
    native.cuda_lib().hds_tp_ctx_destroy(ctx)


def tp_arm(ctx: int, wait: bool, signal: bool) -> None:
    """Attach a flag wait / signal to the NEXT GEMM-family kernel this thread launches (see ``hds_tp_arm``)."""
    native.check(native.cuda_lib().hds_tp_arm(ctx, int(wait), int(signal)), "hds_tp_arm")


def tp_disarm() -> bool:
    return bool(native.cuda_lib().hds_tp_disarm())


def tp_signal(ctx: int) -> None:
    native.check(native.cuda_lib().hds_tp_signal(ctx, _stream()), "hds_tp_signal")
    _count()


def tp_wait(ctx: int) -> None:
    native.check(native.cuda_lib().hds_tp_wait(ctx, _stream()), "hds_tp_wait")
    _count()


def tp_gather_cols(ctx: int, out: torch.Tensor, src_ptrs: Sequence[int], ld_src: int, col0: Sequence[int],
                   width: Sequence[int], rows: int, wait: bool = True) -> torch.Tensor:
    """``out[:rows, col0[p]:col0[p]+width[p]] = rank p's slice`` for every rank (vocabulary-sharded logits all-gather)."""
    n = len(src_ptrs)
    c0 = (ctypes.c_int32 * n)(*col0)
    wd = (ctypes.c_int32 * n)(*width)
    rc = native.cuda_lib().hds_tp_gather_cols(ctx, int(wait), _p(out), out.stride(0), _ptr_array(src_ptrs), ld_src, c0, wd, rows,
                                              out.element_size(), _stream())
    native.check(rc, "hds_tp_gather_cols")
    _count()
    return out


def rope_kv_append(qkv: torch.Tensor, q_out: torch.Tensor, kv_cache_layer: torch.Tensor, tok_pos: torch.Tensor,
                   tok_slot: torch.Tensor, inv_freq: Optional[torch.Tensor], n_q_heads: int, n_kv_heads: int,
                   head_dim: int, rotary_dim: int, block_size: int) -> None:
    _check_2d(qkv, "qkv")
    lib = native.cuda_lib()
    rc = lib.hds_rope_kv_append(_p(qkv), qkv.stride(0), _p(q_out), _p(kv_cache_layer), _p(tok_pos), _p(tok_slot),
                                _p(inv_freq), qkv.shape[0], n_q_heads, n_kv_heads, head_dim, rotary_dim, block_size,
                                0 if inv_freq is None else 1, _dt(qkv), _stream())
    native.check(rc, "hds_rope_kv_append")
    _count()


def embed(ids: torch.Tensor, weight: torch.Tensor, out: torch.Tensor, pos_weight: Optional[torch.Tensor] = None,
          tok_pos: Optional[torch.Tensor] = None, pos_offset: int = 0) -> torch.Tensor:
    assert ids.dtype == torch.int32
    lib = native.cuda_lib()
    rc = lib.hds_embed(_p(ids), _p(weight), _p(pos_weight), _p(tok_pos), pos_offset,
                       0 if pos_weight is None else pos_weight.shape[0], _p(out), ids.numel(), weight.shape[1],
                       weight.shape[0], _dt(weight), _stream())
    native.check(rc, "hds_embed")
    _count()
    return out


def gather_rows(x: torch.Tensor, row_idx: torch.Tensor, out: torch.Tensor) -> torch.Tensor:
    lib = native.cuda_lib()
    rc = lib.hds_gather_rows(_p(x), _p(row_idx), _p(out), row_idx.numel(), x.shape[-1], _stream())
    native.check(rc, "hds_gather_rows")
    _count()
    return out


def argmax(logits: torch.Tensor, out_ids: Optional[torch.Tensor] = None) -> torch.Tensor:
    _check_2d(logits, "logits")
    if out_ids is None:
        out_ids = torch.empty((logits.shape[0],), dtype=torch.int32, device=logits.device)
    code = 2 if logits.dtype == torch.float32 else _dt(logits)
    lib = native.cuda_lib()
    rc = lib.hds_argmax(_p(logits), logits.stride(0), logits.shape[0], logits.shape[1], _p(out_ids), None, code,
                        _stream())
    native.check(rc, "hds_argmax")
    _count()
    return out_ids


def kv_blocks_copy(cache: torch.Tensor, staging: torch.Tensor, block_ids: torch.Tensor, to_staging: bool) -> None:
    """cache: [layers, blocks, 2, kvh, bs, d]; staging: [n_sel, layers, 2, kvh, bs, d]."""
    n_layers = cache.shape[0]
    block_bytes = cache[0, 0].numel() * cache.element_size()
    layer_stride = cache.stride(0) * cache.element_size()
    lib = native.cuda_lib()
    rc = lib.hds_kv_blocks_copy(_p(cache), _p(staging), _p(block_ids), block_ids.numel(), n_layers, block_bytes,
                                layer_stride, 1 if to_staging else 0, _stream())
    native.check(rc, "hds_kv_blocks_copy")
    _count()


def p2p_gather_rows(dst: torch.Tensor, peer_ptrs: Sequence[int], rows_per_peer: int, n_rows: int, my_rank: int) -> None:
    """Fill ``dst[:n_rows]`` rows owned by other TP ranks from their symmetric buffers (one launch, NVLink loads)."""
    arr = (ctypes.c_void_p * len(peer_ptrs))(*peer_ptrs)
    lib = native.cuda_lib()
    rc = lib.hds_p2p_gather_rows(_p(dst), arr, len(peer_ptrs), rows_per_peer, n_rows, dst.shape[-1] * dst.element_size(),
                                 my_rank, _stream())
    native.check(rc, "hds_p2p_gather_rows")
    _count()


def cvt_f32(src: torch.Tensor, dst: torch.Tensor, zero_src: bool = False) -> torch.Tensor:
    lib = native.cuda_lib()
    rc = lib.hds_cvt_f32(_p(src), _p(dst), src.numel(), 1 if zero_src else 0, _dt(dst), _stream())
    native.check(rc, "hds_cvt_f32")
    _count()
    return dst


def splitkv_workspace_bytes(n_items: int, n_q_heads: int, n_kv_heads: int, head_dim: int, max_splits: int,
                            rows_per_item: int) -> int:
    return int(native.cuda_lib().hds_paged_attention_splitkv_workspace_bytes(n_items, n_q_heads, n_kv_heads, head_dim,
                                                                             max_splits, rows_per_item))


def _arm_alibi(lib, alibi: Optional[torch.Tensor], n_q_heads: int, window: int = 0) -> None:
    """ALiBi slopes / sliding window for the next attention launch of this thread (``hds_attention_next_alibi``,
    ``hds_attention_next_window``; both are consumed by that launch)."""
    if window:
        native.check(lib.hds_attention_next_window(int(window)), "hds_attention_next_window")
    if alibi is None:
        return
    assert alibi.dtype == torch.float32 and alibi.is_cuda and alibi.is_contiguous() and alibi.numel() == n_q_heads
    native.check(lib.hds_attention_next_alibi(_p(alibi)), "hds_attention_next_alibi")


def paged_attention_splitkv(q: torch.Tensor, out: torch.Tensor, kv_cache_layer: torch.Tensor,
                            block_tables: torch.Tensor, items: torch.Tensor, n_items: int, rows_per_item: int,
                            n_q_heads: int, n_kv_heads: int, head_dim: int, block_size: int, n_splits: int,
                            scale: float, workspace: Optional[torch.Tensor], alibi: Optional[torch.Tensor] = None,
                            window: int = 0) -> None:
    """kv_cache_layer: [blocks, 2, kvh, bs, d]; items: int32 [n, 4] = (seq, q_row0, n_q, pos0). ``alibi``: fp32 [n_q_heads]
    slopes; score += slope * (key position - query position). ``window``: sliding-window attention, keys (p - window, p]."""
    lib = native.cuda_lib()
    _arm_alibi(lib, alibi, n_q_heads, window)
    rc = lib.hds_paged_attention_splitkv(_p(q), _p(out), _p(kv_cache_layer), kv_cache_layer.shape[0], _p(block_tables),
                                         block_tables.stride(0), _p(items), n_items, rows_per_item, n_q_heads,
                                         n_kv_heads, head_dim, block_size, n_splits, scale, _p(workspace),
                                         0 if workspace is None else workspace.numel() * workspace.element_size(),
                                         _dt(q), _stream())
    native.check(rc, "hds_paged_attention_splitkv")
    _count()


def prefill_workspace_bytes(n_items: int, n_q_heads: int, head_dim: int, n_splits: int) -> int:
    return int(native.cuda_lib().hds_paged_attention_prefill_workspace_bytes(n_items, n_q_heads, head_dim, n_splits))


def paged_attention_prefill(q: torch.Tensor, out: torch.Tensor, kv_cache_layer: torch.Tensor,
                            block_tables: torch.Tensor, items: torch.Tensor, n_items: int, n_q_heads: int,
                            n_kv_heads: int, head_dim: int, block_size: int, scale: float, n_splits: int = 1,
                            workspace: Optional[torch.Tensor] = None, packed: bool = False,
                            alibi: Optional[torch.Tensor] = None, window: int = 0) -> None:
    """tcgen05 paged attention. ``packed``: grouped-query decode form — one CTA per (item, KV head) whose MMA rows are
    the group's (token, query head) pairs (``hds_paged_attention_packed``). ``alibi``: fp32 [n_q_heads] slopes;
    ``window``: sliding-window attention."""
    lib = native.cuda_lib()
    _arm_alibi(lib, alibi, n_q_heads, window)
    fn = lib.hds_paged_attention_packed if packed else lib.hds_paged_attention_prefill
    rc = fn(_p(q), _p(out), _p(kv_cache_layer), kv_cache_layer.shape[0], _p(block_tables),
                                         block_tables.stride(0), _p(items), n_items, q.shape[0], n_q_heads, n_kv_heads,
                                         head_dim, block_size, scale, n_splits, _p(workspace),
                                         0 if workspace is None else workspace.numel() * workspace.element_size(),
                                         _dt(q), _stream())
    native.check(rc, "hds_paged_attention_prefill")
    _count()


# ------------------------------------------------------------------------------------------------
# mixture of experts (ref: ragged_ops/top_k_gating, moe_scatter, moe_gather, cutlass_ops/moe_gemm)
# ------------------------------------------------------------------------------------------------
def moe_padded_rows(n_tokens: int, n_experts: int, top_k: int) -> int:
    """Row capacity of the scattered activation buffer (each expert's range is padded to 128 rows)."""
    return int(native.cuda_lib().hds_moe_padded_rows(n_tokens, n_experts, top_k))


def moe_topk_gating(logits: torch.Tensor, top_k: int, normalize: bool, scores: torch.Tensor, expert_ids: torch.Tensor,
                    offsets: torch.Tensor, counts: torch.Tensor) -> None:
    """Softmax + top-k per token; ``counts`` (int32 [E], zero on entry) is incremented per assignment."""
    T, E = logits.shape
    code = 2 if logits.dtype == torch.float32 else _dt(logits)
    rc = native.cuda_lib().hds_moe_topk_gating(_p(logits), logits.stride(0), T, E, top_k, int(normalize), code, _p(scores),
                                               _p(expert_ids), _p(offsets), _p(counts), _stream())
    native.check(rc, "hds_moe_topk_gating")
    _count()


def moe_scatter(x: torch.Tensor, xs: torch.Tensor, expert_ids: torch.Tensor, offsets: torch.Tensor, counts: torch.Tensor,
                pad_off: torch.Tensor, mapped_slots: torch.Tensor, top_k: int) -> None:
    """Build ``pad_off`` (int32 [E+1]) from ``counts`` (re-zeroed) and copy every token row to its k expert rows."""
    T, H = x.shape
    E = counts.numel()
    rc = native.cuda_lib().hds_moe_scatter(_p(x), x.stride(0), _p(xs), xs.stride(0), _p(expert_ids), _p(offsets), _p(counts),
                                           _p(pad_off), _p(mapped_slots), T, E, top_k, H, _stream())
    native.check(rc, "hds_moe_scatter")
    _count(2)


def moe_gemm(a: torch.Tensor, w_stacked: torch.Tensor, pad_off: torch.Tensor, n_experts: int, out: torch.Tensor,
             act: int = ACT_IDENTITY, gated: bool = False) -> torch.Tensor:
    """Grouped GEMM: rows of ``a`` in expert e's range times ``w_stacked[e*N:(e+1)*N]``ᵀ (``w_stacked``: [E*N, K])."""
    _check_2d(a, "a")
    _check_2d(out, "out")
    rows, K = a.shape
    N = w_stacked.shape[0] // n_experts
    assert w_stacked.shape[1] == K and rows % 128 == 0 and out.shape[0] >= rows
    rc = native.cuda_lib().hds_moe_gemm(_p(a), a.stride(0), _p(w_stacked), w_stacked.stride(0), _p(pad_off), n_experts, rows,
                                        N, K, _dt(a), EPI_GATED if gated else EPI_STORE, act, _p(out), out.stride(0),
                                        _stream())
    native.check(rc, "hds_moe_gemm")
    _count()
    return out


def gemm_2sm(a: torch.Tensor, w: torch.Tensor, out: Optional[torch.Tensor] = None, bias: Optional[torch.Tensor] = None,
             act: int = ACT_IDENTITY, gated: bool = False, bn: int = 0) -> torch.Tensor:
    """``out = epilogue(a @ w.T + bias)`` on the CTA-pair kernel (``tcgen05.mma.cta_group::2``: two SMs per 256 x bn tile)."""
    _check_2d(a, "a")
    _check_2d(w, "w")
    M, K = a.shape
    N = w.shape[0]
    if out is None:
        out = torch.empty(M, N // 2 if gated else N, dtype=a.dtype, device=a.device)
    rc = native.cuda_lib().hds_gemm_2sm(_p(a), a.stride(0), _p(w), w.stride(0), M, N, K, _dt(a),
                                        EPI_GATED if gated else EPI_STORE, act, _p(bias), _p(out), out.stride(0), bn, _stream())
    native.check(rc, "hds_gemm_2sm")
    _count()
    return out


def quantize_fp8_rows(x: torch.Tensor, q: Optional[torch.Tensor] = None,
                      scales: Optional[torch.Tensor] = None) -> Tuple[torch.Tensor, torch.Tensor]:
    """Row-wise E4M3 quantisation: ``q[r] = e4m3(x[r] / scales[r])`` with ``scales[r] = amax(x[r]) / 448``.
    Rows are tokens for activations and output channels for weights. ``q`` is ``uint8`` storage of the fp8 codes."""
    _check_2d(x, "x")
    rows, K = x.shape
    if q is None:
        q = torch.empty(rows, K, dtype=torch.uint8, device=x.device)
    if scales is None:
        scales = torch.empty(rows, dtype=torch.float32, device=x.device)
    rc = native.cuda_lib().hds_quantize_fp8_rows(_p(x), x.stride(0), rows, K, _dt(x), _p(q), q.stride(0), _p(scales),
                                                 _stream())
    native.check(rc, "hds_quantize_fp8_rows")
    _count()
    return q, scales


def gemm_fp8(aq: torch.Tensor, a_scale: torch.Tensor, wq: torch.Tensor, w_scale: torch.Tensor, out: torch.Tensor,
             bias: Optional[torch.Tensor] = None, act: int = ACT_IDENTITY, gated: bool = False, bn: int = 0,
             workspace: Optional[torch.Tensor] = None, splits: int = 0, w_fmt: int = 0) -> torch.Tensor:
    """``out = epilogue((aq @ wq.T) * a_scale[:, None] * w_scale[None, :] + bias)`` with E4M3 operands on
    ``tcgen05.mma.kind::f8f6f4`` (fp32 accumulation in TMEM). ``aq``: [M, K] uint8, ``wq``: [N, K] uint8.
    M <= 64 with ``bn == 0`` takes the feature-major weight-streaming kernel (split-K through ``workspace``, a zeroed
    uint8 buffer of at least :func:`gemm_fp8_skinny_workspace_bytes`)."""
    _check_2d(aq, "aq")
    _check_2d(wq, "wq")
    _check_2d(out, "out")
    M, K = aq.shape
    N = wq.shape[0]
    # w_fmt 1: E3M2 (FP6) weights packed 6 bits each — [N, 3K/4] bytes, unpacked by the TMA on the way to shared memory
    assert wq.shape[1] == (K * 3 // 4 if w_fmt == 1 else K) and aq.dtype == torch.uint8 and wq.dtype == torch.uint8
    assert a_scale.dtype == torch.float32 and w_scale.dtype == torch.float32
    lib = native.cuda_lib()
    if w_fmt:
        native.check(lib.hds_fp8_next_weight_format(w_fmt), "hds_fp8_next_weight_format")
    if M <= 64 and bn == 0:
        rc = lib.hds_gemm_fp8_skinny(_p(aq), aq.stride(0), _p(a_scale), _p(wq), wq.stride(0), _p(w_scale), M, N, K, _dt(out),
                                     EPI_GATED if gated else EPI_STORE, act, _p(bias), _p(out), out.stride(0), None, 0,
                                     _p(workspace), 0 if workspace is None else workspace.numel(), splits, _stream())
        native.check(rc, "hds_gemm_fp8_skinny")
    else:
        rc = lib.hds_gemm_fp8(_p(aq), aq.stride(0), _p(a_scale), _p(wq), wq.stride(0), _p(w_scale), M, N, K, _dt(out),
                              EPI_GATED if gated else EPI_STORE, act, _p(bias), _p(out), out.stride(0), bn, _stream())
        native.check(rc, "hds_gemm_fp8")
    _count()
    return out


def gemm_fp8_skinny_workspace_bytes(M: int, N: int) -> int:
    nt = 16 if M <= 16 else 32 if M <= 32 else 64
    n_tiles = (N + 127) // 128
    return ((n_tiles * 128 * nt * 4 + 255) & ~255) + n_tiles * 16


def gemm_fp8_accum(aq: torch.Tensor, a_scale: torch.Tensor, wq: torch.Tensor, w_scale: torch.Tensor, accum: torch.Tensor,
                   bias: Optional[torch.Tensor] = None, out_dtype: torch.dtype = torch.bfloat16, splits: int = 0,
                   w_fmt: int = 0) -> torch.Tensor:
    """``accum (fp32) += (aq @ wq.T) * scales (+ bias once)`` for M <= 64 (decode): split-K partial sums go straight
    into ``accum`` with fp32 atomics, no scratch."""
    M, K = aq.shape
    N = wq.shape[0]
    assert accum.dtype == torch.float32 and M <= 64
    dt = 0 if out_dtype == torch.bfloat16 else 1
    if w_fmt:
        native.check(native.cuda_lib().hds_fp8_next_weight_format(w_fmt), "hds_fp8_next_weight_format")
    rc = native.cuda_lib().hds_gemm_fp8_skinny(_p(aq), aq.stride(0), _p(a_scale), _p(wq), wq.stride(0), _p(w_scale), M, N, K, dt,
                                               2, ACT_IDENTITY, _p(bias), None, 0, _p(accum), accum.stride(0), None, 0, splits,
                                               _stream())
    native.check(rc, "hds_gemm_fp8_skinny")
    _count()
    return accum


def moe_gather(y: torch.Tensor, mapped_slots: torch.Tensor, scores: torch.Tensor, out: torch.Tensor) -> None:
    """``out[t] += sum_j scores[t, j] * y[mapped_slots[t, j]]`` (``out`` fp32)."""
    T, k = mapped_slots.shape
    assert out.dtype == torch.float32
    rc = native.cuda_lib().hds_moe_gather(_p(y), y.stride(0), _p(mapped_slots), _p(scores), _p(out), out.stride(0), T, k,
                                          y.shape[1], _dt(y), _stream())
    native.check(rc, "hds_moe_gather")
    _count()
