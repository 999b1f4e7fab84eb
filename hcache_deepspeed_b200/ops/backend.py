"""Two interchangeable op back ends with one signature set.

``CudaOps``  — the sm_100a kernels (``ops/kernels.py``); used whenever tensors live on a GPU.
``CpuOps``   — the PyTorch restatements (``ops/reference.py``); used for the CPU configuration
               (``init_inference`` on CPU, CPU-only CI). Never used on a CUDA device.

The model executor (:mod:`hcache_deepspeed_b200.models.transformer`) is written once against this
interface. Both back ends write into caller-provided buffers so steady-state forwards allocate nothing
(the role of the reference's ``empty_from`` view allocator, ``deepspeed/inference/v2/allocator.py:14-28``).
"""
from __future__ import annotations

from typing import Optional

import torch

from . import reference as ref
from .fp8 import Fp8Weight, quantize_rows_reference


def _fp8_emulate(a, w):
    """CPU restatement of the W8A8 path: fp8-round the activations per token, de-quantise both operands to fp32."""
    if not isinstance(w, Fp8Weight):
        return a, w
    q, s = quantize_rows_reference(a)
    return q.view(torch.float8_e4m3fn).float() * s[:, None], w.dequantize()


class CpuOps:
    name = "cpu-reference"

    # ---- linear ---------------------------------------------------------------------------------
    def gemm(self, a, w, out, bias=None, act=0, gated=False, workspace=None, comm=None, tp_wait=False, tp_signal=False):
        a, w = _fp8_emulate(a, w)
        out.copy_(ref.gemm(a, w, bias=bias, act=act, gated=gated, out_dtype=out.dtype))
        return out

    def gemm_accum(self, a, w, accum, bias=None, comm=None, tp_signal=False):
        a, w = _fp8_emulate(a, w)
        accum += ref.gemm(a, w, bias=bias, out_dtype=torch.float32)

    def reserve_fp8(self, max_tokens, ks, device) -> None:
        pass

    # ---- norm -----------------------------------------------------------------------------------
    def norm(self, x, gamma, out, beta=None, delta=None, zero_delta=False, res_out=None, eps=1e-5, layer_norm=False,
             comm=None, zero_buf=None, scatter_rows=None):
        if delta is not None:
            s = (x.float() + delta.float()).to(x.dtype)
            if zero_delta and delta.dtype == torch.float32:
                delta.zero_()
            if res_out is not None:
                res_out.copy_(s)
            x = s
        out.copy_(ref.norm(x, gamma, beta, eps, layer_norm))
        return out

    # ---- attention-side -------------------------------------------------------------------------------
    def qkv_rope_kv_append(self, x, w_qkv, b_qkv, qkv_buf, q_out, kv_layer, tok_pos, tok_slot, inv_freq, n_q, n_kv, d,
                           rot_dim, block_size, workspace=None, comm=None, tp_wait=False):
        T = x.shape[0]
        dt = x.dtype
        x, w_qkv = _fp8_emulate(x, w_qkv)
        qkv = ref.gemm(x, w_qkv, bias=b_qkv, out_dtype=dt)
        q, k, v = qkv.split([n_q * d, n_kv * d, n_kv * d], dim=1)
        q, k, v = q.reshape(T, n_q, d), k.reshape(T, n_kv, d), v.reshape(T, n_kv, d)
        if inv_freq is not None:
            q = ref.rope_rotate_half(q, tok_pos, inv_freq, rot_dim)
            k = ref.rope_rotate_half(k, tok_pos, inv_freq, rot_dim)
        q_out.copy_(q.reshape(T, n_q * d))
        ref.kv_append(kv_layer, k, v, tok_slot)

    def restore_kv_proj(self, latents, w_kv, b_kv, kv_layer, tok_pos, tok_slot, inv_freq, n_kv, d, rot_dim, block_size,
                        scratch=None):
        T = latents.shape[0]
        dt = latents.dtype
        latents, w_kv = _fp8_emulate(latents, w_kv)
        kv = ref.gemm(latents, w_kv, bias=b_kv, out_dtype=dt)
        k, v = kv[:, :n_kv * d].reshape(T, n_kv, d), kv[:, n_kv * d:].reshape(T, n_kv, d)
        if inv_freq is not None:
            k = ref.rope_rotate_half(k, tok_pos, inv_freq, rot_dim)
        ref.kv_append(kv_layer, k, v, tok_slot)

    def attention(self, q, out, kv_layer, batch, n_q, n_kv, d, scale, ws_small=None, ws_large=None, alibi=None, window=0):
        T = q.shape[0]
        q3, o3 = q.view(T, n_q, d), out.view(T, n_q, d)
        for items in (batch.items_small(), batch.items_large()):
            if items.shape[0]:
                ref.paged_attention(q3, kv_layer, batch.block_tables(), items, n_q, n_kv, scale, out=o3, alibi=alibi,
                                    window=window)

    # ---- misc ---------------------------------------------------------------------------------------
    def embed(self, ids, weight, out, pos_weight=None, tok_pos=None, pos_offset=0):
        out.copy_(ref.embed(ids, weight, pos_weight, tok_pos, pos_offset))
        return out

    def gather_rows(self, x, row_idx, out):
        out.copy_(x[row_idx.long()])
        return out

    def argmax(self, logits):
        return logits.float().argmax(-1).to(torch.int32)


class CudaOps:
    name = "sm_100a"

    #: at or below this many tokens the QKV projection runs as split-K GEMM + RoPE/append kernel (fills all
    #: SMs while streaming weights); above it the RoPE/KV epilogue is fused into the tcgen05 GEMM
    FUSED_QKV_MIN_TOKENS = 129

    def __init__(self):
        from . import kernels
        self.k = kernels

    def _fp8_act(self, a):
        """Per-token E4M3 quantisation of ``a`` into persistent buffers (no allocation in steady state)."""
        T, K = a.shape
        bufs = self.__dict__.setdefault("_fp8_bufs", {})
        q, s = bufs.get(K, (None, None))
        if q is None or q.shape[0] < T or q.device != a.device:
            cap = max(T, 0 if q is None else 2 * q.shape[0], 64)
            q = torch.empty(cap, K, dtype=torch.uint8, device=a.device)
            s = torch.empty(cap, dtype=torch.float32, device=a.device)
            bufs[K] = (q, s)
        return self.k.quantize_fp8_rows(a, q[:T], s[:T])

    def reserve_fp8(self, max_tokens: int, ks, device) -> None:
        """Size the activation-quantisation buffers up front (before any CUDA-graph capture)."""
        bufs = self.__dict__.setdefault("_fp8_bufs", {})
        for K in ks:
            bufs[K] = (torch.empty(max_tokens, K, dtype=torch.uint8, device=device),
                       torch.empty(max_tokens, dtype=torch.float32, device=device))

    def gemm(self, a, w, out, bias=None, act=0, gated=False, workspace=None, comm=None, tp_wait=False, tp_signal=False):
        """``tp_wait``: ``a`` was all-gathered into this GPU's buffer by the peers' norm kernels — the first kernel that
        reads it acquires their flags. ``tp_signal``: the GEMM's last CTA releases ``out`` to the peers."""
        if isinstance(w, Fp8Weight):
            if tp_wait:
                comm.arm(wait=True)          # consumed by the activation quantiser, the first reader of `a`
            aq, a_s = self._fp8_act(a)
            if tp_signal:
                comm.arm(signal=True)
            return self.k.gemm_fp8(aq, a_s, w.q, w.scale, out, bias=bias, act=act, gated=gated,
                                   workspace=workspace if a.shape[0] <= 64 else None, w_fmt=w.fmt_code)
        if tp_wait or tp_signal:
            comm.arm(wait=tp_wait, signal=tp_signal)
        return self.k.gemm(a, w, bias=bias, act=act, gated=gated, out=out,
                           workspace=workspace if a.shape[0] <= 128 else None)

    def gemm_accum(self, a, w, accum, bias=None, comm=None, tp_signal=False):
        """``accum (fp32) += a @ w.T``. ``tp_signal`` (tensor parallel): the kernel's last CTA releases the partial sums
        to the peers; the reduction itself happens in ``norm_tp``."""
        if isinstance(w, Fp8Weight):
            if a.shape[0] > 64:
                raise NotImplementedError("fp8 weights: above 64 tokens the executor routes reducing GEMMs through the "
                                          "16-bit partial-sum buffer")
            aq, a_s = self._fp8_act(a)
            if tp_signal:
                comm.arm(signal=True)
            self.k.gemm_fp8_accum(aq, a_s, w.q, w.scale, accum, bias=bias, out_dtype=a.dtype, w_fmt=w.fmt_code)
            return
        if tp_signal:
            comm.arm(signal=True)
        self.k.gemm_accum(a, w, accum, bias=bias)

    def norm(self, x, gamma, out, beta=None, delta=None, zero_delta=False, res_out=None, eps=1e-5, layer_norm=False,
             comm=None, zero_buf=None, scatter_rows=None):
        """``comm`` set: ``delta`` is this rank's partial-sum buffer of a row-parallel GEMM (fp32 or 16-bit, symmetric
        memory) and the kernel performs the collective (see ``hds_norm_tp``). ``scatter_rows = (row0, n_rows)`` selects
        the reduce-scatter + all-gather shape: ``out`` (symmetric) is written on every rank and the kernel signals."""
        if comm is not None and delta is not None:
            comm.pre_wait()
            kw = {}
            if scatter_rows is not None:
                mc = comm.multicast_ptr(out, "st")
                kw = dict(row0=scatter_rows[0], n_rows=scatter_rows[1], signal=not comm.debug_nosync, out_mc=mc,
                          out_ptrs=None if mc else comm.peer_ptrs_raw(out))
            return self.k.norm_tp(comm.ctx, x, gamma, out, comm.peer_ptrs_raw(delta), src_mc=comm.multicast_ptr(delta, "ld"),
                                  partial_16bit=delta.dtype != torch.float32, zero_buf=zero_buf, beta=beta, res_out=res_out,
                                  eps=eps, layer_norm=layer_norm, wait=not comm.debug_nosync, **kw)
        return self.k.norm(x, gamma, out, beta=beta, delta=delta, zero_delta=zero_delta, res_out=res_out, eps=eps,
                           layer_norm=layer_norm)

    def qkv_rope_kv_append(self, x, w_qkv, b_qkv, qkv_buf, q_out, kv_layer, tok_pos, tok_slot, inv_freq, n_q, n_kv, d,
                           rot_dim, block_size, workspace=None, comm=None, tp_wait=False):
        T = x.shape[0]
        fusable = d in (64, 128) and (inv_freq is None or rot_dim == d)
        if isinstance(w_qkv, Fp8Weight):
            self.gemm(x, w_qkv, qkv_buf, bias=b_qkv, comm=comm, tp_wait=tp_wait)
            self.k.rope_kv_append(qkv_buf, q_out, kv_layer, tok_pos, tok_slot, inv_freq, n_q, n_kv, d, rot_dim, block_size)
        elif fusable and T >= self.FUSED_QKV_MIN_TOKENS:
            if tp_wait:
                comm.arm(wait=True)
            self.k.qkv_rope_proj(x, w_qkv, q_out, kv_layer, tok_pos, tok_slot, inv_freq, n_q, n_kv, d, block_size,
                                 bias_qkv=b_qkv)
        else:
            if tp_wait:
                comm.arm(wait=True)
            self.k.gemm(x, w_qkv, bias=b_qkv, out=qkv_buf, workspace=workspace if T <= 128 else None)
            self.k.rope_kv_append(qkv_buf, q_out, kv_layer, tok_pos, tok_slot, inv_freq, n_q, n_kv, d, rot_dim, block_size)

    def restore_kv_proj(self, latents, w_kv, b_kv, kv_layer, tok_pos, tok_slot, inv_freq, n_kv, d, rot_dim, block_size,
                        scratch=None):
        if isinstance(w_kv, Fp8Weight):
            if scratch is None or scratch.shape[0] < latents.shape[0]:
                scratch = torch.empty(latents.shape[0], w_kv.shape[0], dtype=latents.dtype, device=latents.device)
            scratch = scratch[:latents.shape[0]]
            self.gemm(latents, w_kv, scratch, bias=b_kv)
            self.k.rope_kv_append(scratch, scratch, kv_layer, tok_pos, tok_slot, inv_freq, 0, n_kv, d, rot_dim, block_size)
        elif d in (64, 128) and (inv_freq is None or rot_dim == d):
            self.k.restore_kv_proj(latents, w_kv, kv_layer, tok_pos, tok_slot, inv_freq, n_kv, d, block_size, bias_kv=b_kv)
        else:
            # head sizes the fused epilogue does not cover: plain GEMM into scratch, then RoPE/append with zero
            # query heads (scratch: [T, 2 * n_kv * d])
            self.k.gemm(latents, w_kv, bias=b_kv, out=scratch)
            self.k.rope_kv_append(scratch, scratch, kv_layer, tok_pos, tok_slot, inv_freq, 0, n_kv, d, rot_dim, block_size)

    def attention(self, q, out, kv_layer, batch, n_q, n_kv, d, scale, ws_small=None, ws_large=None, alibi=None, window=0):
        k = self.k
        bs = batch.block_size
        # with a sliding window an item reads at most window + one tile of keys: size the KV split for that
        seen = (lambda n_kv_tokens: min(n_kv_tokens, window + 256)) if window else (lambda n_kv_tokens: n_kv_tokens)
        if batch.n_items_small:
            n = batch.n_items_small
            if use_packed_decode(n_q, n_kv, d):
                # grouped-query decode: the group's query heads share one pass over the KV pages, on the tensor cores
                splits = pick_packed_splits(n, n_kv, seen(batch.max_kv_len_small))
                k.paged_attention_prefill(q, out, kv_layer, batch.block_tables(), batch.items_small(), n, n_q, n_kv, d, bs,
                                          scale, n_splits=splits, workspace=ws_large if splits > 1 else None, packed=True,
                                          alibi=alibi, window=window)
            else:
                splits = pick_decode_splits(n, n_kv, seen(batch.max_kv_len_small))
                k.paged_attention_splitkv(q, out, kv_layer, batch.block_tables(), batch.items_small(), n,
                                          batch.rows_per_small_item, n_q, n_kv, d, bs, splits, scale, ws_small, alibi=alibi,
                                          window=window)
        if batch.n_items_large:
            n = batch.n_items_large
            splits = pick_prefill_splits(n, n_q, seen(batch.max_kv_len_large))
            k.paged_attention_prefill(q, out, kv_layer, batch.block_tables(), batch.items_large(), n, n_q, n_kv, d, bs,
                                      scale, n_splits=splits, workspace=ws_large if splits > 1 else None, alibi=alibi,
                                      window=window)

    def embed(self, ids, weight, out, pos_weight=None, tok_pos=None, pos_offset=0):
        return self.k.embed(ids, weight, out, pos_weight=pos_weight, tok_pos=tok_pos, pos_offset=pos_offset)

    def gather_rows(self, x, row_idx, out):
        return self.k.gather_rows(x, row_idx, out)

    def argmax(self, logits):
        return self.k.argmax(logits)


N_SM = 148
MAX_DECODE_SPLITS = 16
MAX_PREFILL_SPLITS = 16


def pick_decode_splits(n_items: int, n_kv_heads: int, max_kv_len: int) -> int:
    """KV splits for the split-KV kernel. Up to one CTA per SM: split until (items x kv heads x splits) fills the SMs —
    measured (profiles/kernel_bench_attn_warps.txt): 32 CTAs -> 4 splits, 64 -> 2. Up to the 2 x 148 resident slots of
    the 4-warp configuration: one wave, no split. Beyond that a second, mostly empty wave would cost up to half the
    bandwidth (320 CTAs: 3.8 TB/s unsplit vs 5.1 TB/s with 4 splits, profiles/kernel_bench_attn_waves.txt): take the
    smallest split count whose waves are at least 85 % full. Each split keeps at least 2 tiles (256 keys)."""
    ctas = max(1, n_items * n_kv_heads)
    tiles = max(1, (max_kv_len + 127) // 128)
    cap = max(1, min(MAX_DECODE_SPLITS, tiles // 2 if tiles >= 2 else 1))
    slots = 2 * N_SM
    if ctas <= N_SM:
        want = max(1, N_SM // ctas)
    elif ctas <= slots:
        want = 1
    else:
        want, best = 1, 0.0
        for s in range(1, min(cap, 8) + 1):
            eff = ctas * s / (slots * -(-ctas * s // slots))
            if eff >= 0.85:
                want = s
                break
            if eff > best:
                want, best = s, eff
    return int(max(1, min(want, cap)))


#: grouped-query decode goes to the tensor-core kernel from this group size on (HDS_GQA_PACKED=0 keeps the SIMT kernel)
PACKED_MIN_GROUP = 4


def use_packed_decode(n_q_heads: int, n_kv_heads: int, head_dim: int) -> bool:
    import os
    G = n_q_heads // max(1, n_kv_heads)
    return (head_dim == 128 and G >= PACKED_MIN_GROUP and 128 % G == 0 and os.environ.get("HDS_GQA_PACKED", "1") != "0")


def pick_packed_splits(n_items: int, n_kv_heads: int, max_kv_len: int) -> int:
    """KV splits of the head-packed decode kernel (one CTA per SM): as many as keep (items x KV heads x splits) within
    one wave of CTAs; each split keeps at least 2 tiles of 128 keys."""
    ctas = max(1, n_items * n_kv_heads)
    tiles = max(1, (max_kv_len + 127) // 128)
    cap = max(1, min(MAX_PREFILL_SPLITS, tiles // 2 if tiles >= 2 else 1))
    return int(max(1, min(cap, N_SM // ctas)))


def pick_prefill_splits(n_items: int, n_q_heads: int, max_kv_len: int) -> int:
    """KV splits for the tensor-core kernel: one CTA per SM when (items x heads) alone cannot fill the GPU."""
    ctas = max(1, n_items * n_q_heads)
    if ctas >= N_SM:
        return 1
    want = (N_SM + ctas - 1) // ctas
    tiles = max(1, (max_kv_len + 127) // 128)
    return int(max(1, min(want, MAX_PREFILL_SPLITS, tiles // 2 if tiles >= 2 else 1)))


def get_ops(device: torch.device):
    return CudaOps() if device.type == "cuda" else CpuOps()
