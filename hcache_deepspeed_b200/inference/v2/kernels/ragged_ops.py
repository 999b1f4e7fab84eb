"""Ragged-batch op wrappers (``deepspeed/inference/v2/kernels/ragged_ops/``: ``embed``, ``logits_gather``,
``top_k_gating``, ``moe_scatter``, ``moe_gather``, ``blocked_flash``, ``linear_blocked_kv_rotary``, ``atom_builder``).

CUDA tensors go to the sm_100a kernels of ``libhds_b200.so``; host tensors take a torch formulation with the SAME
buffer contract (so the MoE pipeline can be unit-tested on a CPU). The batch metadata is this framework's item list —
``int32 [n, 4]`` rows ``(seq, q_row0, n_q, pos0)`` — instead of the reference's atoms; :class:`AtomBuilder` produces it."""
from __future__ import annotations

from typing import Optional

import torch

from ....ops import reference as ref
from .ds_kernel import DSKernelBase

MOE_ROW_PAD = 128      # each expert's row range is padded to the grouped GEMM's M tile


def _k():
    from ....ops import kernels
    return kernels


class RaggedEmbeddingKernel(DSKernelBase):

    def __init__(self, embed_dtype: torch.dtype, token_dtype: torch.dtype = torch.int32, embed_dim: int = 0):
        self._check_dtype(embed_dtype)
        if token_dtype not in (torch.int32, torch.int64):
            raise ValueError(f"Unsupported token dtype: {token_dtype}")

    def __call__(self, embedded_tokens: torch.Tensor, token_ids: torch.Tensor, embedding_weight: torch.Tensor,
                 position_embed_weight: Optional[torch.Tensor] = None, token_positions: Optional[torch.Tensor] = None,
                 position_embed_offset: int = 0) -> torch.Tensor:
        self._ops(embedding_weight).embed(token_ids, embedding_weight, embedded_tokens, position_embed_weight, token_positions,
                                          position_embed_offset)
        return embedded_tokens


class RaggedLogitsGather(DSKernelBase):
    """Rows of the last token of every sequence -> a dense ``[n_seqs, hidden]`` tensor (only those get logits)."""

    def __init__(self, model_dim: int, fp_dtype: torch.dtype):
        self._check_dtype(fp_dtype)

    def __call__(self, final_token_activations: torch.Tensor, all_activations: torch.Tensor, last_rows: torch.Tensor) -> torch.Tensor:
        self._ops(all_activations).gather_rows(all_activations, last_rows, final_token_activations)
        return final_token_activations


class RaggedTopKGating(DSKernelBase):
    """Softmax over experts + top-k per token. Fills ``scores`` fp32 [T,k], ``assignments`` int32 [T,k], ``offsets``
    int32 [T,k] (position of the token inside its expert's range) and adds to ``expert_counts`` int32 [E]."""

    def __init__(self, logit_dtype: torch.dtype, normalize: bool = True):
        self._check_dtype(logit_dtype)
        self.normalize = normalize

    def __call__(self, expert_counts, scores, assignments, offsets, logits, batch=None):
        k = assignments.shape[1]
        if logits.is_cuda:
            _k().moe_topk_gating(logits, k, self.normalize, scores, assignments, offsets, expert_counts)
            return scores
        probs = torch.softmax(logits.float(), dim=-1)
        topv, topi = torch.topk(probs, k, dim=-1)
        scores.copy_(topv / topv.sum(-1, keepdim=True) if self.normalize else topv)
        assignments.copy_(topi.to(torch.int32))
        flat = topi.reshape(-1)
        for idx in range(flat.numel()):
            e = int(flat[idx])
            offsets.view(-1)[idx] = expert_counts[e]
            expert_counts[e] += 1
        return scores


class MoEScatter(DSKernelBase):
    """Token rows -> their experts' (128-row padded) ranges. Fills ``expert_cumsum`` int32 [E+1] (padded range starts)
    and ``mapped_slots`` int32 [T,k]; ``expert_counts`` is re-zeroed for the next layer."""

    def __init__(self, dtype: torch.dtype, channels: int = 0):
        self._check_dtype(dtype)

    def __call__(self, moe_input, expert_cumsum, mapped_slots, activations, expert_counts, assignments, offsets):
        k = assignments.shape[1]
        if activations.is_cuda:
            _k().moe_scatter(activations, moe_input, assignments, offsets, expert_counts, expert_cumsum, mapped_slots, k)
            return moe_input, expert_cumsum, mapped_slots
        padded = (expert_counts + MOE_ROW_PAD - 1) // MOE_ROW_PAD * MOE_ROW_PAD
        expert_cumsum[0] = 0
        expert_cumsum[1:] = torch.cumsum(padded, 0).to(expert_cumsum.dtype)
        slots = expert_cumsum[:-1][assignments.long()] + offsets
        mapped_slots.copy_(slots.to(mapped_slots.dtype))
        moe_input[slots.reshape(-1).long()] = activations.repeat_interleave(k, dim=0)
        expert_counts.zero_()
        return moe_input, expert_cumsum, mapped_slots


class MoEGather(DSKernelBase):
    """``layer_output[t] += sum_j scores[t, j] * moe_output[mapped_slots[t, j]]`` (fp32 accumulator)."""

    def __init__(self, dtype: torch.dtype, channels: int = 0, normalize_scores: bool = False):
        self._check_dtype(dtype)

    def __call__(self, layer_output, moe_output, scores, mapped_slots, expert_counts=None):
        if moe_output.is_cuda:
            _k().moe_gather(moe_output, mapped_slots, scores, layer_output)
            return layer_output
        rows = moe_output[mapped_slots.reshape(-1).long()].float().view(*mapped_slots.shape, -1)
        layer_output += (rows * scores[..., None]).sum(1).to(layer_output.dtype)
        return layer_output


class AtomBuilder(DSKernelBase):
    """Work items of a ragged batch for the attention kernels: one ``(seq, q_row0, n_q, pos0)`` row per sequence."""

    def __init__(self):
        pass

    def __call__(self, items: torch.Tensor, seq_lens_new, seq_lens_seen, q_block_size: int = 0, kv_block_size: int = 0):
        row = 0
        for i, (n_new, seen) in enumerate(zip(seq_lens_new, seq_lens_seen)):
            items[i, 0], items[i, 1], items[i, 2], items[i, 3] = i, row, int(n_new), int(seen)
            row += int(n_new)
        return items, len(seq_lens_new)


class BlockedRotaryEmbeddings(DSKernelBase):
    """RoPE on q and k + append of k/v to the paged cache, from a packed ``[T, (n_q + 2 n_kv) * d]`` QKV tensor."""

    def __init__(self, head_size: int, n_q_heads: int, n_kv_heads: int, dtype: torch.dtype, rotary_dim: int, theta_base: float = 10000.0):
        self._check_dtype(dtype)
        self.d, self.n_q, self.n_kv, self.rot = head_size, n_q_heads, n_kv_heads, rotary_dim
        self.theta = theta_base
        self._inv = None

    def __call__(self, kv_cache_layer, qkv, q_out, tok_pos, tok_slot, block_size: int, inv_freq: Optional[torch.Tensor] = None):
        if inv_freq is None:
            if self._inv is None or self._inv.device != qkv.device:
                self._inv = ref.default_inv_freq(self.rot, self.theta, device=qkv.device) if self.rot else None
            inv_freq = self._inv
        if qkv.is_cuda:
            _k().rope_kv_append(qkv, q_out, kv_cache_layer, tok_pos, tok_slot, inv_freq, self.n_q, self.n_kv, self.d, self.rot, block_size)
            return q_out
        T = qkv.shape[0]
        q, k, v = qkv.view(T, -1, self.d).split([self.n_q, self.n_kv, self.n_kv], dim=1)
        if self.rot:
            q, k = ref.rope_rotate_half(q, tok_pos, inv_freq, self.rot), ref.rope_rotate_half(k, tok_pos, inv_freq, self.rot)
        q_out.copy_(q.reshape(q_out.shape))
        ref.kv_append(kv_cache_layer, k, v, tok_slot)
        return q_out


class LinearBlockedKVCopy(BlockedRotaryEmbeddings):
    """Append without rotation (models with learned / ALiBi positions)."""

    def __init__(self, head_size: int, n_q_heads: int, n_kv_heads: int, dtype: torch.dtype):
        super().__init__(head_size, n_q_heads, n_kv_heads, dtype, rotary_dim=0)


class BlockedFlashAttn(DSKernelBase):
    """Causal attention of a ragged batch against the paged KV cache."""

    def __init__(self, head_size: int, dtype: torch.dtype, n_q_heads: int = 0, n_kv_heads: int = 0, scale: Optional[float] = None):
        self._check_dtype(dtype)
        if head_size % 16:
            raise ValueError("Head size must be divisible by 16 (configured with {})".format(head_size))
        self.d, self.n_q, self.n_kv = head_size, n_q_heads, n_kv_heads
        self.scale = scale if scale is not None else ref.attention_scale(head_size)

    def __call__(self, out, q, kv_cache_layer, block_tables, items, batch=None):
        n_q = self.n_q or q.shape[1]
        n_kv = self.n_kv or kv_cache_layer.shape[-2]
        if q.is_cuda:
            if batch is None:
                raise ValueError("the CUDA path needs the engine's ragged batch (device block tables + work items)")
            self._ops(q).attention(q, out, kv_cache_layer, batch, n_q, n_kv, self.d, self.scale)
            return out
        return ref.paged_attention(q, kv_cache_layer, block_tables, items, n_q, n_kv, self.scale, out=out)
