"""ALiBi helpers for auto-TP of BLOOM-style models (``deepspeed/module_inject/auto_tp_model_utils.py``): each rank
needs the ALiBi slopes of ITS heads only."""
from __future__ import annotations

import math

import torch

from .. import comm as dist


def _slopes(n_heads: int) -> torch.Tensor:
    closest = 2 ** math.floor(math.log2(n_heads))
    base = 2.0 ** (-(2.0 ** -(math.log2(closest) - 3)))
    s = torch.pow(torch.tensor(base), torch.arange(1, 1 + closest, dtype=torch.float32))
    if closest != n_heads:
        extra_base = 2.0 ** (-(2.0 ** -(math.log2(2 * closest) - 3)))
        n_rem = min(closest, n_heads - closest)
        s = torch.cat([s, torch.pow(torch.tensor(extra_base), torch.arange(1, 1 + 2 * n_rem, 2, dtype=torch.float32))])
    return s


def build_bloom_alibi_tensor(attention_mask: torch.Tensor, num_heads: int, dtype: torch.dtype, mp_size=None,
                             mp_rank=None) -> torch.Tensor:
    """``[batch * local_heads, 1, seq]`` ALiBi bias; under tensor parallelism only this rank's head slice (the same
    head split as the fused-QKV slicing: ``tp_shard.get_shard_size_list``)."""
    from .tp_shard import get_shard_size_list
    batch, seq = attention_mask.shape
    slopes = _slopes(num_heads).to(attention_mask.device)
    pos = ((attention_mask.cumsum(dim=-1) - 1) * attention_mask)[:, None, :]
    alibi = slopes[None, :, None] * pos                       # [batch, heads, seq]
    if mp_size is None and dist.is_initialized():
        mp_size, mp_rank = dist.get_world_size(), dist.get_rank()
    if mp_size and mp_size > 1:
        sizes = get_shard_size_list(num_heads, mp_size)
        start = sum(sizes[:mp_rank])
        alibi = alibi[:, start:start + sizes[mp_rank]]
        num_heads = sizes[mp_rank]
    return alibi.reshape(batch * num_heads, 1, seq).to(dtype)


def slice_heads(t: torch.Tensor, dim: int, rank: int, world: int) -> torch.Tensor:
    """This rank's heads along ``dim`` of a per-head tensor (same uneven split as the attention sharding)."""
    if t is None or world <= 1:
        return t
    from .tp_shard import get_shard_size_list
    sizes = get_shard_size_list(t.shape[dim], world)
    return t.narrow(dim, sum(sizes[:rank]), sizes[rank])


def _world():
    return (dist.get_rank(), dist.get_world_size()) if dist.is_initialized() else (0, 1)


def get_alibi_mask(self, tensor, seq_length_with_past):
    """Baichuan-style cached ALiBi mask ``[heads, seq, seq]``, sliced to this rank's heads; installed as the class's
    ``get_alibi_mask`` with the original kept as ``get_alibi_mask_orig`` (ref: auto_tp_model_utils.py:63)."""
    mask = self.get_alibi_mask_orig(tensor, seq_length_with_past)
    return slice_heads(mask, 0, *_world())


def build_mpt_alibi_tensor(self, *args, **kwargs):
    """MPT's ``[1, heads, 1, seq]`` ALiBi tensor for this rank's heads (original kept as ``build_mpt_alibi_tensor_orig``)."""
    return slice_heads(self.build_mpt_alibi_tensor_orig(*args, **kwargs), 1, *_world())


def build_mpt_atten_bias_tensor(self, device, dtype, attention_mask=None, prefix_mask=None, sequence_id=None):
    """MPT's ``_attn_bias`` for this rank's heads (original kept as ``_attn_bias_orig``)."""
    bias, mask = self._attn_bias_orig(device, dtype, attention_mask=attention_mask, prefix_mask=prefix_mask,
                                      sequence_id=sequence_id)
    return (slice_heads(bias, 1, *_world()) if bias is not None else None), mask
