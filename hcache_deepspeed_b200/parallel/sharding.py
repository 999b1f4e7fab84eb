"""Tensor-parallel sharding rules (Megatron style) for the serving path.

Same *rules* as the reference (``deepspeed/inference/v2/model_implementations/sharding/
{utils,qkv,attn,attn_out,mlp,unembed,embedding}.py``), restated over head-index lists:

* QKV: column split by head. MHA heads may split unevenly (first ``n % tp`` ranks get one extra).
  GQA with ``n_kv >= tp`` needs ``n_kv % tp == 0`` and gives each rank whole groups; with
  ``n_kv < tp`` (needs ``tp % n_kv == 0``) each KV head is replicated over ``tp / n_kv`` ranks and
  its query group is split among them.
* attention output / MLP down projection: row (input-dim) split matching the producing shard;
  biases live on rank 0 only.
* MLP up projection: column split with granularity 32 (gated: the gate and up halves are split
  identically so they can be interleaved per rank).
* unembedding: vocabulary split; embedding: replicated.
"""
from __future__ import annotations

from typing import List, Optional, Tuple

import torch

DEFAULT_SHARD_GRANULARITY = 32


def get_shard_endpoints(dim_size: int, shard_rank: int, num_shards: int,
                        granularity: int = DEFAULT_SHARD_GRANULARITY) -> Tuple[int, int]:
    """[start, end) of ``shard_rank``'s slice of a dimension, aligned to ``granularity``; the first
    ``chunks % num_shards`` ranks receive one extra chunk."""
    if dim_size % granularity != 0:
        raise ValueError(f"Dimension size {dim_size} must be divisible by granularity {granularity}")
    chunks = dim_size // granularity
    base, rem = divmod(chunks, num_shards)
    start = shard_rank * base + min(shard_rank, rem)
    end = start + base + (1 if shard_rank < rem else 0)
    return start * granularity, end * granularity


def local_head_indices(shard_rank: int, num_shards: int, n_heads_q: int,
                       n_heads_kv: Optional[int] = None) -> Tuple[List[int], List[int]]:
    """Global indices of the query heads and KV heads owned by a rank."""
    if n_heads_q < num_shards:
        raise ValueError("There must be at least as many attention heads as there are shards.")
    if n_heads_kv is None or n_heads_kv == n_heads_q:
        s, e = get_shard_endpoints(n_heads_q, shard_rank, num_shards, granularity=1)
        heads = list(range(s, e))
        return heads, list(heads)
    if n_heads_q % n_heads_kv != 0:
        raise ValueError("Must be an even ratio between query and key/value heads.")
    ratio = n_heads_q // n_heads_kv
    if n_heads_kv >= num_shards:
        if n_heads_kv % num_shards != 0:
            raise ValueError("If parallelizing groups, must be able to evenly distribute them.")
        per = n_heads_kv // num_shards
        kv = list(range(shard_rank * per, (shard_rank + 1) * per))
        q = [g * ratio + i for g in kv for i in range(ratio)]
        return q, kv
    if num_shards % n_heads_kv != 0:
        raise ValueError("If splitting a group across multiple shards, we must be able to distribute the groups evenly.")
    ranks_per_group = num_shards // n_heads_kv
    group = shard_rank // ranks_per_group
    s, e = get_shard_endpoints(ratio, shard_rank % ranks_per_group, ranks_per_group, granularity=1)
    return [group * ratio + i for i in range(s, e)], [group]


def get_local_heads(shard_rank: int, num_shards: int, n_heads_q: int, n_heads_kv: Optional[int] = None) -> Tuple[int, int]:
    q, kv = local_head_indices(shard_rank, num_shards, n_heads_q, n_heads_kv)
    return len(q), len(kv)


def _rows_of_heads(heads: List[int], head_size: int, base: int = 0) -> torch.Tensor:
    idx = torch.tensor(heads, dtype=torch.long)[:, None] * head_size + torch.arange(head_size)[None, :] + base
    return idx.reshape(-1)


def shard_qkv_param(param: Optional[torch.Tensor], shard_rank: int, num_shards: int, head_size: int,
                    n_heads_q: int, n_heads_kv: Optional[int] = None) -> Optional[torch.Tensor]:
    """Slice a fused ``[(n_q + 2 n_kv) * head_size, ...]`` QKV weight/bias (rows: Q heads, K heads, V heads)."""
    n_kv = n_heads_kv if n_heads_kv is not None else n_heads_q
    q_heads, kv_heads = local_head_indices(shard_rank, num_shards, n_heads_q, n_heads_kv)
    if param is None:
        return None
    if param.shape[0] != head_size * (n_heads_q + 2 * n_kv):
        raise ValueError(f"QKV param has {param.shape[0]} rows, expected {head_size * (n_heads_q + 2 * n_kv)}")
    if num_shards == 1:
        return param
    rows = torch.cat([
        _rows_of_heads(q_heads, head_size),
        _rows_of_heads(kv_heads, head_size, base=n_heads_q * head_size),
        _rows_of_heads(kv_heads, head_size, base=(n_heads_q + n_kv) * head_size),
    ])
    return param[rows.to(param.device)].contiguous()


def shard_attn_out_param(param: Optional[torch.Tensor], shard_rank: int, num_shards: int, head_size: int,
                         n_heads_q: Optional[int] = None, n_heads_kv: Optional[int] = None) -> Optional[torch.Tensor]:
    """Row-parallel slice of the attention output projection ``[H, n_q * head_size]`` (columns follow the
    rank's query heads). A 1-D bias survives on rank 0 only."""
    if param is None:
        return None
    if param.dim() == 1:
        return param if shard_rank == 0 else None
    if num_shards == 1:
        return param
    n_q = n_heads_q if n_heads_q is not None else param.shape[1] // head_size
    q_heads, _ = local_head_indices(shard_rank, num_shards, n_q, n_heads_kv)
    cols = _rows_of_heads(q_heads, head_size)
    return param[:, cols.to(param.device)].contiguous()


def shard_mlp_1_param(param: Optional[torch.Tensor], shard_rank: int, num_shards: int, gated: bool = False,
                      is_moe: bool = False) -> Optional[torch.Tensor]:
    """Column split of the MLP up projection. ``gated``: rows are [gate; up] halves, each split identically."""
    if param is None:
        return None
    if num_shards == 1:
        return param
    dim = 1 if is_moe else 0
    n = param.shape[dim]
    if gated:
        half = n // 2
        s, e = get_shard_endpoints(half, shard_rank, num_shards)
        idx = torch.cat([torch.arange(s, e), torch.arange(half + s, half + e)])
        return param.index_select(dim, idx.to(param.device)).contiguous()
    s, e = get_shard_endpoints(n, shard_rank, num_shards)
    return param.narrow(dim, s, e - s).contiguous()


def shard_mlp_2_param(param: Optional[torch.Tensor], shard_rank: int, num_shards: int,
                      is_moe: bool = False) -> Optional[torch.Tensor]:
    """Row-parallel slice of the MLP down projection ``[H, I]``; 1-D bias on rank 0 only."""
    if param is None:
        return None
    bias_dims = 2 if is_moe else 1
    if param.dim() == bias_dims:
        return param if shard_rank == 0 else None
    if num_shards == 1:
        return param
    s, e = get_shard_endpoints(param.shape[-1], shard_rank, num_shards)
    return param[..., s:e].contiguous()


def sharded_intermediate_dim(intermediate: int, num_shards: int, shard_rank: int) -> int:
    s, e = get_shard_endpoints(intermediate, shard_rank, num_shards)
    return e - s


def shard_unembed_param(param: torch.Tensor, shard_rank: int, num_shards: int) -> torch.Tensor:
    """Vocabulary split (granularity 1, as the reference: ``sharding/unembed.py:12-41``)."""
    if num_shards == 1:
        return param
    s, e = get_shard_endpoints(param.shape[0], shard_rank, num_shards, granularity=1)
    return param[s:e].contiguous()


def sharded_unembed_dim(vocab_size: int, shard_rank: int, num_shards: int) -> int:
    s, e = get_shard_endpoints(vocab_size, shard_rank, num_shards, granularity=1)
    return e - s


def shard_embedding_param(param: torch.Tensor, shard_rank: int, num_shards: int) -> torch.Tensor:
    """Embeddings are replicated (``inference_transformer_base.py:239-245``)."""
    return param


def qkv_out_features(in_features: int, shard_rank: int, num_shards: int, head_size: int, n_heads_q: Optional[int] = None,
                     n_heads_kv: Optional[int] = None) -> int:
    n_q = n_heads_q if n_heads_q is not None else in_features // head_size
    lq, lkv = get_local_heads(shard_rank, num_shards, n_q, n_heads_kv)
    return head_size * (lq + 2 * lkv)


def attn_out_in_features(out_features: int, shard_rank: int, num_shards: int, head_size: int,
                         n_heads_q: Optional[int] = None, n_heads_kv: Optional[int] = None) -> int:
    n_q = n_heads_q if n_heads_q is not None else out_features // head_size
    lq, _ = get_local_heads(shard_rank, num_shards, n_q, n_heads_kv)
    return head_size * lq
