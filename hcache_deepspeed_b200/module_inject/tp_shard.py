"""Shard-size arithmetic for tensor parallelism with head counts that do not divide evenly: sizes are assigned per KV
head so a rank never owns a fraction of a head. Reference: ``deepspeed/module_inject/tp_shard.py:9-87``."""
from __future__ import annotations

from typing import List, Optional

num_kv_heads: Optional[int] = None
n_embd: Optional[int] = None
tp_grain_size = 1


def set_num_kv_heads(n: Optional[int]):
    global num_kv_heads
    num_kv_heads = n


def set_num_attention_heads(n):
    global num_attention_heads
    num_attention_heads = n


def set_n_embd(n):
    global n_embd
    n_embd = n


def set_tp_grain_size(n):
    global tp_grain_size
    tp_grain_size = max(1, int(n))


def get_num_kv_heads():
    return num_kv_heads


def get_num_attention_heads():
    return globals().get("num_attention_heads")


def get_n_embd():
    return n_embd


def get_shard_size(total_size: int, mp_size: int, name: Optional[str] = None, rank: Optional[int] = None) -> int:
    """Rows / columns owned by ``rank`` of ``mp_size``. With a known KV-head count (and when the dimension is a multiple
    of it) the split follows whole heads: the first ``heads % mp`` ranks get one extra head. The LM head and the
    embedding are split on the grain size instead."""
    if rank is None:
        from .. import comm as dist
        rank = dist.get_rank() % mp_size if dist.is_initialized() else 0
    last_linear = ("lm_head", "embed_out")
    if num_kv_heads is not None and total_size % num_kv_heads == 0 and (name is None or not any(l in name for l in last_linear)):
        my_slices = num_kv_heads // mp_size + (1 if rank < (num_kv_heads % mp_size) else 0)
        return total_size * my_slices // num_kv_heads
    grains = total_size // tp_grain_size
    if total_size >= tp_grain_size and total_size % tp_grain_size == 0:
        return (grains // mp_size + (1 if rank < (grains % mp_size) else 0)) * tp_grain_size
    return total_size // mp_size + (1 if rank < (total_size % mp_size) else 0)


def get_shard_size_list(total_size: int, mp_size: int, name: Optional[str] = None) -> List[int]:
    return [get_shard_size(total_size, mp_size, name, r) for r in range(mp_size)]
