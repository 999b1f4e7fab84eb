"""Process-group topology helpers: data / tensor(model) / expert / sequence parallel groups and the ZeRO++ hpZ
secondary groups. Same public functions as the reference (``deepspeed/utils/groups.py``), built on one rank-grid
routine: ranks are laid out ``[dp][sp][tp]`` with tp fastest, so a group of any axis is a strided slice.

All creators are collective (every rank must call them in the same order) and idempotent.
"""
from __future__ import annotations

from typing import Dict, List, Optional

from .. import comm as dist

_WORLD_GROUP = None
_DATA_PARALLEL_GROUP = None
_MODEL_PARALLEL_GROUP = None  # tensor parallel
_SEQUENCE_PARALLEL_GROUP = None
_SEQUENCE_DATA_PARALLEL_GROUP = None
_EXPERT_PARALLEL_GROUP: Dict[str, object] = {}
_EXPERT_DATA_PARALLEL_GROUP: Dict[str, object] = {}
_ZERO_PARAM_INTRA_PARALLEL_GROUP = None
_ALL_TO_ALL_GROUP: Dict[str, object] = {}
mpu = None
mesh_device = None
expert_tensor_parallel_world_size = 1


def _ensure_divisibility(numerator: int, denominator: int):
    assert numerator % denominator == 0, f"{numerator} is not divisible by {denominator}"


def rank_grid(world_size: int, tp: int = 1, sp: int = 1) -> List[List[List[int]]]:
    """``grid[d][s][t]`` = global rank; tp fastest, then sp, then dp."""
    _ensure_divisibility(world_size, tp * sp)
    dp = world_size // (tp * sp)
    return [[[(d * sp + s) * tp + t for t in range(tp)] for s in range(sp)] for d in range(dp)]


def _new_groups(rank_lists: List[List[int]]):
    """Create one group per rank list (collectively) and return the one containing this rank."""
    mine = None
    me = dist.get_rank()
    for ranks in rank_lists:
        g = dist.new_group(ranks)
        if me in ranks:
            mine = g
    return mine


def initialize(ep_size: int = 1, mpu_=None):
    """Deprecated entry kept for API parity: creates expert groups of size ``ep_size``."""
    global mpu
    mpu = mpu_
    _create_expert_and_data_parallel(ep_size)


def _create_model_parallel(model_parallel_size_: int, sequence_parallel_size_: int = 1):
    """Create tensor-parallel + data-parallel (+ sequence-parallel) groups over the world."""
    global _DATA_PARALLEL_GROUP, _MODEL_PARALLEL_GROUP, _SEQUENCE_PARALLEL_GROUP, _SEQUENCE_DATA_PARALLEL_GROUP
    ws = dist.get_world_size()
    tp, sp = min(model_parallel_size_, ws), sequence_parallel_size_
    grid = rank_grid(ws, tp, sp)
    dp = len(grid)
    _MODEL_PARALLEL_GROUP = _new_groups([grid[d][s] for d in range(dp) for s in range(sp)])
    _DATA_PARALLEL_GROUP = _new_groups([[grid[d][s][t] for d in range(dp)] for s in range(sp) for t in range(tp)])
    if sp > 1:
        _SEQUENCE_PARALLEL_GROUP = _new_groups([[grid[d][s][t] for s in range(sp)] for d in range(dp) for t in range(tp)])
        _SEQUENCE_DATA_PARALLEL_GROUP = _new_groups([[grid[d][s][t] for d in range(dp) for s in range(sp)]
                                                     for t in range(tp)])
    return _DATA_PARALLEL_GROUP, _MODEL_PARALLEL_GROUP


def expert_group_ranks(world_size: int, ep_size: int, tp: int = 1, use_data_before_expert_parallel: bool = False):
    """(expert-parallel rank lists, expert-data-parallel rank lists) for ``world_size`` ranks with tp fastest.
    Expert parallelism runs over data-parallel ranks: consecutive dp ranks form an EP group (or strided ones when
    ``use_data_before_expert_parallel``)."""
    _ensure_divisibility(world_size, tp)
    dp = world_size // tp
    _ensure_divisibility(dp, ep_size)
    ep_groups, edp_groups = [], []
    for t in range(tp):
        dp_ranks = [d * tp + t for d in range(dp)]
        if use_data_before_expert_parallel:
            stride = dp // ep_size
            eps = [dp_ranks[i::stride] for i in range(stride)]
        else:
            eps = [dp_ranks[i:i + ep_size] for i in range(0, dp, ep_size)]
        ep_groups += eps
        edp_groups += [list(x) for x in zip(*eps)]
    return ep_groups, edp_groups


def _create_expert_and_data_parallel(expert_parallel_size_: int, use_data_before_expert_parallel_: bool = False):
    ws = dist.get_world_size()
    tp = _get_model_parallel_world_size() if (_MODEL_PARALLEL_GROUP is not None or mpu is not None) else 1
    name = f"ep_size_{expert_parallel_size_}"
    if name in _EXPERT_PARALLEL_GROUP:
        return
    ep, edp = expert_group_ranks(ws, expert_parallel_size_, tp, use_data_before_expert_parallel_)
    _EXPERT_PARALLEL_GROUP[name] = _new_groups(ep)
    _EXPERT_DATA_PARALLEL_GROUP[name] = _new_groups(edp)


def _create_expert_data_and_model_parallel(expert_parallel_size_: int, mpu_, use_data_before_expert_parallel_=False):
    global mpu, expert_tensor_parallel_world_size
    mpu = mpu_
    expert_tensor_parallel_world_size = mpu_.get_model_parallel_world_size()
    _create_expert_and_data_parallel(expert_parallel_size_, use_data_before_expert_parallel_)


def _get_max_expert_size() -> int:
    assert _EXPERT_PARALLEL_GROUP, "no expert parallel group has been created"
    return max(int(k.split("_")[-1]) for k in _EXPERT_PARALLEL_GROUP)


def _get_max_expert_size_name() -> str:
    return f"ep_size_{_get_max_expert_size()}"


def _get_max_expert_parallel_group():
    return _EXPERT_PARALLEL_GROUP[_get_max_expert_size_name()]


def _get_expert_parallel_group(group_name: str):
    assert group_name in _EXPERT_PARALLEL_GROUP, f"expert parallel group {group_name} is not initialized"
    return _EXPERT_PARALLEL_GROUP[group_name]


def _get_expert_parallel_group_dict():
    return _EXPERT_PARALLEL_GROUP


def _get_expert_data_parallel_group(group_name: str):
    assert group_name in _EXPERT_DATA_PARALLEL_GROUP, f"expert data parallel group {group_name} is not initialized"
    return _EXPERT_DATA_PARALLEL_GROUP[group_name]


def _get_expert_data_parallel_group_dict():
    return _EXPERT_DATA_PARALLEL_GROUP


def _clone_world_group():
    global _WORLD_GROUP
    if _WORLD_GROUP is None:
        _WORLD_GROUP = dist.new_group(ranks=list(range(dist.get_world_size())))
    return _WORLD_GROUP


def _get_data_parallel_group():
    if mesh_device is not None:
        return mesh_device.get_group(mesh_dim="data_parallel")
    if mpu is not None:
        return mpu.get_data_parallel_group()
    if _DATA_PARALLEL_GROUP is not None:
        return _DATA_PARALLEL_GROUP
    return _clone_world_group()


def _get_broadcast_src_rank() -> int:
    return dist.get_global_rank(_get_sequence_data_parallel_group(), 0)


def _get_model_parallel_group():
    if mpu is not None:
        return mpu.get_model_parallel_group()
    return _MODEL_PARALLEL_GROUP


def _get_model_parallel_world_size() -> int:
    if mpu is not None:
        return mpu.get_model_parallel_world_size()
    return dist.get_world_size(group=_MODEL_PARALLEL_GROUP) if _MODEL_PARALLEL_GROUP is not None else 1


def _get_model_parallel_rank() -> int:
    if mpu is not None:
        return mpu.get_model_parallel_rank()
    return dist.get_rank(group=_MODEL_PARALLEL_GROUP) if _MODEL_PARALLEL_GROUP is not None else 0


# ---- the mpu-style public names (the module itself can be handed to code that expects an ``mpu``; ref: groups.py:60-130) ----
_tp_world_size_override = None
_tp_rank_override = None


def get_model_parallel_group():
    return _get_model_parallel_group()


def get_data_parallel_group():
    return _get_data_parallel_group()


def set_tensor_model_parallel_world_size(world_size) -> None:
    """Pin the value ``get_tensor_model_parallel_world_size`` reports (``None`` = ask the group again)."""
    global _tp_world_size_override
    _tp_world_size_override = world_size


def set_tensor_model_parallel_rank(rank) -> None:
    global _tp_rank_override
    _tp_rank_override = rank


def get_model_parallel_world_size() -> int:
    return _tp_world_size_override if _tp_world_size_override is not None else _get_model_parallel_world_size()


def get_model_parallel_rank() -> int:
    return _tp_rank_override if _tp_rank_override is not None else _get_model_parallel_rank()


def get_tensor_model_parallel_src_rank() -> int:
    """Global rank of the first member of this rank's tensor-parallel group."""
    ws = get_model_parallel_world_size()
    return (dist.get_rank() // ws) * ws if dist.is_initialized() else 0


def get_data_parallel_world_size() -> int:
    return _get_data_parallel_world_size()


def get_data_parallel_rank() -> int:
    return _get_data_parallel_rank()


get_tensor_model_parallel_group = _get_model_parallel_group
get_tensor_model_parallel_world_size = get_model_parallel_world_size
get_tensor_model_parallel_rank = get_model_parallel_rank


def _get_data_parallel_world_size() -> int:
    return dist.get_world_size(group=_get_data_parallel_group())


def _get_data_parallel_rank() -> int:
    return dist.get_rank(group=_get_data_parallel_group())


def _get_expert_parallel_world_size(group_name: str) -> int:
    return dist.get_world_size(group=_get_expert_parallel_group(group_name))


def _get_expert_parallel_rank(group_name: str) -> int:
    return dist.get_rank(group=_get_expert_parallel_group(group_name))


def _get_expert_data_parallel_world_size(group_name: str) -> int:
    return dist.get_world_size(group=_get_expert_data_parallel_group(group_name))


def _get_expert_data_parallel_rank(group_name: str) -> int:
    return dist.get_rank(group=_get_expert_data_parallel_group(group_name))


def _get_expert_parallel_src_rank(group_name: str) -> int:
    return dist.get_global_rank(_get_expert_parallel_group(group_name), 0)


def _get_sequence_parallel_group():
    if mpu is not None and hasattr(mpu, "get_sequence_parallel_group"):
        return mpu.get_sequence_parallel_group()
    return _SEQUENCE_PARALLEL_GROUP


def _get_sequence_parallel_world_size() -> int:
    g = _get_sequence_parallel_group()
    return dist.get_world_size(group=g) if g is not None else 1


def _get_sequence_parallel_rank() -> int:
    g = _get_sequence_parallel_group()
    return dist.get_rank(group=g) if g is not None else 0


def _get_sequence_data_parallel_group():
    if mpu is not None and hasattr(mpu, "get_sequence_data_parallel_group"):
        return mpu.get_sequence_data_parallel_group()
    if _SEQUENCE_DATA_PARALLEL_GROUP is not None:
        return _SEQUENCE_DATA_PARALLEL_GROUP
    return _get_data_parallel_group()


def _get_sequence_data_parallel_world_size() -> int:
    return dist.get_world_size(group=_get_sequence_data_parallel_group())


def _get_sequence_data_parallel_rank() -> int:
    return dist.get_rank(group=_get_sequence_data_parallel_group())


def _create_zero_param_parallel_group(group_size: int):
    """ZeRO++ hpZ: secondary parameter partition inside groups of ``group_size`` consecutive ranks (one node)."""
    global _ZERO_PARAM_INTRA_PARALLEL_GROUP
    if _ZERO_PARAM_INTRA_PARALLEL_GROUP is not None:
        return _ZERO_PARAM_INTRA_PARALLEL_GROUP
    ws = dist.get_world_size()
    _ensure_divisibility(ws, group_size)
    _ZERO_PARAM_INTRA_PARALLEL_GROUP = _new_groups([list(range(i, i + group_size)) for i in range(0, ws, group_size)])
    return _ZERO_PARAM_INTRA_PARALLEL_GROUP


def _get_zero_param_intra_parallel_group():
    return _ZERO_PARAM_INTRA_PARALLEL_GROUP


def _zero_param_parallel_is_initialized() -> bool:
    return _ZERO_PARAM_INTRA_PARALLEL_GROUP is not None


def _get_zero_param_intra_parallel_rank_in_mygroup() -> int:
    return dist.get_rank(group=_ZERO_PARAM_INTRA_PARALLEL_GROUP)


def _get_zero_param_intra_parallel_group_world_size() -> int:
    return dist.get_world_size(group=_ZERO_PARAM_INTRA_PARALLEL_GROUP)


def _get_zero_param_intra_parallel_group_ranks() -> Optional[List[int]]:
    g = _ZERO_PARAM_INTRA_PARALLEL_GROUP
    return dist.get_all_ranks_from_group(g) if g is not None else None


def _get_local_all_to_all_group():
    """Intra-node / inter-node groups for the hierarchical (quantized) all-to-all reduction of ZeRO++ qgZ."""
    if _ALL_TO_ALL_GROUP:
        return _ALL_TO_ALL_GROUP
    import os
    ws = dist.get_world_size()
    local = int(os.environ.get("LOCAL_WORLD_SIZE", ws))
    local = local if ws % local == 0 else ws
    nodes = ws // local
    for n in range(nodes):
        g = dist.new_group(list(range(n * local, (n + 1) * local)))
        _ALL_TO_ALL_GROUP[f"local_{n}"] = g
    for i in range(local):
        g = dist.new_group([i + n * local for n in range(nodes)])
        _ALL_TO_ALL_GROUP[f"global_{i}"] = g
    return _ALL_TO_ALL_GROUP


def reset():
    """Forget every cached group (tests)."""
    global _WORLD_GROUP, _DATA_PARALLEL_GROUP, _MODEL_PARALLEL_GROUP, _SEQUENCE_PARALLEL_GROUP
    global _SEQUENCE_DATA_PARALLEL_GROUP, _ZERO_PARAM_INTRA_PARALLEL_GROUP, mpu, mesh_device
    _WORLD_GROUP = _DATA_PARALLEL_GROUP = _MODEL_PARALLEL_GROUP = _SEQUENCE_PARALLEL_GROUP = None
    _SEQUENCE_DATA_PARALLEL_GROUP = _ZERO_PARAM_INTRA_PARALLEL_GROUP = None
    _EXPERT_PARALLEL_GROUP.clear()
    _EXPERT_DATA_PARALLEL_GROUP.clear()
    _ALL_TO_ALL_GROUP.clear()
    mpu = mesh_device = None
