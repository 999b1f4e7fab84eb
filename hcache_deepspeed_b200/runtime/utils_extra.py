"""More of the reference's ``deepspeed/runtime/utils.py`` helper surface (re-exported by ``runtime/utils.py``):
device moves of nested structures, norms over parameters / gradients, zero counting, a tensor partitioned over a
process group (what pipeline stages exchange under ``partition_activations``), memory reporting, ZeRO all-gathers
over per-group flat buffers, and a few small classes."""
from __future__ import annotations

import contextlib
import math
from typing import Callable, Dict, List, Optional, Sequence, Union

import torch

from .. import comm as dist
from ..utils.logging import logger


class DummyOptim:
    """Presents the model parameters as one param group: lets ZeRO-3 wrap a model that is trained without an optimizer
    (parameter partitioning only; ref: runtime/utils.py:39)."""

    def __init__(self, params):
        self.param_groups = [{"params": list(params)}]
        self.state = {}

    def step(self, closure=None):
        return None

    def zero_grad(self, set_to_none: bool = True):
        for p in self.param_groups[0]["params"]:
            p.grad = None

    def state_dict(self):
        return {}

    def load_state_dict(self, sd):
        return None


@contextlib.contextmanager
def noop_context():
    yield


# ---- CUDA-graph wrapped helper calls (ref: graph_process) ------------------------------------------------------------------
graph_cache: Dict[str, object] = {}


def graph_process(replay_first_step: bool, func: Callable, *args, **kwargs):
    """Run ``func(*args, **kwargs)`` through a CUDA graph captured on its first call (keyed by the function's name); the
    arguments must keep their addresses between calls. On CPU the function simply runs."""
    if not (torch.cuda.is_available() and any(torch.is_tensor(a) and a.is_cuda for a in list(args) + list(kwargs.values()))):
        return func(*args, **kwargs)
    key = func.__name__
    if key not in graph_cache:
        side = torch.cuda.Stream()
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):
            func(*args, **kwargs)                          # warm-up outside the capture
        torch.cuda.current_stream().wait_stream(side)
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            func(*args, **kwargs)
        graph_cache[key] = g
        if replay_first_step:
            g.replay()
    else:
        graph_cache[key].replay()


# ---- nested structures --------------------------------------------------------------------------------------------------------
def _map_tensors(item, fn: Callable, criterion: Callable):
    if torch.is_tensor(item):
        return fn(item) if criterion(item) else item
    if isinstance(item, list):
        return [_map_tensors(v, fn, criterion) for v in item]
    if isinstance(item, tuple):
        return tuple(_map_tensors(v, fn, criterion) for v in item)
    if isinstance(item, dict):
        return {k: _map_tensors(v, fn, criterion) for k, v in item.items()}
    return item


def copy_to_device(item, device, criterion_func: Callable = lambda t: True):
    """A copy of every tensor in a nested list / tuple / dict that satisfies ``criterion_func``, on ``device``."""
    return _map_tensors(item, lambda t: t.to(device), criterion_func)


def move_to_device(item, device, criterion_func: Callable = lambda t: True):
    """Like ``copy_to_device`` but in place: the tensors keep their identity, their storage moves."""

    def move(t):
        t.data = t.data.to(device)
        return t

    return _map_tensors(item, move, criterion_func)


def compare_tensors_in_structures(inputs1: Union[List, Dict], inputs2: Union[List, Dict]) -> bool:
    """Deep equality of two lists / dicts that may hold tensors (exact values, same devices)."""
    if type(inputs1) is not type(inputs2):
        return False
    if isinstance(inputs1, (list, tuple)):
        return len(inputs1) == len(inputs2) and all(compare_tensors_in_structures(a, b) if isinstance(a, (list, tuple, dict))
                                                    else _leaf_equal(a, b) for a, b in zip(inputs1, inputs2))
    if isinstance(inputs1, dict):
        return inputs1.keys() == inputs2.keys() and all(
            compare_tensors_in_structures(inputs1[k], inputs2[k]) if isinstance(inputs1[k], (list, tuple, dict))
            else _leaf_equal(inputs1[k], inputs2[k]) for k in inputs1)
    return _leaf_equal(inputs1, inputs2)


def _leaf_equal(a, b) -> bool:
    if torch.is_tensor(a) and torch.is_tensor(b):
        return a.device == b.device and a.shape == b.shape and bool(torch.equal(a, b))
    if torch.is_tensor(a) or torch.is_tensor(b):
        return False
    return a == b


# ---- norms / counts -------------------------------------------------------------------------------------------------------------
def _mp_reduce(value: float, op, mpu, device) -> float:
    if mpu is None or not dist.is_initialized():
        return value
    t = torch.tensor([float(value)], dtype=torch.float32, device=device)
    dist.all_reduce(t, op=op, group=mpu.get_model_parallel_group())
    return float(t[0])


def _counts_here(p, mpu) -> bool:
    """Tensor-parallel replicas of a parameter are counted on model-parallel rank 0 only."""
    if mpu is None:
        return True
    return bool(getattr(p, "model_parallel", False) or getattr(p, "tensor_model_parallel", False)) or \
        mpu.get_model_parallel_rank() == 0


def _norm_of(tensors: List[torch.Tensor], norm_type: float, mpu, device) -> float:
    if norm_type == math.inf:
        local = max((float(t.abs().max()) for t in tensors if t.numel()), default=0.0)
        total = _mp_reduce(local, dist.ReduceOp.MAX, mpu, device)
    else:
        local = sum(float(t.float().norm(norm_type)) ** norm_type for t in tensors if t.numel())
        total = _mp_reduce(local, dist.ReduceOp.SUM, mpu, device) ** (1.0 / norm_type)
    return -1.0 if (math.isinf(total) or math.isnan(total)) else total


def get_weight_norm(parameters, norm_type: float = 2, mpu=None) -> float:
    """Norm of the parameter VALUES over all model-parallel ranks (-1 for inf / nan)."""
    ps = [parameters] if torch.is_tensor(parameters) else list(parameters)
    dev = ps[0].device if ps else torch.device("cpu")
    return _norm_of([p.data for p in ps if _counts_here(p, mpu)], float(norm_type), mpu, dev)


def get_flattened_grad_norm(parameters, norm_type: float = 2, mpu=None, grad_norm_mask=None) -> float:
    """Gradient norm of (flat) parameters; ``grad_norm_mask[i]`` lists ``[start, end)`` ranges of parameter ``i``'s flat
    gradient to leave out (elements that are replicated on other model-parallel ranks)."""
    ps = [parameters] if torch.is_tensor(parameters) else list(parameters)
    ps = [p for p in ps if p.grad is not None]
    dev = ps[0].device if ps else torch.device("cpu")
    grads = []
    for i, p in enumerate(ps):
        g = p.grad.data.reshape(-1)
        ranges = grad_norm_mask[i] if grad_norm_mask is not None and i < len(grad_norm_mask) else None
        if ranges is not None and len(ranges):
            keep = torch.ones(g.numel(), dtype=torch.bool, device=g.device)
            for lo, hi in (ranges.tolist() if torch.is_tensor(ranges) else ranges):
                keep[int(lo):int(hi)] = False
            g = g[keep]
        grads.append(g)
    return _norm_of(grads, float(norm_type), mpu, dev)


def get_grad_zeros(parameters, mpu=None) -> int:
    """Number of gradient elements that are exactly zero, over all model-parallel ranks."""
    ps = [parameters] if torch.is_tensor(parameters) else list(parameters)
    ps = [p for p in ps if p.grad is not None]
    dev = ps[0].device if ps else torch.device("cpu")
    local = sum(int(p.grad.numel() - torch.count_nonzero(p.grad)) for p in ps if _counts_here(p, mpu))
    return int(_mp_reduce(float(local), dist.ReduceOp.SUM, mpu, dev))


def get_norm_with_moe_layers(non_expert_norm: float, mpu, expert_tensors: Dict[str, List[torch.Tensor]], norm_type: float = 2) -> float:
    """Global norm from the norm of the shared parameters and the gradients of each expert group (each group's norm is
    summed over its expert-parallel group, since ranks hold different experts); -1 if any part is inf / nan."""
    from ..utils import groups
    parts = [float(non_expert_norm)]
    for name, tensors in expert_tensors.items():
        tensors = [t for t in tensors if t is not None and t.numel()]
        dev = tensors[0].device if tensors else torch.device("cpu")
        ep = groups._get_expert_parallel_group(name) if dist.is_initialized() else None
        if norm_type == math.inf:
            v = torch.tensor([max((float(t.abs().max()) for t in tensors), default=0.0)], device=dev)
            if ep is not None:
                dist.all_reduce(v, op=dist.ReduceOp.MAX, group=ep)
            parts.append(float(v[0]))
        else:
            v = torch.tensor([sum(float(t.float().norm(norm_type)) ** norm_type for t in tensors)], device=dev)
            if ep is not None:
                dist.all_reduce(v, op=dist.ReduceOp.SUM, group=ep)
            parts.append(_mp_reduce(float(v[0]), dist.ReduceOp.SUM, mpu, dev) ** (1.0 / norm_type))
    if any(p == -1 or math.isinf(p) or math.isnan(p) for p in parts):
        return -1
    return max(parts) if norm_type == math.inf else sum(p ** norm_type for p in parts) ** (1.0 / norm_type)


def get_norm_with_moe_layers_fast(all_groups_norm, group):
    """Scale a norm computed over expert parameters for the size of ``group`` they are replicated over."""
    scale = float(dist.get_world_size(group=group)) if dist.is_initialized() else 1.0
    return all_groups_norm * (1.0 / math.sqrt(scale))


def mask_nan_or_inf_with_val_inplace(input: torch.Tensor, device=None, val: float = -1.0) -> None:
    bad = input.isnan().logical_or(input.isinf())
    input.masked_fill_(bad, val)


def get_inactive_params(param_list):
    from .zero.partition_parameters import ZeroParamStatus
    return [p for p in param_list if hasattr(p, "ds_id") and p.ds_status == ZeroParamStatus.NOT_AVAILABLE]


# ---- a tensor split over a process group ----------------------------------------------------------------------------------
class PartitionedTensor:
    """Each rank of ``group`` keeps one contiguous slice of a tensor's flat data; ``full()`` reassembles it. A pipeline
    stage under ``partition_activations`` sends ``to_meta()`` + ``data()`` instead of the whole activation, and the
    receiving stage rebuilds it with ``from_meta(...).full()`` (ref: runtime/utils.py:624)."""

    def __init__(self, tensor: torch.Tensor, group, partition_meta=None):
        from .utils import partition_uniform
        self.group = group
        self.num_parts = dist.get_world_size(group=group) if dist.is_initialized() else 1
        self.rank = dist.get_rank(group=group) if dist.is_initialized() else 0
        self.orig_size = list(tensor.size())
        self.orig_device = tensor.device
        self.partition = partition_uniform(num_items=tensor.numel(), num_parts=self.num_parts)
        lo, hi = self.partition[self.rank], self.partition[self.rank + 1]
        self.local_data = tensor.detach().contiguous().view(-1)[lo:hi].clone()
        self.even_split = tensor.numel() % self.num_parts == 0

    @classmethod
    def from_meta(cls, meta: torch.Tensor, local_part: torch.Tensor, group, device=None) -> "PartitionedTensor":
        assert meta.dtype == torch.long
        vals = meta.tolist()
        ndim = vals[0]
        obj = cls.__new__(cls)
        obj.group = group
        obj.orig_size = vals[1:1 + ndim]
        rest = vals[1 + ndim:]
        obj.num_parts, obj.rank, obj.partition = rest[0], rest[1], rest[2:]
        have = (dist.get_world_size(group=group), dist.get_rank(group=group)) if dist.is_initialized() else (1, 0)
        assert (obj.num_parts, obj.rank) == have, f"partition meta is for rank {obj.rank}/{obj.num_parts}, this is {have[1]}/{have[0]}"
        obj.orig_device = torch.device(device) if device is not None else local_part.device
        obj.local_data = local_part.detach()
        numel = 1
        for s in obj.orig_size:
            numel *= s
        obj.even_split = numel % obj.num_parts == 0
        return obj

    def full(self, device=None) -> torch.Tensor:
        device = device if device is not None else self.orig_device
        numel = self.partition[-1]
        flat = torch.zeros(numel, dtype=self.local_data.dtype, device=device)
        if self.num_parts == 1:
            flat.copy_(self.local_data)
        elif self.even_split:
            dist.all_gather_into_tensor(flat, self.local_data.to(device).contiguous(), group=self.group)
        else:
            ranks = torch.distributed.get_process_group_ranks(self.group) if self.group is not None else None
            for part in range(self.num_parts):
                buf = flat[self.partition[part]:self.partition[part + 1]]
                if part == self.rank:
                    buf.copy_(self.local_data)
                src = ranks[part] if ranks is not None else part
                dist.broadcast(buf, src, group=self.group)
        return flat.view(self.orig_size).clone().detach()

    def to_meta(self) -> torch.Tensor:
        """``[ndim, *shape, num_parts, rank, *partition boundaries]`` as a LongTensor on the tensor's device."""
        return torch.tensor([len(self.orig_size), *self.orig_size, self.num_parts, self.rank, *self.partition],
                            dtype=torch.long, device=self.orig_device)

    def data(self) -> torch.Tensor:
        return self.local_data

    def local_size(self):
        return self.local_data.size()

    def full_size(self):
        return self.orig_size


# ---- memory ---------------------------------------------------------------------------------------------------------------------
def get_ma_status() -> int:
    """Bytes currently allocated on the GPU (0 without one, or off rank 0 in a distributed job)."""
    if dist.is_initialized() and dist.get_rank() != 0:
        return 0
    return torch.cuda.memory_allocated() if torch.cuda.is_available() else 0


def empty_cache() -> None:
    if torch.cuda.is_available():
        torch.cuda.empty_cache()
        torch.cuda.reset_peak_memory_stats()


_mem = {"alloc": 0, "cached": 0}


def memory_status(msg: str, print_rank: int = -1, reset_max: bool = False):
    """Print allocator numbers and their change since the last call on ``print_rank`` (-1: every rank)."""
    rank = dist.get_rank() if dist.is_initialized() else 0
    if print_rank not in (-1, rank) or not torch.cuda.is_available():
        return None
    torch.cuda.synchronize()
    if reset_max:
        torch.cuda.reset_peak_memory_stats()
    gb = 1024 ** 3
    alloc, cached = torch.cuda.memory_allocated(), torch.cuda.memory_reserved()
    d_alloc, d_cached = alloc - _mem["alloc"], cached - _mem["cached"]
    _mem["alloc"], _mem["cached"] = alloc, cached
    logger.info(f"RANK={rank} MEMSTATS {msg} current alloc={alloc / gb:0.4f}GB (delta={d_alloc / gb:0.4f}GB max="
                f"{torch.cuda.max_memory_allocated() / gb:0.4f}GB) current cache={cached / gb:0.4f}GB (delta={d_cached / gb:0.4f}GB "
                f"max={torch.cuda.max_memory_reserved() / gb:0.4f}GB)")
    return alloc, cached


# ---- ZeRO all-gathers over per-group flat buffers --------------------------------------------------------------------------
def all_gather_into_tensor_dp_groups(groups_flat: Sequence[torch.Tensor], partitioned_param_groups, dp_process_group) -> None:
    """For every parameter group: all-gather this rank's slice into the group's flat buffer (one collective per group)."""
    for gid, (flat, parts) in enumerate(zip(groups_flat, partitioned_param_groups)):
        group = dp_process_group[gid] if isinstance(dp_process_group, (list, tuple)) else dp_process_group
        ws = dist.get_world_size(group=group)
        if ws == 1:
            continue
        rank = dist.get_rank(group=group)
        dist.all_gather_into_tensor(flat, parts[rank].contiguous().clone(), group=group)


def all_gather_dp_groups(groups_flat, partitioned_param_groups, dp_process_group, start_alignment_factor: int = 1,
                         allgather_bucket_size: int = 0) -> None:
    """As above; ``allgather_bucket_size`` > 0 splits each group's gather into shards of at most that many elements."""
    if not allgather_bucket_size:
        return all_gather_into_tensor_dp_groups(groups_flat, partitioned_param_groups, dp_process_group)
    for gid, parts in enumerate(partitioned_param_groups):
        group = dp_process_group[gid] if isinstance(dp_process_group, (list, tuple)) else dp_process_group
        ws = dist.get_world_size(group=group)
        if ws == 1:
            continue
        psize = parts[0].numel()
        shard = max(start_alignment_factor, (allgather_bucket_size // ws) // start_alignment_factor * start_alignment_factor)
        for lo in range(0, psize, shard):
            n = min(shard, psize - lo)
            outs = [p.narrow(0, lo, n) for p in parts]
            dist.all_gather(outs, outs[dist.get_rank(group=group)].clone(), group=group)


class TLinear(torch.nn.Linear):
    """A linear layer built from one whose weight is stored ``[in, out]`` (transposed), e.g. HuggingFace ``Conv1D``."""

    def __init__(self, orig_layer, name: str = ""):
        self.name = name
        w = orig_layer.weight
        super().__init__(w.shape[0], w.shape[1], bias=orig_layer.bias is not None)
        self.weight.data = w.data.t().contiguous()
        self.bias = orig_layer.bias
