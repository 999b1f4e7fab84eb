"""Resume from a universal checkpoint under ANY data-parallel degree: every rank cuts its own fp32 / moment slices out
of the per-parameter files. Reference: ``deepspeed/checkpoint/universal_checkpoint.py:13-150`` (load_hp_checkpoint_state)."""
from __future__ import annotations

import dataclasses
import os
from typing import Dict, List, Optional, Tuple

import torch


@dataclasses.dataclass
class SubparamShape:
    """A parameter that is several sub-parameters stacked along ``partition_dim`` (fused QKV with unequal parts, ...):
    ``shape`` has a tuple of sub-sizes at that position; ``patterns`` are the parameter-name regexes it applies to."""
    patterns: List[str]
    shape: Tuple
    partition_dim: int


def _tp_slice(full: torch.Tensor, meta: dict, tp_rank: int, tp_size: int, padded_vocab: Optional[int] = None) -> torch.Tensor:
    """This tensor-parallel rank's part of a universal (full) tensor, by the metadata the converter stored next to it."""
    if meta.get("vocab_tensor") and padded_vocab and full.shape[0] < padded_vocab:
        pad = full.new_zeros((padded_vocab - full.shape[0], ) + tuple(full.shape[1:]))
        full = torch.cat([full, pad], dim=0)
    if tp_size <= 1:
        return full
    sub = meta.get("sub_param_shape")
    if sub is not None:
        dim = sub.partition_dim
        sizes = sub.shape[dim] if isinstance(sub.shape[dim], (tuple, list)) else (sub.shape[dim], )
        whole = [(sum(d) if isinstance(d, (tuple, list)) else d) for d in sub.shape]
        view, parts, off = full.reshape(whole), [], 0
        for size in sizes:
            parts.append(view.narrow(dim, off, size).chunk(tp_size, dim=dim)[tp_rank])
            off += size
        return torch.cat(parts, dim=dim)
    if "cat_dim" not in meta:
        return full                                     # replicated / averaged parameters: every rank takes it whole
    dim = meta["cat_dim"]
    n_sub = meta.get("param_n_sub_params", 1)
    if n_sub > 1:
        return torch.cat([c.chunk(tp_size, dim=dim)[tp_rank] for c in full.chunk(n_sub, dim=dim)], dim=dim)
    return full.chunk(tp_size, dim=dim)[tp_rank] if full.dim() > dim else full


def load_hp_checkpoint_state(folder: str, name: str, key: str = "fp32", tp_rank: int = 0, tp_world_size: int = 1,
                             padded_vocab_size: Optional[int] = None) -> Optional[torch.Tensor]:
    """One parameter's fp32 weight or moment from a universal checkpoint; under tensor parallelism this rank's slice (the
    file records how the tensor was joined: ``cat_dim``, stacked sub-parameters, vocabulary padding)."""
    path = os.path.join(folder, "zero", name, f"{key}.pt")
    if not os.path.exists(path):
        return None
    obj = torch.load(path, map_location="cpu", weights_only=False)
    if not isinstance(obj, dict):
        return obj
    return _tp_slice(obj["param"], obj, tp_rank, tp_world_size, padded_vocab_size)


def enable_universal_checkpoint(param_list) -> None:
    """Give every low-precision parameter the ``load_hp_checkpoint_state(folder, tp_rank, tp_world_size)`` method the
    reference's optimizers call on it; here it returns the tensors instead of copying into a linked fragment."""
    import types

    def load(self, folder, tp_rank=0, tp_world_size=1):
        name = os.path.basename(folder.rstrip("/"))
        root = os.path.dirname(os.path.dirname(folder.rstrip("/")))
        return {k: load_hp_checkpoint_state(root, name, k, tp_rank, tp_world_size) for k in ("fp32", "exp_avg", "exp_avg_sq")}

    for p in param_list:
        p.load_hp_checkpoint_state = types.MethodType(load, p)


def load_universal_into_optimizer(opt, folder: str, load_optimizer_states: bool = True,
                                  lp_groups_name: str = "bit16_groups") -> int:
    """This rank's fp32 master slices and Adam moments of a ZeRO-1/2 / BF16 optimizer, cut out of the per-parameter
    files (``base_optimizer.py:20-63`` of the reference does this per lp parameter)."""
    restored = 0
    if True:
        for gi, params in enumerate(getattr(opt, lp_groups_name)):
            if not params:
                continue
            flat = opt.bit16_groups_flat[gi]
            psize = flat.numel() // opt.partition_count
            lo = opt.rank * psize
            master = opt.single_partition_of_fp32_groups[gi]
            state = opt.optimizer.state.get(master, {})
            cursor = 0
            for p in params:
                name = opt.param_names.get(p)
                n = p.numel()
                s, e = max(cursor, lo), min(cursor + n, lo + psize)
                if name is not None and s < e:
                    full = load_hp_checkpoint_state(folder, name, "fp32")
                    if full is not None:
                        master.data[s - lo:e - lo].copy_(full.reshape(-1)[s - cursor:e - cursor])
                        restored += 1
                    if load_optimizer_states:
                        for key in ("exp_avg", "exp_avg_sq"):
                            t = load_hp_checkpoint_state(folder, name, key)
                            if t is not None and key in state:
                                state[key].data[s - lo:e - lo].copy_(t.reshape(-1)[s - cursor:e - cursor])
                        step = None
                        sp = os.path.join(folder, "zero", name, "step.pt")
                        if os.path.exists(sp):
                            step = torch.load(sp, map_location="cpu", weights_only=False)
                        if step is not None and "step" in state:
                            state["step"] = step if not torch.is_tensor(state["step"]) else torch.as_tensor(float(step))
                cursor += n
            flat.narrow(0, lo, psize).copy_(master.detach().to(flat.device))
            if opt.partition_count > 1:
                from .. import comm as dist
                dist.all_gather_into_tensor(flat, flat.narrow(0, lo, psize).clone(), group=opt.dp_process_group)
    return restored


def load_universal_into_stage3(opt, folder: str, load_optimizer_states: bool = True) -> int:
    """ZeRO-3: every rank takes ITS slice of each parameter (padded equal cut) for the fp32 master, the moments and the
    low-precision partition."""
    from ..runtime.zero.partition_parameters import ZeroParamStatus, all_gather_params, write_partition
    restored, ws, rank = 0, opt.partition_count, opt.rank
    if load_optimizer_states and not any("exp_avg" in st for st in opt.optimizer.state.values()):
        opt.initialize_optimizer_states()          # moments exist only after a first step: create them to load into
    with opt._Resident(opt):
        for gi, params in enumerate(opt.fp16_groups):
            master = opt.fp32_partitioned_groups_flat[gi]
            state = opt.optimizer.state.get(master, {})
            off = 0
            for p in params:
                name, n = opt.param_names.get(p), p.ds_partition_numel

                def my_slice(full):
                    flat = full.reshape(-1).float()
                    lo, hi = rank * n, min(flat.numel(), (rank + 1) * n)
                    out = torch.zeros(n)
                    if lo < hi:
                        out[:hi - lo] = flat[lo:hi]
                    return out

                full = load_hp_checkpoint_state(folder, name, "fp32") if name is not None else None
                if full is not None:
                    master.data.narrow(0, off, n).copy_(my_slice(full))
                    write_partition(p, master.detach().narrow(0, off, n))
                    if p.ds_status == ZeroParamStatus.AVAILABLE:
                        p.ds_status = ZeroParamStatus.NOT_AVAILABLE
                        all_gather_params([p])
                    restored += 1
                    if load_optimizer_states:
                        for key in ("exp_avg", "exp_avg_sq"):
                            t = load_hp_checkpoint_state(folder, name, key)
                            if t is not None and key in state:
                                state[key].data.narrow(0, off, n).copy_(my_slice(t))
                        sp = os.path.join(folder, "zero", name, "step.pt")
                        if os.path.exists(sp) and "step" in state:
                            step = torch.load(sp, map_location="cpu", weights_only=False)
                            state["step"] = step if not torch.is_tensor(state["step"]) else torch.as_tensor(float(step))
                off += n
    return restored


def load_universal_into_engine(engine, folder: str, load_optimizer_states: bool = True) -> int:
    """Fill the module's parameters, and — for ZeRO-1/2 / BF16 optimizers — this rank's fp32 master slices and Adam
    moments, from a universal checkpoint. Returns the number of parameters restored."""
    from ..runtime.zero.stage3 import DeepSpeedZeroOptimizer_Stage3
    from ..runtime.zero.stage_1_and_2 import DeepSpeedZeroOptimizer
    restored = 0
    opt = engine.optimizer
    if isinstance(opt, DeepSpeedZeroOptimizer):
        return load_universal_into_optimizer(opt, folder, load_optimizer_states)
    if isinstance(opt, DeepSpeedZeroOptimizer_Stage3):
        return load_universal_into_stage3(opt, folder, load_optimizer_states)
    for name, p in engine.module.named_parameters():
        full = load_hp_checkpoint_state(folder, name, "fp32")
        if full is not None:
            p.data.copy_(full.to(p.dtype).reshape(p.shape))
            restored += 1
    return restored
