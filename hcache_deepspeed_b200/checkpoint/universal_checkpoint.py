"""Resume from a universal checkpoint under ANY data-parallel degree: every rank cuts its own fp32 / moment slices out
of the per-parameter files. Reference: ``deepspeed/checkpoint/universal_checkpoint.py:13-150`` (load_hp_checkpoint_state)."""
from __future__ import annotations

import os
from typing import Dict, Optional

import torch


def load_hp_checkpoint_state(folder: str, name: str, key: str = "fp32") -> Optional[torch.Tensor]:
    path = os.path.join(folder, "zero", name, f"{key}.pt")
    if not os.path.exists(path):
        return None
    obj = torch.load(path, map_location="cpu", weights_only=False)
    return obj["param"] if isinstance(obj, dict) else obj


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


def load_universal_into_engine(engine, folder: str, load_optimizer_states: bool = True) -> int:
    """Fill the module's parameters, and — for ZeRO-1/2 / BF16 optimizers — this rank's fp32 master slices and Adam
    moments, from a universal checkpoint. Returns the number of parameters restored."""
    from ..runtime.zero.stage_1_and_2 import DeepSpeedZeroOptimizer
    restored = 0
    opt = engine.optimizer
    if isinstance(opt, DeepSpeedZeroOptimizer):
        return load_universal_into_optimizer(opt, folder, load_optimizer_states)
    for name, p in engine.module.named_parameters():
        full = load_hp_checkpoint_state(folder, name, "fp32")
        if full is not None:
            p.data.copy_(full.to(p.dtype).reshape(p.shape))
            restored += 1
    return restored
