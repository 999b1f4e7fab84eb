"""Full-precision views of a parameter whose master copy / gradient / optimizer state is partitioned by ZeRO:
``safe_get_full_fp32_param``, ``safe_get_full_grad``, ``safe_get_full_optimizer_state``, their setters, and the
``*_local_*`` variants that touch only this rank's partition (ZeRO-3).

Reference: ``deepspeed/utils/tensor_fragment.py`` (safe_get_full_fp32_param :118, safe_set_full_fp32_param :138,
safe_get_full_optimizer_state :158, safe_set_full_optimizer_state :180, safe_get_full_grad :203, safe_set_full_grad :230,
the local variants :260-330) — there through per-parameter fragment mappings attached to every parameter; here by slicing
the flat partitions the optimizers keep (ZeRO-1/2 and the BF16 optimizer: one flat fp32 partition per group; ZeRO-3: the
concatenation of each parameter's own partition).

The ``full`` getters / setters are COLLECTIVE under ZeRO (every rank of the data-parallel group must call them). The
optimizer is found through ``param._ds_optimizer`` (set by the engine) or the ``optimizer=`` argument.
"""
from __future__ import annotations

from typing import Optional, Tuple

import torch

from .. import comm as dist


# ---- locating a parameter inside the optimizers' flat buffers -----------------------------------------------------------
def _optimizer_of(param, optimizer):
    opt = optimizer if optimizer is not None else getattr(param, "_ds_optimizer", None)
    if opt is not None and hasattr(opt, "global_steps") and hasattr(opt, "module"):      # an engine was passed
        opt = opt.optimizer
    return opt


def _kind(opt) -> str:
    from ..runtime.zero.stage3 import DeepSpeedZeroOptimizer_Stage3
    from ..runtime.zero.stage_1_and_2 import DeepSpeedZeroOptimizer
    if isinstance(opt, DeepSpeedZeroOptimizer_Stage3):
        return "z3"
    if isinstance(opt, DeepSpeedZeroOptimizer):
        return "z12"
    return "plain"


def _find12(opt, param) -> Tuple[Optional[int], Optional[int]]:
    for gi, params in enumerate(opt.bit16_groups):
        off = 0
        for p in params:
            if p is param:
                return gi, off
            off += p.numel()
    return None, None


def _find3(opt, param) -> Tuple[Optional[int], Optional[int]]:
    for gi, params in enumerate(opt.fp16_groups):
        off = 0
        for p in params:
            if p is param:
                return gi, off
            off += p.ds_partition_numel
    return None, None


def _gather_flat12(opt, gi: int, local: torch.Tensor) -> torch.Tensor:
    if opt.partition_count == 1:
        return local
    dev = opt.bit16_groups_flat[gi].device
    full = torch.empty(local.numel() * opt.partition_count, dtype=local.dtype, device=dev)
    dist.all_gather_into_tensor(full, local.to(dev).contiguous(), group=opt.dp_process_group)
    return full


def _gather3(param, local: torch.Tensor) -> torch.Tensor:
    """Full tensor of a ZeRO-3 parameter from every rank's partition-sized ``local`` slice."""
    group = param.ds_process_group
    ws = dist.get_world_size(group=group) if dist.is_initialized() else 1
    local = local.detach().contiguous()
    if ws > 1:
        full = torch.empty(local.numel() * ws, dtype=local.dtype, device=local.device)
        dist.all_gather_into_tensor(full, local, group=group)
    else:
        full = local
    return full[:param.ds_numel].view(param.ds_shape).clone()


def _my_slice3(param, value: torch.Tensor) -> torch.Tensor:
    """This rank's partition (zero-padded to the partition size) of a FULL tensor ``value``."""
    group = param.ds_process_group
    rank = dist.get_rank(group=group) if dist.is_initialized() else 0
    psize = param.ds_partition_numel
    flat = value.reshape(-1)
    out = flat.new_zeros(psize)
    lo, hi = rank * psize, min(param.ds_numel, (rank + 1) * psize)
    if lo < hi:
        out[:hi - lo].copy_(flat[lo:hi])
    return out


class _resident:
    """ZeRO-3 masters may live on NVMe between steps: bring them in for the duration."""

    def __init__(self, opt):
        self.ctx = opt._Resident(opt) if hasattr(opt, "_Resident") else None

    def __enter__(self):
        if self.ctx is not None:
            self.ctx.__enter__()

    def __exit__(self, *exc):
        if self.ctx is not None:
            self.ctx.__exit__(*exc)
        return False


# ---- fp32 master parameter --------------------------------------------------------------------------------------------
def safe_get_full_fp32_param(param, optimizer=None) -> Optional[torch.Tensor]:
    """Collective: every rank gets the full fp32 master value of ``param``."""
    opt = _optimizer_of(param, optimizer)
    kind = _kind(opt) if opt is not None else "plain"
    if kind == "z12":
        gi, off = _find12(opt, param)
        if gi is not None:
            full = _gather_flat12(opt, gi, opt.single_partition_of_fp32_groups[gi].detach())
            return full[off:off + param.numel()].view(param.shape).clone()
    elif kind == "z3":
        gi, off = _find3(opt, param)
        if gi is not None:
            with _resident(opt):
                local = opt.fp32_partitioned_groups_flat[gi].detach().narrow(0, off, param.ds_partition_numel).clone()
            return _gather3(param, local)
        from ..runtime.zero.partition_parameters import GatheredParameters     # frozen parameter: no master copy
        with GatheredParameters([param]):
            return param.data.detach().float().clone()
    return param.detach().float()


def safe_set_full_fp32_param(param, value: torch.Tensor, optimizer=None):
    """Collective: write ``value`` (full shape) into the fp32 master and the 16-bit copy of ``param``."""
    opt = _optimizer_of(param, optimizer)
    kind = _kind(opt) if opt is not None else "plain"
    if kind == "z12":
        gi, off = _find12(opt, param)
        if gi is not None:
            psize = opt.bit16_groups_flat[gi].numel() // opt.partition_count
            lo = opt.rank * psize
            s, e = max(off, lo), min(off + param.numel(), lo + psize)
            if s < e:
                opt.single_partition_of_fp32_groups[gi].data[s - lo:e - lo].copy_(value.reshape(-1)[s - off:e - off])
        param.data.copy_(value.to(param.dtype))
        return
    if kind == "z3":
        from ..runtime.zero.partition_parameters import ZeroParamStatus, all_gather_params, write_partition
        mine = _my_slice3(param, value.detach().float())
        gi, off = _find3(opt, param)
        if gi is not None:
            with _resident(opt):
                opt.fp32_partitioned_groups_flat[gi].data.narrow(0, off, param.ds_partition_numel).copy_(mine)
        write_partition(param, mine)
        if param.ds_status == ZeroParamStatus.AVAILABLE:          # a gathered / persistent copy is refreshed
            param.ds_status = ZeroParamStatus.NOT_AVAILABLE
            all_gather_params([param])
        return
    param.data.copy_(value.to(param.dtype))


# ---- optimizer state ------------------------------------------------------------------------------------------------------
def _state_of(opt, master):
    return opt.optimizer.state.get(master, {})


def safe_get_full_optimizer_state(param, optim_state_key: str, optimizer=None) -> Optional[torch.Tensor]:
    opt = _optimizer_of(param, optimizer)
    kind = _kind(opt) if opt is not None else "plain"
    if kind == "z12":
        gi, off = _find12(opt, param)
        if gi is None:
            return None
        st = _state_of(opt, opt.single_partition_of_fp32_groups[gi])
        if optim_state_key not in st or not torch.is_tensor(st[optim_state_key]) or st[optim_state_key].dim() == 0:
            return None
        full = _gather_flat12(opt, gi, st[optim_state_key].detach())
        return full[off:off + param.numel()].view(param.shape).clone()
    if kind == "z3":
        gi, off = _find3(opt, param)
        if gi is None:
            return None
        with _resident(opt):
            st = _state_of(opt, opt.fp32_partitioned_groups_flat[gi])
            if optim_state_key not in st or not torch.is_tensor(st[optim_state_key]) or st[optim_state_key].dim() == 0:
                return None
            local = st[optim_state_key].detach().narrow(0, off, param.ds_partition_numel).clone()
        return _gather3(param, local)
    if opt is not None and hasattr(opt, "state"):
        st = opt.state.get(param, {})
        return st[optim_state_key].detach().float().clone() if optim_state_key in st and torch.is_tensor(st[optim_state_key]) else None
    return None


def safe_set_full_optimizer_state(param, value: torch.Tensor, optim_state_key: str, optimizer=None):
    opt = _optimizer_of(param, optimizer)
    kind = _kind(opt) if opt is not None else "plain"
    if kind == "z12":
        gi, off = _find12(opt, param)
        if gi is None:
            return
        st = _state_of(opt, opt.single_partition_of_fp32_groups[gi])
        if optim_state_key not in st:
            return
        psize = opt.bit16_groups_flat[gi].numel() // opt.partition_count
        lo = opt.rank * psize
        s, e = max(off, lo), min(off + param.numel(), lo + psize)
        if s < e:
            st[optim_state_key].data[s - lo:e - lo].copy_(value.reshape(-1)[s - off:e - off])
    elif kind == "z3":
        gi, off = _find3(opt, param)
        if gi is None:
            return
        with _resident(opt):
            st = _state_of(opt, opt.fp32_partitioned_groups_flat[gi])
            if optim_state_key in st:
                st[optim_state_key].data.narrow(0, off, param.ds_partition_numel).copy_(_my_slice3(param, value.detach().float()))
    elif opt is not None and hasattr(opt, "state") and optim_state_key in opt.state.get(param, {}):
        opt.state[param][optim_state_key].data.copy_(value)


# ---- gradients (between backward and step) ----------------------------------------------------------------------------------
def safe_get_full_grad(param, optimizer=None) -> Optional[torch.Tensor]:
    """The averaged, unscaled gradient of ``param`` (gathered from the partitions under ZeRO-2 / ZeRO-3)."""
    opt = _optimizer_of(param, optimizer)
    kind = _kind(opt) if opt is not None else "plain"
    if kind == "z12":
        gi, off = _find12(opt, param)
        if gi is None or opt.grad_partitions[gi] is None:
            # stage 1 (and the BF16 optimizer) reduce at step time: average this micro-batch's gradients here
            have = torch.tensor([0 if param.grad is None else 1])
            if dist.is_initialized() and opt.partition_count > 1:
                dist.all_reduce(have, group=opt.dp_process_group)
            if int(have) == 0:
                return None
            g = (param.grad.detach().float() if param.grad is not None else torch.zeros(param.shape)).clone()
            if dist.is_initialized() and opt.partition_count > 1:
                dist.all_reduce(g, group=opt.dp_process_group)
                g /= opt.partition_count
            return g / opt.loss_scale
        full = _gather_flat12(opt, gi, opt.grad_partitions[gi].detach())
        return full[off:off + param.numel()].view(param.shape).float().clone() / opt.loss_scale
    if kind == "z3":
        g = opt.grad_partitions.get(param.ds_id)
        have = torch.tensor([0 if g is None else 1])
        if dist.is_initialized():
            dist.all_reduce(have, group=param.ds_process_group)
        if int(have) == 0:
            return None
        if g is None:
            g = torch.zeros(param.ds_partition_numel, dtype=torch.float32)
        return _gather3(param, g.float()) / opt.loss_scale
    return None if param.grad is None else param.grad.detach().float()


def safe_set_full_grad(param, value: torch.Tensor, optimizer=None):
    """Replace the (averaged, unscaled) gradient of ``param`` before ``step()``."""
    opt = _optimizer_of(param, optimizer)
    kind = _kind(opt) if opt is not None else "plain"
    if kind == "z12":
        gi, off = _find12(opt, param)
        if gi is not None and opt.grad_partitions[gi] is not None:
            part = opt.grad_partitions[gi]
            psize = part.numel()
            lo = opt.rank * psize
            s, e = max(off, lo), min(off + param.numel(), lo + psize)
            if s < e:
                part.data[s - lo:e - lo].copy_((value.reshape(-1)[s - off:e - off] * opt.loss_scale).to(part.dtype))
            return
        param.grad = (value * getattr(opt, "loss_scale", 1.0)).to(param.dtype).reshape(param.shape).clone()
        return
    if kind == "z3":
        mine = _my_slice3(param, value.detach().float()) * opt.loss_scale
        opt.grad_partitions[param.ds_id] = mine.to(opt.gradient_accumulation_dtype)
        return
    param.grad = value.to(param.dtype).reshape(param.shape).clone()


# ---- this rank's partition only (ZeRO-3; no communication) --------------------------------------------------------------
def safe_get_local_fp32_param(param, optimizer=None) -> Optional[torch.Tensor]:
    opt = _optimizer_of(param, optimizer)
    if opt is not None and _kind(opt) == "z3":
        gi, off = _find3(opt, param)
        if gi is None:
            return None
        with _resident(opt):
            return opt.fp32_partitioned_groups_flat[gi].detach().narrow(0, off, param.ds_partition_numel).clone()
    return safe_get_full_fp32_param(param, optimizer)


def safe_set_local_fp32_param(param, value: torch.Tensor, optimizer=None):
    opt = _optimizer_of(param, optimizer)
    if opt is not None and _kind(opt) == "z3":
        from ..runtime.zero.partition_parameters import ZeroParamStatus, all_gather_params, write_partition
        gi, off = _find3(opt, param)
        if gi is not None:
            with _resident(opt):
                opt.fp32_partitioned_groups_flat[gi].data.narrow(0, off, param.ds_partition_numel).copy_(value.reshape(-1))
        write_partition(param, value.reshape(-1))
        if param.ds_status == ZeroParamStatus.AVAILABLE:
            param.ds_status = ZeroParamStatus.NOT_AVAILABLE
            all_gather_params([param])                     # collective: every rank set its slice, all refresh the copy
        return
    safe_set_full_fp32_param(param, value, optimizer)


def safe_get_local_grad(param, optimizer=None) -> Optional[torch.Tensor]:
    opt = _optimizer_of(param, optimizer)
    if opt is not None and _kind(opt) == "z3":
        g = opt.grad_partitions.get(param.ds_id)
        return None if g is None else g.detach().float().clone() / opt.loss_scale
    return safe_get_full_grad(param, optimizer)


def safe_set_local_grad(param, value: torch.Tensor, optimizer=None):
    opt = _optimizer_of(param, optimizer)
    if opt is not None and _kind(opt) == "z3":
        opt.grad_partitions[param.ds_id] = (value.reshape(-1).float() * opt.loss_scale).to(opt.gradient_accumulation_dtype)
        return
    safe_set_full_grad(param, value, optimizer)


def safe_get_local_optimizer_state(param, optim_state_key: str, optimizer=None) -> Optional[torch.Tensor]:
    opt = _optimizer_of(param, optimizer)
    if opt is not None and _kind(opt) == "z3":
        gi, off = _find3(opt, param)
        if gi is None:
            return None
        with _resident(opt):
            st = _state_of(opt, opt.fp32_partitioned_groups_flat[gi])
            if optim_state_key not in st or not torch.is_tensor(st[optim_state_key]) or st[optim_state_key].dim() == 0:
                return None
            return st[optim_state_key].detach().narrow(0, off, param.ds_partition_numel).clone()
    return safe_get_full_optimizer_state(param, optim_state_key, optimizer)


def safe_set_local_optimizer_state(param, value: torch.Tensor, optim_state_key: str, optimizer=None):
    opt = _optimizer_of(param, optimizer)
    if opt is not None and _kind(opt) == "z3":
        gi, off = _find3(opt, param)
        if gi is None:
            return
        with _resident(opt):
            st = _state_of(opt, opt.fp32_partitioned_groups_flat[gi])
            if optim_state_key in st:
                st[optim_state_key].data.narrow(0, off, param.ds_partition_numel).copy_(value.reshape(-1))
        return
    safe_set_full_optimizer_state(param, value, optim_state_key, optimizer)


# ---- the fragment-linkage form of the accessors (ref: tensor_fragment.py:13-130, 312): functions that take the low-precision
#      parameter as ``self`` - ``link_hp_params`` binds them as methods for optimizers that keep ``_hp_mapping`` fragments --------
from .mixed_precision_linkage import fragment_address, get_hp_fragment_mapping, tensor_fragment  # noqa: E402,F401


def _frag_full(self, pick) -> torch.Tensor:
    """Every rank writes its fragment into a zero buffer of the parameter's size; the sum over the group is the tensor."""
    buf = torch.zeros(self.numel(), dtype=torch.float32, device=self.device)
    m = getattr(self, "_hp_mapping", None)
    if m is not None:
        a = m.lp_fragment_address
        buf.narrow(0, a.start, a.numel).copy_(pick(m).detach().float().reshape(-1))
    group = getattr(self, "_dp_group", None)
    if dist.is_initialized() and dist.get_world_size(group=group) > 1:
        dist.all_reduce(buf, group=group)
    return buf.reshape(self.shape)


def get_full_hp_param(self, optim_state_key=None):
    """Full fp32 value (or optimizer state ``optim_state_key``) of a parameter whose master copy is spread over the group
    as ``_hp_mapping`` fragments; parameters without the linkage go through ``safe_get_full_*``."""
    if not hasattr(self, "_hp_mapping"):
        return safe_get_full_fp32_param(self) if optim_state_key is None else safe_get_full_optimizer_state(self, optim_state_key)
    return _frag_full(self, lambda m: m.get_hp_fragment(optim_state_key))


def set_full_hp_param(self, value, optim_state_key=None) -> None:
    if not hasattr(self, "_hp_mapping"):
        if optim_state_key is None:
            return safe_set_full_fp32_param(self, value)
        return safe_set_full_optimizer_state(self, value, optim_state_key)
    m = self._hp_mapping
    if m is not None:
        a = m.lp_fragment_address
        m.get_hp_fragment(optim_state_key).data.copy_(value.reshape(-1).narrow(0, a.start, a.numel).to(m.hp_fragment.dtype))


def _lp_grad_fragment(self):
    m = self._hp_mapping
    grads = getattr(m, "gradient_dict", None)
    if grads and m.param_group_index in grads and grads[m.param_group_index] is not None:
        return grads[m.param_group_index][self._index_in_param_group]
    if self.grad is None:
        raise ValueError("Gradients are only available immediately after backward and before engine step")
    a = m.lp_fragment_address
    return self.grad.reshape(-1).narrow(0, a.start, a.numel)


def get_full_hp_grad(self):
    if not hasattr(self, "_hp_mapping"):
        return safe_get_full_grad(self)
    return _frag_full(self, lambda m: _lp_grad_fragment(self)) if self._hp_mapping is not None else _frag_full(self, None)


def set_full_hp_grad(self, value) -> None:
    if not hasattr(self, "_hp_mapping"):
        return safe_set_full_grad(self, value)
    if self._hp_mapping is not None:
        frag = _lp_grad_fragment(self)
        a = self._hp_mapping.lp_fragment_address
        frag.data.copy_(value.reshape(-1).narrow(0, a.start, a.numel).reshape(frag.shape).to(frag.dtype))


def map_to_flat_opt_states(flat_hp_tensor, lp_tensors, optim_state, opt_keys) -> None:
    """Rebuild the flat optimizer-state tensors ``optim_state[flat_hp_tensor][key]`` from the fragments the low-precision
    parameters hold, and re-point the fragments into the new flat buffers."""
    for key in opt_keys:
        buf = torch.zeros_like(flat_hp_tensor)
        for lp in lp_tensors:
            m = getattr(lp, "_hp_mapping", None)
            if m is None:
                continue
            a = m.hp_fragment_address
            piece = buf.narrow(0, a.start, a.numel)
            piece.copy_(m.get_hp_fragment(optim_state_key=key).data)
            m.optim_fragment[key] = piece
        optim_state[flat_hp_tensor][key] = buf
