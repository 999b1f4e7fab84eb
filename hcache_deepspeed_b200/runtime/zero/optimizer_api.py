"""User-facing methods shared by the ZeRO optimizer wrappers (stages 1/2 and 3).

The reference's two ZeRO optimizers (``runtime/zero/stage_1_and_2.py``, ``runtime/zero/stage3.py``) carry, next to their
internals, a set of methods that trainers and integrations call from outside: learning-rate access, an externally
managed loss scale, overflow checks, gradient-norm computation over partitioned gradients, un-scale-and-clip, partition
views, the hp/lp accessors behind ``deepspeed.utils.safe_get_*``, and state offload. Both wrappers here get them from
this one mixin; the per-stage part is reduced to three hooks:

* ``_api_master_tensors()`` - this rank's flat fp32 master tensors, one per parameter group
* ``_api_local_grads()``    - this rank's (partitioned, still loss-scaled) gradient tensors
* ``_api_params()``         - the low-precision parameters the optimizer owns
"""
from __future__ import annotations

import math
from typing import Iterable, List, Optional

import torch

from ... import comm as dist


def _tf():
    """The accessor module (``utils.tensor_fragment`` the attribute is a class of the same name, hence by path)."""
    import importlib
    return importlib.import_module(__package__.rsplit(".", 2)[0] + ".utils.tensor_fragment")


class ZeroOptimizerAPI:

    # ---- learning rate --------------------------------------------------------------------------------------------------
    def set_lr(self, lr: float) -> None:
        """Set the learning rate of every parameter group."""
        for g in self.optimizer.param_groups:
            g["lr"] = lr

    def get_lr(self) -> float:
        """Learning rate of the first parameter group."""
        return self.optimizer.param_groups[0]["lr"]

    # ---- loss scale --------------------------------------------------------------------------------------------------------
    def override_loss_scale(self, loss_scale: float) -> None:
        """From now on the loss is scaled by this externally managed value; the built-in (dynamic) scaler stops updating."""
        self.custom_loss_scaler = True
        self.external_loss_scale = loss_scale

    # ---- overflow / norms ------------------------------------------------------------------------------------------------
    @staticmethod
    def _tensor_has_inf_or_nan(t: torch.Tensor) -> bool:
        s = float(t.float().sum())
        return s in (float("inf"), -float("inf")) or s != s

    def has_overflow_serial(self, tensors: Iterable[torch.Tensor]) -> bool:
        """inf / nan in any of ``tensors`` (their ``.grad`` when they are parameters), on this rank only."""
        for t in tensors:
            g = t.grad if isinstance(t, torch.nn.Parameter) else t
            if g is not None and g.numel() and self._tensor_has_inf_or_nan(g):
                return True
        return False

    def has_overflow_partitioned_grads_serial(self) -> bool:
        return self.has_overflow_serial(g for g in self._api_local_grads() if g is not None)

    def check_overflow(self, partition_gradients: bool = True) -> bool:
        """Collective: sets and returns ``self.overflow`` for the gradients currently held."""
        local = 1.0 if self.has_overflow_partitioned_grads_serial() else 0.0
        self.overflow = self._all_reduce_scalar(local, dist.ReduceOp.MAX) > 0
        return self.overflow

    def get_grad_norm_direct(self, gradients: Iterable[torch.Tensor], params=None, norm_type: float = 2) -> float:
        """Norm of partitioned gradients: every rank passes ITS slices, the result is the norm of the whole gradient
        (summed over the partition group and the model-parallel group). inf / nan come back as ``-1``."""
        gradients = [g for g in gradients if g is not None and g.numel()]
        if norm_type == float("inf"):
            local = max((float(g.abs().max()) for g in gradients), default=0.0)
            total = self._all_reduce_scalar(local, dist.ReduceOp.MAX)
        else:
            local = sum(float(g.float().norm(norm_type)) ** norm_type for g in gradients)
            total = self._all_reduce_scalar(local, dist.ReduceOp.SUM) ** (1.0 / norm_type)
        return -1.0 if (math.isinf(total) or math.isnan(total)) else total

    def unscale_and_clip_grads(self, grad_groups_flat: List[torch.Tensor], total_norm: float) -> float:
        """Divide flat gradients in place by the loss scale, and further when ``total_norm`` (of the scaled gradients)
        exceeds ``clip_grad`` after un-scaling. Returns the divisor used."""
        scale = float(self.loss_scale)
        combined = scale
        if self.clip_grad > 0.0:
            clip = (total_norm / scale + 1e-6) / self.clip_grad
            if clip > 1:
                combined = clip * scale
        for g in grad_groups_flat:
            if isinstance(g, (list, tuple)):
                for t in g:
                    t.data.mul_(1.0 / combined)
            else:
                g.data.mul_(1.0 / combined)
        return combined

    # ---- partition geometry ------------------------------------------------------------------------------------------------
    def get_data_parallel_partitions(self, tensor: torch.Tensor, group_id: Optional[int] = None) -> List[torch.Tensor]:
        """Views of a flat tensor, one per rank of the partition group (the first ``numel % ranks`` are one element longer)."""
        n, ws = tensor.numel(), self.partition_count
        base, rem = divmod(n, ws)
        out, start = [], 0
        for r in range(ws):
            size = base + (1 if r < rem else 0)
            out.append(tensor.narrow(0, start, size))
            start += size
        return out

    def get_partition_info(self, tensor_list: List[torch.Tensor], partition_size: int, partition_id: int):
        """Which tensors of ``tensor_list`` (laid out back to back) fall inside partition ``partition_id``: returns
        (tensors inside, tensors outside, offset of the first inside tensor's first owned element)."""
        lo, hi = partition_size * partition_id, partition_size * (partition_id + 1)
        inside, outside, first_offset, pos = [], [], 0, 0
        for t in tensor_list:
            n = t.numel()
            if pos < hi and pos + n > lo:
                if not inside:
                    first_offset = max(lo - pos, 0)
                inside.append(t)
            else:
                outside.append(t)
            pos += n
        return inside, outside, first_offset

    def get_flat_partition(self, tensor_list, first_offset: int, partition_size: int, dtype=None, device=None,
                           return_tensor_list: bool = False):
        """This rank's ``partition_size`` elements of the gradients of ``tensor_list`` (as returned by
        ``get_partition_info``), missing gradients read as zero, zero-padded at the end."""
        pieces, have = [], 0
        for i, t in enumerate(tensor_list):
            g = t.grad if getattr(t, "grad", None) is not None else torch.zeros_like(t)
            flat = g.reshape(-1)
            if i == 0 and first_offset:
                flat = flat[first_offset:]
            flat = flat[:partition_size - have]
            pieces.append(flat.to(dtype=dtype or flat.dtype, device=device or flat.device))
            have += flat.numel()
            if have >= partition_size:
                break
        if have < partition_size:
            ref = pieces[0] if pieces else torch.zeros(0, dtype=dtype or torch.float32, device=device or "cpu")
            pieces.append(torch.zeros(partition_size - have, dtype=ref.dtype, device=ref.device))
        return pieces if return_tensor_list else torch.cat(pieces)

    def free_grad_in_param_list(self, param_list) -> None:
        for p in param_list:
            p.grad = None

    def set_none_gradients_to_zero(self, i: int, partition_id: int = 0) -> None:
        groups = getattr(self, "bit16_groups", None) or getattr(self, "fp16_groups", [])
        for p in groups[i]:
            if p.grad is None and p.numel():
                p.grad = torch.zeros_like(p)

    # ---- optimizer state bring-up -----------------------------------------------------------------------------------------
    def initialize_optimizer_states(self) -> None:
        """Create the inner optimizer's state tensors now (one step on zero gradients) so that memory is reserved up
        front and a checkpoint can be loaded into them. Weight decay, if any, is suspended for that step."""
        masters = [m for m in self._api_master_tensors() if m.numel()]
        saved = [m.detach().clone() for m in masters]
        decays = [g.get("weight_decay", None) for g in self.optimizer.param_groups]
        for g in self.optimizer.param_groups:
            if "weight_decay" in g:
                g["weight_decay"] = 0.0
        for m in masters:
            m.grad = torch.zeros_like(m)
        self.optimizer.step()
        for m, s in zip(masters, saved):
            m.grad = None
            m.data.copy_(s)
        for g, d in zip(self.optimizer.param_groups, decays):
            if d is not None:
                g["weight_decay"] = d
        for st in self.optimizer.state.values():           # the warm-up step must not count as a real one
            if "step" in st:
                st["step"] = st["step"] * 0 if torch.is_tensor(st["step"]) else 0

    def reset_cpu_buffers(self) -> None:
        """Per-step host-side accumulators of the reference's CPU-offload path; gradients here are reduced straight into
        the owned slice, so there is nothing to clear beyond the norm bookkeeping."""
        self._global_grad_norm = 0.0

    # ---- hp / lp accessors (what deepspeed.utils.safe_get_* dispatch to; tensor_fragment.py has the logic) ----------------
    def get_full_hp_param(self, param, optim_state_key: Optional[str] = None):
        tf = _tf()
        return tf.safe_get_full_fp32_param(param, optimizer=self) if optim_state_key is None else tf.safe_get_full_optimizer_state(param, optim_state_key, optimizer=self)

    def set_full_hp_param(self, value, param, optim_state_key: Optional[str] = None) -> None:
        tf = _tf()
        if optim_state_key is None:
            tf.safe_set_full_fp32_param(param, value, optimizer=self)
        else:
            tf.safe_set_full_optimizer_state(param, value, optim_state_key, optimizer=self)

    def get_fp32_grad_for_param(self, param):
        tf = _tf()
        return tf.safe_get_full_grad(param, optimizer=self)

    def set_fp32_grad_for_param(self, value, param) -> None:
        tf = _tf()
        tf.safe_set_full_grad(param, value, optimizer=self)

    def get_local_fp32_param(self, param, optim_state_key: Optional[str] = None):
        tf = _tf()
        return tf.safe_get_local_fp32_param(param, optimizer=self) if optim_state_key is None else tf.safe_get_local_optimizer_state(param, optim_state_key, optimizer=self)

    def set_local_hp_param(self, value, param, optim_state_key: Optional[str] = None) -> None:
        tf = _tf()
        if optim_state_key is None:
            tf.safe_set_local_fp32_param(param, value, optimizer=self)
        else:
            tf.safe_set_local_optimizer_state(param, value, optim_state_key, optimizer=self)

    def get_local_fp32_grad_for_param(self, param):
        tf = _tf()
        return tf.safe_get_local_grad(param, optimizer=self)

    def set_local_grad_for_param(self, value, param) -> None:
        tf = _tf()
        tf.safe_set_local_grad(param, value, optimizer=self)

    # ---- misc ----------------------------------------------------------------------------------------------------------------
    def get_lean_optimizer_state(self):
        """Optimizer state of this rank with the alignment padding of the flat partitions cut off, per parameter group."""
        out = []
        pads = getattr(self, "groups_padding", None)
        for gi, master in enumerate(self._api_master_tensors()):
            st = self.optimizer.state.get(master, {})
            pad = pads[gi] if pads is not None and self.rank == self.partition_count - 1 else 0
            n = master.numel() - pad
            out.append({k: (v.narrow(0, 0, n) if torch.is_tensor(v) and v.dim() and v.numel() == master.numel() else v)
                        for k, v in st.items()})
        return out

    def get_groups_without_padding(self, groups_with_padding):
        pads = getattr(self, "groups_padding", None)
        out = []
        for gi, t in enumerate(groups_with_padding):
            pad = pads[gi] if pads is not None and self.rank == self.partition_count - 1 else 0
            out.append(t.narrow(0, 0, t.numel() - pad))
        return out

    def destroy(self) -> None:
        """Take the optimizer's hooks off the module (the engine calls this from ``engine.destroy()``)."""
        off = getattr(self, "parameter_offload", None)
        if off is not None and hasattr(off, "destroy"):
            off.destroy()
        for h in getattr(self, "_grad_hooks", []) or []:
            h.remove()
        self._grad_hooks = []
