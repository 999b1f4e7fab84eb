"""FusedAdam: Adam / AdamW whose whole step is one multi-tensor kernel launch per <= 24 tensors
(``ops/csrc/optimizers.cu``). On CPU tensors (tests, CPU configuration) the same update runs through
``torch._foreach`` ops. API follows the reference (``deepspeed/ops/adam/fused_adam.py:18-196``): ``adam_w_mode``,
``bias_correction``, ``set_grad_none``; ``amsgrad`` is rejected."""
from __future__ import annotations

import torch

from .. import _mta


class FusedAdam(torch.optim.Optimizer):

    def __init__(self, params, lr=1e-3, bias_correction=True, betas=(0.9, 0.999), eps=1e-8, adam_w_mode=True,
                 weight_decay=0.0, amsgrad=False, set_grad_none=True):
        if amsgrad:
            raise RuntimeError("FusedAdam does not support the AMSGrad variant.")
        super().__init__(params, dict(lr=lr, bias_correction=bias_correction, betas=betas, eps=eps,
                                      weight_decay=weight_decay))
        self.adam_w_mode = 1 if adam_w_mode else 0
        self.set_grad_none = set_grad_none

    def zero_grad(self, set_to_none: bool = None):
        super().zero_grad(set_to_none=self.set_grad_none if set_to_none is None else set_to_none)

    @torch.no_grad()
    def step(self, closure=None, grads=None, output_params=None, scale=None, grad_norms=None, grad_scaler=None):
        if any(x is not None for x in (grads, output_params, scale, grad_norms)):
            raise RuntimeError("FusedAdam has been updated; use the plain step() (no explicit grads/scale arguments).")
        loss = closure() if closure is not None else None
        for group in self.param_groups:
            if not group["params"]:
                continue
            beta1, beta2 = group["betas"]
            group["step"] = group.get("step", 0) + 1
            ps, gs, ms, vs = [], [], [], []
            for p in group["params"]:
                if p.grad is None:
                    continue
                if p.grad.is_sparse:
                    raise RuntimeError("FusedAdam does not support sparse gradients")
                st = self.state[p]
                if not st:
                    st["exp_avg"] = torch.zeros_like(p, dtype=torch.float32)
                    st["exp_avg_sq"] = torch.zeros_like(p, dtype=torch.float32)
                ps.append(p.data)
                gs.append(p.grad.data.contiguous())
                ms.append(st["exp_avg"])
                vs.append(st["exp_avg_sq"])
            if not ps:
                continue
            if ps[0].is_cuda:
                for bp, bg, bm, bv in _mta.group_by_dtype(ps, gs, ms, vs):
                    _mta.multi_tensor_adam(bp, bg, bm, bv, group["lr"], beta1, beta2, group["eps"],
                                           group["weight_decay"], group["step"], self.adam_w_mode,
                                           group["bias_correction"])
            else:
                _foreach_adam(ps, gs, ms, vs, group["lr"], beta1, beta2, group["eps"], group["weight_decay"],
                              group["step"], self.adam_w_mode, group["bias_correction"])
        return loss


def _foreach_adam(ps, gs, ms, vs, lr, beta1, beta2, eps, wd, step, adamw, bias_correction):
    """The kernel's update restated with torch ops (CPU path and test oracle)."""
    bc1 = 1 - beta1 ** step if bias_correction else 1.0
    bc2 = 1 - beta2 ** step if bias_correction else 1.0
    for p, g, m, v in zip(ps, gs, ms, vs):
        w = p.float()
        g = g.float()
        if not adamw and wd != 0:
            g = g + wd * w
        m.mul_(beta1).add_(g, alpha=1 - beta1)
        v.mul_(beta2).addcmul_(g, g, value=1 - beta2)
        upd = (m / bc1) / ((v / bc2).sqrt() + eps)
        if adamw and wd != 0:
            upd = upd + wd * w
        p.copy_((w - lr * upd).to(p.dtype))
