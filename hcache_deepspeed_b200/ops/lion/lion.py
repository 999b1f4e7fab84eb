"""Lion (sign-momentum) optimizers: ``FusedLion`` (multi-tensor CUDA kernel / torch on CPU tensors) and
``DeepSpeedCPULion`` (host loop for ZeRO-Offload). Ref: ``deepspeed/ops/lion/fused_lion.py``, ``cpu_lion.py``."""
from __future__ import annotations

import torch

from .. import _mta, native


class FusedLion(torch.optim.Optimizer):

    def __init__(self, params, lr=1e-3, betas=(0.9, 0.999), weight_decay=0.0, set_grad_none=True):
        super().__init__(params, dict(lr=lr, betas=betas, weight_decay=weight_decay))
        self.set_grad_none = set_grad_none

    def zero_grad(self, set_to_none: bool = None):
        super().zero_grad(set_to_none=self.set_grad_none if set_to_none is None else set_to_none)

    @torch.no_grad()
    def step(self, closure=None):
        loss = closure() if closure is not None else None
        for group in self.param_groups:
            beta1, beta2 = group["betas"]
            ps, gs, ms = [], [], []
            for p in group["params"]:
                if p.grad is None:
                    continue
                st = self.state[p]
                if not st:
                    st["exp_avg"] = torch.zeros_like(p, dtype=torch.float32)
                ps.append(p.data)
                gs.append(p.grad.data.contiguous())
                ms.append(st["exp_avg"])
            if not ps:
                continue
            if ps[0].is_cuda:
                for bp, bg, bm in _mta.group_by_dtype(ps, gs, ms):
                    _mta.multi_tensor_lion(bp, bg, bm, group["lr"], beta1, beta2, group["weight_decay"])
            else:
                for p, g, m in zip(ps, gs, ms):
                    lion_reference(p, g, m, group["lr"], beta1, beta2, group["weight_decay"])
        return loss


def lion_reference(p, g, m, lr, beta1, beta2, wd):
    w, g = p.float(), g.float()
    upd = (beta1 * m + (1 - beta1) * g).sign()
    p.copy_((w * (1 - lr * wd) - lr * upd).to(p.dtype))
    m.mul_(beta2).add_(g, alpha=1 - beta2)


class DeepSpeedCPULion(torch.optim.Optimizer):

    def __init__(self, model_params, lr=1e-3, betas=(0.9, 0.999), weight_decay=0, threads: int = 0):
        super().__init__(model_params, dict(lr=lr, betas=betas, weight_decay=weight_decay))
        self.threads = threads
        self._lib = native.host_lib()

    @torch.no_grad()
    def step(self, closure=None, fp16_param_groups=None):
        loss = closure() if closure is not None else None
        for gi, group in enumerate(self.param_groups):
            beta1, beta2 = group["betas"]
            for pi, p in enumerate(group["params"]):
                if p.grad is None:
                    continue
                assert p.device.type == "cpu" and p.dtype == torch.float32 and p.is_contiguous()
                st = self.state[p]
                if not st:
                    st["exp_avg"] = torch.zeros_like(p)
                g = p.grad.data.float().contiguous()
                shadow = fp16_param_groups[gi][pi].data if fp16_param_groups is not None else None
                host_shadow = shadow if (shadow is not None and shadow.device.type == "cpu") else None
                out_ptr, out_dt = _mta.host_copy_out(host_shadow)
                rc = self._lib.hds_cpu_lion_step(p.data_ptr(), g.data_ptr(), st["exp_avg"].data_ptr(), p.numel(),
                                                 group["lr"], beta1, beta2, group["weight_decay"], out_ptr, out_dt,
                                                 self.threads)
                if rc != 0:
                    raise RuntimeError("hds_cpu_lion_step failed")
                if shadow is not None and host_shadow is None:
                    shadow.copy_(p.data, non_blocking=True)
        return loss
