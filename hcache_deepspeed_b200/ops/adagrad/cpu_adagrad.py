"""DeepSpeedCPUAdagrad: Adagrad over host fp32 master parameters (ZeRO-Offload). Ref: ``deepspeed/ops/adagrad/cpu_adagrad.py``."""
from __future__ import annotations

import torch

from .. import _mta, native


class DeepSpeedCPUAdagrad(torch.optim.Optimizer):

    def __init__(self, model_params, lr=1e-2, eps=1e-10, weight_decay=0, amsgrad=False, fp32_optimizer_states=True,
                 threads: int = 0):
        super().__init__(model_params, dict(lr=lr, eps=eps, weight_decay=weight_decay, amsgrad=amsgrad))
        self.threads = threads
        self._lib = native.host_lib()

    @torch.no_grad()
    def step(self, closure=None, fp16_param_groups=None):
        loss = closure() if closure is not None else None
        for gi, group in enumerate(self.param_groups):
            for pi, p in enumerate(group["params"]):
                if p.grad is None:
                    continue
                assert p.device.type == "cpu" and p.dtype == torch.float32 and p.is_contiguous()
                st = self.state[p]
                if not st:
                    st["step"] = 0
                    st["exp_avg_sq"] = torch.zeros_like(p)
                st["step"] += 1
                if p.grad.is_sparse:
                    g = p.grad.coalesce().to_dense().float().contiguous()
                else:
                    g = p.grad.data.float().contiguous()
                shadow = fp16_param_groups[gi][pi].data if fp16_param_groups is not None else None
                host_shadow = shadow if (shadow is not None and shadow.device.type == "cpu") else None
                out_ptr, out_dt = _mta.host_copy_out(host_shadow)
                rc = self._lib.hds_cpu_adagrad_step(p.data_ptr(), g.data_ptr(), st["exp_avg_sq"].data_ptr(), p.numel(),
                                                    group["lr"], group["eps"], group["weight_decay"], out_ptr, out_dt,
                                                    self.threads)
                if rc != 0:
                    raise RuntimeError("hds_cpu_adagrad_step failed")
                if shadow is not None and host_shadow is None:
                    shadow.copy_(p.data, non_blocking=True)
        return loss
