"""DeepSpeedCPUAdam: Adam / AdamW over fp32 master parameters in host memory (ZeRO-Offload), executed by the
vectorised, multi-threaded loop in ``libhds_host.so`` (``csrc_host/cpu_optimizers.cpp``). ``step`` can write the
updated values straight into a bf16/fp16 shadow (``fp16_param_groups``) in the same pass.
API follows ``deepspeed/ops/adam/cpu_adam.py:17-186``."""
from __future__ import annotations

import torch

from .. import _mta, native


class DeepSpeedCPUAdam(torch.optim.Optimizer):
    optimizer_id = 0

    def __init__(self, model_params, lr=1e-3, bias_correction=True, betas=(0.9, 0.999), eps=1e-8, weight_decay=0,
                 amsgrad=False, adamw_mode=True, fp32_optimizer_states=True, threads: int = 0):
        if amsgrad:
            raise RuntimeError("DeepSpeedCPUAdam does not support the AMSGrad variant.")
        super().__init__(model_params, dict(lr=lr, betas=betas, eps=eps, weight_decay=weight_decay,
                                            bias_correction=bias_correction, amsgrad=amsgrad))
        self.opt_id = DeepSpeedCPUAdam.optimizer_id
        DeepSpeedCPUAdam.optimizer_id += 1
        self.adam_w_mode = adamw_mode
        self.fp32_optimizer_states = fp32_optimizer_states
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
                assert p.device.type == "cpu", "CPUAdam param is on a GPU; offload the optimizer state to use it"
                assert p.dtype == torch.float32 and p.is_contiguous(), "CPUAdam expects contiguous fp32 master params"
                st = self.state[p]
                if not st:
                    st["step"] = 0
                    st["exp_avg"] = torch.zeros_like(p, dtype=torch.float32)
                    st["exp_avg_sq"] = torch.zeros_like(p, dtype=torch.float32)
                st["step"] += 1
                g = p.grad.data
                if g.dtype != torch.float32 or not g.is_contiguous():
                    g = g.float().contiguous()
                shadow = fp16_param_groups[gi][pi].data if fp16_param_groups is not None else None
                host_shadow = shadow if (shadow is not None and shadow.device.type == "cpu") else None
                out_ptr, out_dt = _mta.host_copy_out(host_shadow)
                rc = self._lib.hds_cpu_adam_step(p.data_ptr(), g.data_ptr(), st["exp_avg"].data_ptr(),
                                                 st["exp_avg_sq"].data_ptr(), p.numel(), group["lr"], beta1, beta2,
                                                 group["eps"], group["weight_decay"], st["step"], int(self.adam_w_mode),
                                                 int(group["bias_correction"]), out_ptr, out_dt, self.threads)
                if rc != 0:
                    raise RuntimeError("hds_cpu_adam_step failed")
                if shadow is not None and host_shadow is None:
                    shadow.copy_(p.data, non_blocking=True)
        return loss
