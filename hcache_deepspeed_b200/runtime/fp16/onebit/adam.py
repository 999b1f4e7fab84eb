"""1-bit Adam: plain Adam for ``freeze_step`` warm-up steps (gradients all-reduced densely by the engine), then the
variance is frozen and only the MOMENTUM is communicated — 1 bit per element with error feedback.
Reference: ``deepspeed/runtime/fp16/onebit/adam.py:14-306``."""
from __future__ import annotations

import torch

from ....utils.logging import logger
from ...comm.compressed import CompressedBackend


class OnebitAdam(torch.optim.Optimizer):

    def __init__(self, params, deepspeed=None, lr=1e-3, freeze_step=100000, bias_correction=True, betas=(0.9, 0.999), eps=1e-8,
                 eps_inside_sqrt=False, weight_decay=0.0, max_grad_norm=0.0, amsgrad=False, cuda_aware=False,
                 comm_backend_name="nccl"):
        if amsgrad:
            raise RuntimeError("1-bit Adam does not support the AMSGrad variant.")
        super().__init__(params, dict(lr=lr, bias_correction=bias_correction, betas=betas, eps=eps,
                                      weight_decay=weight_decay, max_grad_norm=max_grad_norm))
        self.eps_mode = 0 if eps_inside_sqrt else 1
        self.deepspeed = deepspeed
        self.freeze_step = freeze_step
        self.adam_freeze_key = False
        self.initialize = False
        self.comm_backend_name = comm_backend_name
        self.comm_backend_handle = CompressedBackend(getattr(deepspeed, "mpu", None))
        self.size = self.comm_backend_handle.size

    @torch.no_grad()
    def step(self, closure=None, grads=None):
        loss = closure() if closure is not None else None
        for group in self.param_groups:
            beta1, beta2 = group["betas"]
            for p in group["params"]:
                if p.grad is None:
                    continue
                g = p.grad.data.float()
                st = self.state[p]
                if not st:
                    st["step"] = 0
                    st["exp_avg"] = torch.zeros_like(p, dtype=torch.float32)
                    st["exp_avg_sq"] = torch.zeros_like(p, dtype=torch.float32)
                    P = self.comm_backend_handle.padded_size(p.numel())
                    st["worker_error"] = torch.zeros(P, dtype=torch.float32, device=p.device)
                    st["server_error"] = torch.zeros(P // max(1, self.size), dtype=torch.float32, device=p.device)
                st["step"] += 1
                m, v = st["exp_avg"], st["exp_avg_sq"]
                if not self.adam_freeze_key:
                    m.mul_(beta1).add_(g, alpha=1 - beta1)
                    v.mul_(beta2).addcmul_(g, g, value=1 - beta2)
                else:
                    # compression phase: local momentum update with the LOCAL gradient, then 1-bit averaged momentum
                    m.mul_(beta1).add_(g, alpha=1 - beta1)
                    if self.size > 1:
                        self.comm_backend_handle.compressed_allreduce(m, st["worker_error"], st["server_error"])
                update = m / (v.sqrt() + group["eps"])
                if group["weight_decay"] > 0.0:
                    update = update + group["weight_decay"] * p.data.float()
                p.data.add_((-group["lr"] * update).to(p.dtype))
                if st["step"] >= self.freeze_step and not self.adam_freeze_key:
                    self.adam_freeze_key = True
                    if self.deepspeed is not None:
                        self.deepspeed.enable_backward_allreduce = False  # gradients stay local from now on
                    logger.info("1-bit Adam: variance frozen, switching to compressed momentum communication")
        return loss

    def load_state_dict(self, state_dict):
        super().load_state_dict(state_dict)
        steps = [st.get("step", 0) for st in self.state.values()]
        self.adam_freeze_key = bool(steps) and max(steps) >= self.freeze_step
        for st in self.state.values():  # error-feedback buffers restart from zero after a resume
            for k in ("worker_error", "server_error"):
                if k in st:
                    st[k].zero_()
