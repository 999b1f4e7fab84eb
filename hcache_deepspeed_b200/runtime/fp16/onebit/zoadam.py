"""0/1 Adam: variance is refreshed only at exponentially spaced steps (``var_update_scaler``), and between
synchronisation points ranks take purely LOCAL steps (``local_step_scaler`` / ``local_step_clipper``), exchanging
1-bit compressed momentum when they do synchronise. Reference: ``deepspeed/runtime/fp16/onebit/zoadam.py:14-359``."""
from __future__ import annotations

import torch

from ...comm.compressed import CompressedBackend


class ZeroOneAdam(torch.optim.Optimizer):

    def __init__(self, params, deepspeed=None, lr=1e-3, bias_correction=True, betas=(0.9, 0.999), eps=1e-8,
                 eps_inside_sqrt=False, weight_decay=0.0, max_grad_norm=0.0, var_freeze_step=100000, var_update_scaler=16,
                 local_step_scaler=32678, local_step_clipper=16, amsgrad=False, cuda_aware=False, comm_backend_name="nccl"):
        if amsgrad:
            raise RuntimeError("0/1 Adam does not support the AMSGrad variant.")
        super().__init__(params, dict(lr=lr, bias_correction=bias_correction, betas=betas, eps=eps, weight_decay=weight_decay,
                                      max_grad_norm=max_grad_norm))
        self.deepspeed = deepspeed
        self.var_freeze_step, self.var_update_scaler = var_freeze_step, var_update_scaler
        self.local_step_scaler, self.local_step_clipper = local_step_scaler, local_step_clipper
        self.comm_backend_handle = CompressedBackend(getattr(deepspeed, "mpu", None))
        self.size = self.comm_backend_handle.size
        self.var_interval, self.var_counter = 1, 0
        self.local_step_interval, self.local_step_counter = 1, 0
        self.freeze_key = False

    @torch.no_grad()
    def step(self, closure=None, grads=None):
        loss = closure() if closure is not None else None
        step_id = None
        for group in self.param_groups:
            beta1, beta2 = group["betas"]
            for p in group["params"]:
                if p.grad is None:
                    continue
                g = p.grad.data.float()
                st = self.state[p]
                if not st:
                    st.update(step=0, exp_avg=torch.zeros_like(p, dtype=torch.float32),
                              exp_avg_sq=torch.zeros_like(p, dtype=torch.float32))
                    P = self.comm_backend_handle.padded_size(p.numel())
                    st["worker_error"] = torch.zeros(P, dtype=torch.float32, device=p.device)
                    st["server_error"] = torch.zeros(P // max(1, self.size), dtype=torch.float32, device=p.device)
                st["step"] += 1
                step_id = st["step"]
                m, v = st["exp_avg"], st["exp_avg_sq"]
                update_var = (not self.freeze_key) and (step_id % self.var_interval == 0)
                m.mul_(beta1).add_(g, alpha=1 - beta1)
                if update_var:
                    v.mul_(beta2).addcmul_(g, g, value=1 - beta2)
                sync_now = self.freeze_key and self.size > 1 and (step_id % self.local_step_interval == 0)
                if sync_now:
                    self.comm_backend_handle.compressed_allreduce(m, st["worker_error"], st["server_error"])
                update = m / (v.sqrt() + group["eps"])
                if group["weight_decay"] > 0.0:
                    update = update + group["weight_decay"] * p.data.float()
                p.data.add_((-group["lr"] * update).to(p.dtype))
        if step_id is not None:
            if not self.freeze_key:
                if step_id % self.var_interval == 0:
                    self.var_counter += 1
                    if self.var_counter == self.var_update_scaler:
                        self.var_counter, self.var_interval = 0, self.var_interval * 2
                if step_id >= self.var_freeze_step:
                    self.freeze_key = True
                    if self.deepspeed is not None:
                        self.deepspeed.enable_backward_allreduce = False
            else:
                self.local_step_counter += 1
                if self.local_step_counter == self.local_step_scaler:
                    self.local_step_counter = 0
                    self.local_step_interval = min(self.local_step_clipper, self.local_step_interval * 2)
        return loss
