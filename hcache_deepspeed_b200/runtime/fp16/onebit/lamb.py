"""1-bit LAMB: LAMB during warm-up; afterwards momentum is exchanged 1-bit and the per-layer trust ratios are frozen
at their warm-up moving average, rescaled by the change of the momentum/variance ratio.
Reference: ``deepspeed/runtime/fp16/onebit/lamb.py:15-443``."""
from __future__ import annotations

import torch

from ...comm.compressed import CompressedBackend


class OnebitLamb(torch.optim.Optimizer):

    def __init__(self, params, deepspeed=None, lr=1e-3, freeze_step=100000, bias_correction=True, betas=(0.9, 0.999), eps=1e-8,
                 eps_inside_sqrt=False, weight_decay=0.0, max_grad_norm=0.0, max_coeff=10.0, min_coeff=0.01, amsgrad=False,
                 cuda_aware=False, comm_backend_name="nccl", coeff_beta=0.9, factor_max=4.0, factor_min=0.5, factor_threshold=0.1):
        if amsgrad:
            raise RuntimeError("1-bit Lamb does not support the AMSGrad variant.")
        super().__init__(params, dict(lr=lr, bias_correction=bias_correction, betas=betas, eps=eps, weight_decay=weight_decay,
                                      max_grad_norm=max_grad_norm, max_coeff=max_coeff, min_coeff=min_coeff))
        self.deepspeed, self.freeze_step, self.lamb_freeze_key = deepspeed, freeze_step, False
        self.coeff_beta, self.factor_max, self.factor_min, self.factor_threshold = coeff_beta, factor_max, factor_min, factor_threshold
        self.comm_backend_handle = CompressedBackend(getattr(deepspeed, "mpu", None))
        self.size = self.comm_backend_handle.size
        self.lamb_coeffs = []

    @torch.no_grad()
    def step(self, closure=None, grads=None):
        loss = closure() if closure is not None else None
        self.lamb_coeffs = []
        for group in self.param_groups:
            beta1, beta2 = group["betas"]
            for p in group["params"]:
                if p.grad is None:
                    continue
                g = p.grad.data.float()
                st = self.state[p]
                if not st:
                    st.update(step=0, exp_avg=torch.zeros_like(p, dtype=torch.float32),
                              exp_avg_sq=torch.zeros_like(p, dtype=torch.float32), lamb_coeff_freeze=0.0, last_factor=1.0)
                    P = self.comm_backend_handle.padded_size(p.numel())
                    st["worker_error"] = torch.zeros(P, dtype=torch.float32, device=p.device)
                    st["server_error"] = torch.zeros(P // max(1, self.size), dtype=torch.float32, device=p.device)
                st["step"] += 1
                m, v = st["exp_avg"], st["exp_avg_sq"]
                w = p.data.float()
                if not self.lamb_freeze_key:
                    m.mul_(beta1).add_(g, alpha=1 - beta1)
                    v.mul_(beta2).addcmul_(g, g, value=1 - beta2)
                    update = m / (v.sqrt() + group["eps"]) + group["weight_decay"] * w
                    wn, un = float(w.norm()), float(update.norm())
                    coeff = min(max(wn / un, group["min_coeff"]), group["max_coeff"]) if wn > 0 and un > 0 else 1.0
                    st["lamb_coeff_freeze"] = self.coeff_beta * st["lamb_coeff_freeze"] + (1 - self.coeff_beta) * coeff
                    if st["step"] == self.freeze_step:
                        st["exp_avg_sq_fresh"] = v.clone()
                else:
                    m.mul_(beta1).add_(g, alpha=1 - beta1)
                    if self.size > 1:
                        self.comm_backend_handle.compressed_allreduce(m, st["worker_error"], st["server_error"])
                    fresh = st.setdefault("exp_avg_sq_fresh", v.clone())
                    g_rec = g  # local gradient as the estimate of the fresh variance direction
                    fresh.mul_(beta2).addcmul_(g_rec, g_rec, value=1 - beta2)
                    update = m / (v.sqrt() + group["eps"]) + group["weight_decay"] * w
                    ratio = float(((v.sqrt() + group["eps"]) / (fresh.sqrt() + group["eps"])).mean())
                    factor = min(max(ratio, self.factor_min), self.factor_max)
                    lo, hi = st["last_factor"] * (1 - self.factor_threshold), st["last_factor"] * (1 + self.factor_threshold)
                    factor = min(max(factor, lo), hi)
                    st["last_factor"] = factor
                    coeff = st["lamb_coeff_freeze"] * factor
                self.lamb_coeffs.append(coeff)
                p.data.add_((-group["lr"] * coeff * update).to(p.dtype))
                if st["step"] >= self.freeze_step and not self.lamb_freeze_key:
                    self.lamb_freeze_key = True
                    if self.deepspeed is not None:
                        self.deepspeed.enable_backward_allreduce = False
        return loss

    def get_lamb_coeffs(self):
        return list(self.lamb_coeffs)
