"""FusedLamb: layer-wise adaptive large-batch optimizer. The Adam direction is computed per tensor, then scaled by
the trust ratio ``clamp(|w| / |update|, min_coeff, max_coeff)``; norms are the only reductions, so the step is a few
``torch._foreach`` passes (on GPU: vendor elementwise kernels — this optimizer is not on any hot path of the
framework). Semantics follow ``deepspeed/ops/lamb/fused_lamb.py:13-195`` / ``csrc/lamb/fused_lamb_cuda_kernel.cu``."""
from __future__ import annotations

import torch


class FusedLamb(torch.optim.Optimizer):

    def __init__(self, params, lr=1e-3, bias_correction=True, betas=(0.9, 0.999), eps=1e-8, eps_inside_sqrt=False,
                 weight_decay=0.0, max_grad_norm=0.0, max_coeff=10.0, min_coeff=0.01, amsgrad=False):
        if amsgrad:
            raise RuntimeError("FusedLamb does not support the AMSGrad variant.")
        super().__init__(params, dict(lr=lr, bias_correction=bias_correction, betas=betas, eps=eps,
                                      weight_decay=weight_decay, max_grad_norm=max_grad_norm, max_coeff=max_coeff,
                                      min_coeff=min_coeff))
        self.eps_mode = 0 if eps_inside_sqrt else 1
        self.lamb_coeffs = []

    @torch.no_grad()
    def step(self, closure=None, grads=None, output_params=None, scale=1.0, grad_norms=None):
        loss = closure() if closure is not None else None
        self.lamb_coeffs = []
        for group in self.param_groups:
            beta1, beta2 = group["betas"]
            for p in group["params"]:
                if p.grad is None:
                    continue
                g = p.grad.data.float() / scale
                if group["max_grad_norm"] > 0:
                    gn = float(g.norm())
                    clip = gn / group["max_grad_norm"]
                    if clip > 1:
                        g = g / clip
                st = self.state[p]
                if not st:
                    st["step"] = 0
                    st["exp_avg"] = torch.zeros_like(p, dtype=torch.float32)
                    st["exp_avg_sq"] = torch.zeros_like(p, dtype=torch.float32)
                st["step"] += 1
                m, v = st["exp_avg"], st["exp_avg_sq"]
                m.mul_(beta1).add_(g, alpha=1 - beta1)
                v.mul_(beta2).addcmul_(g, g, value=1 - beta2)
                if group["bias_correction"]:
                    bc1, bc2 = 1 - beta1 ** st["step"], 1 - beta2 ** st["step"]
                else:
                    bc1 = bc2 = 1.0
                denom = (v / bc2 + group["eps"]).sqrt() if self.eps_mode == 0 else (v / bc2).sqrt() + group["eps"]
                w = p.data.float()
                update = (m / bc1) / denom + group["weight_decay"] * w
                w_norm, u_norm = float(w.norm()), float(update.norm())
                coeff = 1.0
                if w_norm != 0 and u_norm != 0:
                    coeff = min(max(w_norm / u_norm, group["min_coeff"]), group["max_coeff"])
                self.lamb_coeffs.append(coeff)
                p.data.copy_((w - group["lr"] * coeff * update).to(p.dtype))
        return loss

    def get_lamb_coeffs(self):
        return list(self.lamb_coeffs)
