"""Largest Hessian eigenvalue per layer block by power iteration on Hessian-vector products (used to modulate
quantization schedules). Reference: ``deepspeed/runtime/eigenvalue.py:13-149``."""
from __future__ import annotations

import torch

from ..utils.logging import log_dist


class Eigenvalue:

    def __init__(self, verbose=False, max_iter=100, tol=1e-2, stability=0, gas_boundary_resolution=1,
                 layer_name="", layer_num=0):
        self.verbose, self.max_iter, self.tol, self.stability = verbose, max_iter, tol, stability
        self.gas_boundary_resolution, self.layer_name, self.layer_num = gas_boundary_resolution, layer_name, layer_num
        assert len(layer_name) > 0 and layer_num > 0, "eigenvalue needs layer_name and layer_num"

    @staticmethod
    def nan_to_num(x):
        return torch.nan_to_num(x, nan=0.0, posinf=0.0, neginf=0.0)

    def normalize(self, vs):
        norm = sum(float((v * v).sum()) for v in vs) ** 0.5
        return [self.nan_to_num(v / (norm + self.stability)) for v in vs]

    def inner_product(self, xs, ys):
        return sum(float((x * y).sum()) for x, y in zip(xs, ys))

    def get_layers(self, module):
        obj = module
        for name in self.layer_name.split("."):
            obj = getattr(obj, name)
        return obj

    def compute_eigenvalue(self, module, device=None, scale: float = 1.0):
        """Returns {param_id: (eigenvalue / max eigenvalue, layer index)}; requires gradients created with
        ``create_graph=True`` on the last backward."""
        block_eigenvalue = []
        layers = self.get_layers(module)
        for block in range(self.layer_num):
            params = [p for p in layers[block].parameters() if p.grad is not None and p.grad.grad_fn is not None]
            if not params:
                block_eigenvalue.append(1.0)
                continue
            grads = [p.grad for p in params]
            v = self.normalize([torch.randn_like(p) for p in params])
            cur, prev, i = 1.0, 0.0, 0
            while i < self.max_iter and abs(cur) > 0 and abs((cur - prev) / cur) >= self.tol:
                prev = cur
                hv = torch.autograd.grad(grads, params, grad_outputs=v, only_inputs=True, retain_graph=True)
                hv = [self.nan_to_num(h.float()) for h in hv]
                cur = self.inner_product(hv, v)
                v = self.normalize(hv)
                v = [x / scale for x in v]
                i += 1
            block_eigenvalue.append(cur * scale)
            if self.verbose:
                log_dist(f"block: {block}, power iteration: {i}, eigenvalue: {cur * scale}", ranks=[0])
        block_eigenvalue = self.post_process(block_eigenvalue)
        out = {}
        for block in range(self.layer_num):
            for p in layers[block].parameters():
                out[id(p)] = (block_eigenvalue[block], block)
        return out

    def post_process(self, values):
        m = max((abs(v) for v in values), default=1.0) or 1.0
        return [abs(v) / m if v != 0.0 else 1.0 for v in values]
