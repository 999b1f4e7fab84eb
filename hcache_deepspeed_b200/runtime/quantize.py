"""MoQ (Mixture-of-Quantization): progressively lower the weight precision during training from ``start_bits`` to
``target_bits``, halving-period schedule optionally modulated by the Hessian eigenvalue of each block.
Reference: ``deepspeed/runtime/quantize.py:14-180``."""
from __future__ import annotations

import torch

from ..ops.quantizer import ds_quantizer

TWO_D_PARAMS = 6


class Quantizer:

    def __init__(self, q_groups=1, q_mixed_fp16=False, q_change_ratio=0.01, q_type=0, q_rounding=0, q_verbose=False,
                 q_eigenvalue=False, use_quantizer_kernel=False, layer_num=0):
        self.q_groups, self.q_mixed_fp16, self.q_change_ratio, self.q_type = q_groups, q_mixed_fp16, q_change_ratio, q_type
        self.qsteps, self.quantize_real_ratio = 0, 1.0
        self.q_verbose, self.q_eigenvalue, self.use_quantizer_kernel, self.q_rounding = q_verbose, q_eigenvalue, use_quantizer_kernel, q_rounding
        self.layer_num = layer_num

    def any_precision_switch(self) -> bool:
        if self.layer_num == 0:
            return True
        return any(p.start_bits != p.target_bits for p in getattr(self, "_params", []))

    def quantize(self, parameter_group, overflow: bool, eigenvalue_enabled: bool, block_eigenvalue=None):
        if overflow and not eigenvalue_enabled:
            return
        self.step()
        self.update_fp16_ratio()
        for group in parameter_group:
            for p in group:
                if len(p.size()) > 1 and hasattr(p, "start_bits") and p.start_bits:
                    factor = 1
                    if block_eigenvalue and id(p) in block_eigenvalue:
                        ev, _ = block_eigenvalue[id(p)]
                        factor = 1 + int(ev * 4) if ev is not None else 1
                    p.data = self.compute_quantization(p, factor).data

    def step(self):
        self.qsteps += 1

    def update_fp16_ratio(self):
        if self.q_mixed_fp16:
            self.quantize_real_ratio = max(0.0, self.quantize_real_ratio - self.q_change_ratio)

    def compute_quantization(self, p, factor: int = 1):
        if p.start_bits != p.target_bits and self.qsteps >= p.q_period:
            p.start_bits -= 1
            p.q_period = p.q_period * 2 * factor
            if self.q_verbose:
                print(f"Quantization settings: current bit-precision = {p.start_bits}, step = {self.qsteps}, period = {p.q_period}")
        assert p.start_bits >= p.target_bits, "Quantization bit is lower than target precision bits!"
        q = ds_quantizer(p.data.clone(), self.q_groups, p.start_bits, sr=self.q_rounding == 1, asym=self.q_type == 1)
        if self.q_mixed_fp16 and p.start_bits >= p.target_bits - 1:
            q = p.data * self.quantize_real_ratio + (1 - self.quantize_real_ratio) * q
        return q
