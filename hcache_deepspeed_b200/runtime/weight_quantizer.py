"""Weight quantiser used while loading Megatron checkpoints for int8 inference: per-group symmetric int8 with extra
grouping for MLP weights; keeps the scales so they can be merged per layer.
Reference: ``deepspeed/runtime/weight_quantizer.py:11-150``."""
from __future__ import annotations

import torch


class WeightQuantization:

    def __init__(self, mlp_extra_grouping: bool = True, mp_size: int = 1):
        self.dense_scales, self.qkv_scales, self.mlp4hh_scales, self.mlph4h_scales = [], [], [], []
        self.mlp_extra_grouping, self.mp_size = mlp_extra_grouping, mp_size

    def quantize_data(self, data: torch.Tensor, quantize_bits: int, groups: int, key=None):
        g = torch.split(data.float().flatten(), data.numel() // groups)
        absmax = torch.stack([t.abs().max() for t in g]).clamp_min(1e-12)
        q_range = 2 ** quantize_bits
        scale = q_range / (2 * absmax)
        q = torch.cat([(t * s).round().clamp(-q_range // 2, q_range // 2 - 1) for t, s in zip(g, scale)])
        return q.reshape(data.shape).to(torch.int8), scale

    def is_mlp(self, data, merge_count: int = 1) -> bool:
        return (self.mp_size * data.shape[0] * merge_count) / data.shape[1] == 4 or \
            (self.mp_size * data.shape[1] * merge_count) / data.shape[0] == 4

    def is_qkv(self, data) -> bool:
        return (self.mp_size * data.shape[0]) / data.shape[1] == 3 or (self.mp_size * data.shape[1]) / data.shape[0] == 3

    def Quantize(self, value_list, quantize_bits: int, groups: int, key: str, merge_dim: int = 0):
        if self.mlp_extra_grouping and self.is_mlp(value_list[0], merge_count=len(value_list)):
            groups *= 2
        q_scale = []
        for i, data in enumerate(value_list):
            dq, s = self.quantize_data(data, quantize_bits, groups, key)
            q_scale.append(s)
            value_list[i] = dq
        scales = 1 / torch.cat(q_scale, dim=merge_dim).view(-1).unsqueeze(0)
        if "mlp.dense_4h_to_h.weight" in key:
            self.mlp4hh_scales.append(scales)
        elif "mlp.dense_h_to_4h.weight" in key:
            self.mlph4h_scales.append(scales)
        elif "attention.query_key_value.weight" in key:
            self.qkv_scales.append(scales)
        else:
            self.dense_scales.append(scales)
        return value_list

    def merge_layer_scales(self, layer_scales):
        m = max(s.numel() for s in layer_scales)
        return torch.cat([torch.cat((s, torch.zeros((1, m - s.numel()), device=s.device)), dim=-1) if s.numel() < m else s
                          for s in layer_scales]).unsqueeze(0)

    def merge_scales(self):
        return torch.cat([self.merge_layer_scales([q, d, h4, hh]) for d, q, h4, hh in
                          zip(self.dense_scales, self.qkv_scales, self.mlph4h_scales, self.mlp4hh_scales)])

    def sd_quantize_megatron(self, sd, quantize_bits: int, groups: int):
        for key in list(sd.keys()):
            if any(k in key for k in ("attention.dense.weight", "mlp.dense_4h_to_h.weight", "mlp.dense_h_to_4h.weight",
                                      "attention.query_key_value.weight")):
                sd[key] = self.Quantize([sd[key]], quantize_bits, groups, key=key)[0]
        return sd, self.merge_scales()
