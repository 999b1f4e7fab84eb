"""Integer group quantisation (int8 / int4, symmetric / asymmetric) — CUDA kernels on device tensors
(``ops/csrc/quantize.cu``), the same arithmetic in torch on CPU tensors — plus the ZeRO++ helpers built on it
(swizzled quantisation for the hierarchical all-to-all, dequant-reduce-requant).
Reference: ``deepspeed/ops/quantizer/quantizer.py``, ``csrc/quantization/{quantize,dequantize,swizzled_quantize,
quant_reduce,fake_quantizer}.cu``."""
from __future__ import annotations

from typing import Tuple

import torch

from .. import _mta, native


def _groups(t: torch.Tensor, groups: int) -> Tuple[int, int]:
    n = t.numel()
    assert n % groups == 0, f"tensor of {n} elements cannot be split into {groups} groups"
    return groups, n // groups


def quantize_int(x: torch.Tensor, groups: int, bits: int = 8, symmetric: bool = True):
    """Returns (codes int8 [numel * bits / 8], params fp32 [groups, 2] = (scale, offset))."""
    g, gs = _groups(x, groups)
    assert bits in (4, 8) and gs % 2 == 0
    x = x.contiguous()
    codes = torch.empty(x.numel() * bits // 8, dtype=torch.int8, device=x.device)
    params = torch.empty(g, 2, dtype=torch.float32, device=x.device)
    if x.is_cuda:
        rc = native.cuda_lib().hds_quantize_int(x.data_ptr(), _mta.dtype_code(x), g, gs, bits, int(symmetric), codes.data_ptr(),
                                                params.data_ptr(), _mta.stream_ptr())
        native.check(rc, "hds_quantize_int")
        return codes, params
    v = x.float().reshape(g, gs)
    qmax, qmin = float((1 << (bits - 1)) - 1), -float(1 << (bits - 1))
    if symmetric:
        a = v.abs().amax(dim=1, keepdim=True)
        scale = torch.where(a > 0, a / qmax, torch.ones_like(a))
        offset = torch.zeros_like(scale)
    else:
        mx, mn = v.amax(dim=1, keepdim=True), v.amin(dim=1, keepdim=True)
        scale = torch.where(mx > mn, (mx - mn) / float((1 << bits) - 1), torch.ones_like(mx))
        offset = mn - qmin * scale
    q = torch.clamp(torch.round((v - offset) / scale), qmin, qmax).to(torch.int32)
    params.copy_(torch.cat([scale, offset], dim=1))
    if bits == 8:
        codes.copy_(q.reshape(-1).to(torch.int8))
    else:
        q = q.reshape(-1, 2)
        codes.copy_(((q[:, 0] & 0xf) | ((q[:, 1] & 0xf) << 4)).to(torch.uint8).view(torch.int8))
    return codes, params


def dequantize_int(codes: torch.Tensor, params: torch.Tensor, bits: int = 8, dtype=torch.float16, shape=None) -> torch.Tensor:
    g = params.shape[0]
    n = codes.numel() * 8 // bits
    gs = n // g
    out = torch.empty(n, dtype=dtype, device=codes.device)
    if codes.is_cuda:
        rc = native.cuda_lib().hds_dequantize_int(codes.data_ptr(), params.contiguous().data_ptr(), g, gs, bits, out.data_ptr(),
                                                  _mta.dtype_code(out), _mta.stream_ptr())
        native.check(rc, "hds_dequantize_int")
    else:
        if bits == 8:
            q = codes.to(torch.int32)
        else:
            b = codes.view(torch.uint8).to(torch.int32)
            lo, hi = b & 0xf, (b >> 4) & 0xf
            lo, hi = torch.where(lo > 7, lo - 16, lo), torch.where(hi > 7, hi - 16, hi)
            q = torch.stack([lo, hi], dim=1).reshape(-1)
        v = q.float().reshape(g, gs) * params[:, 0:1] + params[:, 1:2]
        out.copy_(v.reshape(-1).to(dtype))
    return out.reshape(shape) if shape is not None else out


def ds_quantizer(x: torch.Tensor, groups: int = 1, bit_num: int = 8, sr: bool = False, asym: bool = False) -> torch.Tensor:
    """Fake quantisation (quantise + dequantise in place, same dtype) — ``ops/quantizer/quantizer.py:21``. ``sr``
    (stochastic rounding) adds uniform noise before rounding."""
    if sr:
        g, gs = _groups(x, groups)
        v = x.float().reshape(g, gs)
        a = v.abs().amax(dim=1, keepdim=True).clamp_min(1e-12)
        step = a / float((1 << (bit_num - 1)) - 1)
        x = (v + (torch.rand_like(v) - 0.5) * step).reshape(x.shape).to(x.dtype)
    bits = 8 if bit_num > 4 else 4
    if bit_num not in (4, 8):  # arbitrary widths: torch arithmetic
        g, gs = _groups(x, groups)
        v = x.float().reshape(g, gs)
        qmax = float((1 << (bit_num - 1)) - 1)
        a = v.abs().amax(dim=1, keepdim=True).clamp_min(1e-12)
        return (torch.clamp(torch.round(v / a * qmax), -qmax - 1, qmax) * a / qmax).reshape(x.shape).to(x.dtype)
    codes, params = quantize_int(x, groups, bits, symmetric=not asym)
    return dequantize_int(codes, params, bits, dtype=x.dtype, shape=x.shape)


class Quantizer:
    """Stateful wrapper (group count, bit width, symmetric) used by weight-quantised layers."""

    def __init__(self, q_bits: int = 8, groups: int = 1, symmetric: bool = True):
        self.q_bits, self.groups, self.symmetric = q_bits, groups, symmetric

    def quantize(self, x):
        return quantize_int(x, self.groups, self.q_bits, self.symmetric)

    def dequantize(self, codes, params, dtype=torch.float16, shape=None):
        return dequantize_int(codes, params, self.q_bits, dtype, shape)


def swizzled_quantize(x: torch.Tensor, groups: int, bits: int, pipeline_size: int, nodes: int, devices_per_node: int):
    """ZeRO++ qgZ first hop: quantise ``x`` and lay the groups out so that the chunk destined for (node n, device d)
    becomes contiguous per *device-major* order — after the intra-node all-to-all every device holds what it must
    forward inter-node. Returns (codes, params) in swizzled group order."""
    parts = nodes * devices_per_node
    assert groups % (parts * pipeline_size) == 0
    codes, params = quantize_int(x, groups, bits, symmetric=True)
    gs_bytes = codes.numel() // groups
    per = groups // (parts * pipeline_size)
    c = codes.reshape(pipeline_size, nodes, devices_per_node, per, gs_bytes).permute(0, 2, 1, 3, 4).contiguous()
    p = params.reshape(pipeline_size, nodes, devices_per_node, per, 2).permute(0, 2, 1, 3, 4).contiguous()
    return c.reshape(-1), p.reshape(-1, 2)


def quantized_reduction(codes: torch.Tensor, params: torch.Tensor, in_groups: int, out_groups: int, bits: int,
                        out_bits: int, devices_per_node: int):
    """ZeRO++ qgZ between the two hops: dequantise the ``devices_per_node`` received chunks, reduce them (sum) and
    requantise the result into ``out_groups`` groups."""
    n = codes.numel() * 8 // bits
    full = dequantize_int(codes, params, bits, dtype=torch.float32).reshape(devices_per_node, n // devices_per_node)
    red = full.sum(dim=0)
    return quantize_int(red, out_groups, out_bits, symmetric=True)
