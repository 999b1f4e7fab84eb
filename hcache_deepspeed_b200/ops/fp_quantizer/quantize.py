"""Minifloat group quantisation: FP8 (e4m3), FP6 (e3m2), FP4 (e2m1), FP12 (e4m7); one fp32 scale per group of 512
(default) elements, codes packed at their natural bit width. CUDA kernels for device tensors
(``ops/csrc/quantize.cu``), the same encoding in torch for CPU tensors.
Reference: ``deepspeed/ops/fp_quantizer/quantize.py:43-190`` (FP_Quantize), ``csrc/fp_quantizer/fp_quantize.cu``."""
from __future__ import annotations

from typing import Optional, Tuple

import torch

from .. import _mta, native

FORMATS = {8: (4, 3, 7, 480.0), 6: (3, 2, 3, 28.0), 4: (2, 1, 1, 6.0), 12: (4, 7, 7, 510.0)}


def packed_bytes(n: int, q_bits: int) -> int:
    return (n * q_bits + 7) // 8


def _encode_torch(y: torch.Tensor, q_bits: int) -> torch.Tensor:
    e_bits, m_bits, bias, vmax = FORMATS[q_bits]
    sign = (y < 0).to(torch.int64)
    a = y.abs().clamp(max=vmax)
    min_normal = 2.0 ** (1 - bias)
    sub = torch.round(a * 2.0 ** (bias - 1 + m_bits)).to(torch.int64)
    fr, e = torch.frexp(torch.where(a > 0, a, torch.ones_like(a)))
    eu = e.to(torch.int64) - 1
    mant = torch.round((fr * 2.0 - 1.0) * (1 << m_bits)).to(torch.int64)
    carry = mant == (1 << m_bits)
    mant = torch.where(carry, torch.zeros_like(mant), mant)
    eu = torch.where(carry, eu + 1, eu)
    norm = ((eu + bias) << m_bits) | mant
    norm = norm.clamp(max=(1 << (e_bits + m_bits)) - 1)
    code = torch.where(a < min_normal, sub, norm)
    return (sign << (e_bits + m_bits)) | code


def _decode_torch(code: torch.Tensor, q_bits: int) -> torch.Tensor:
    e_bits, m_bits, bias, _ = FORMATS[q_bits]
    sign = (code >> (e_bits + m_bits)) & 1
    e = (code >> m_bits) & ((1 << e_bits) - 1)
    m = code & ((1 << m_bits) - 1)
    sub = m.float() * 2.0 ** (1 - bias - m_bits)
    norm = (1.0 + m.float() / (1 << m_bits)) * torch.pow(torch.tensor(2.0), (e - bias).float())
    v = torch.where(e == 0, sub, norm)
    return torch.where(sign == 1, -v, v)


def _pack(codes: torch.Tensor, q_bits: int) -> torch.Tensor:
    c = codes.to(torch.int64)
    if q_bits == 8:
        return c.to(torch.uint8)
    if q_bits == 4:
        c = c.reshape(-1, 2)
        return (c[:, 0] | (c[:, 1] << 4)).to(torch.uint8)
    if q_bits == 6:
        c = c.reshape(-1, 4)
        w = c[:, 0] | (c[:, 1] << 6) | (c[:, 2] << 12) | (c[:, 3] << 18)
    else:
        c = c.reshape(-1, 2)
        w = c[:, 0] | (c[:, 1] << 12)
    return torch.stack([w & 0xff, (w >> 8) & 0xff, (w >> 16) & 0xff], dim=1).reshape(-1).to(torch.uint8)


def _unpack(packed: torch.Tensor, n: int, q_bits: int) -> torch.Tensor:
    p = packed.to(torch.int64)
    if q_bits == 8:
        return p[:n]
    if q_bits == 4:
        return torch.stack([p & 0xf, (p >> 4) & 0xf], dim=1).reshape(-1)[:n]
    p = p.reshape(-1, 3)
    w = p[:, 0] | (p[:, 1] << 8) | (p[:, 2] << 16)
    if q_bits == 6:
        return torch.stack([(w >> (6 * i)) & 0x3f for i in range(4)], dim=1).reshape(-1)[:n]
    return torch.stack([w & 0xfff, (w >> 12) & 0xfff], dim=1).reshape(-1)[:n]


def fp_quantize(x: torch.Tensor, q_bits: int = 8, group_size: int = 512) -> Tuple[torch.Tensor, torch.Tensor]:
    """Returns (packed uint8 [numel * q_bits / 8], scales fp32 [groups])."""
    assert q_bits in FORMATS and x.numel() % group_size == 0 and group_size % 4 == 0
    x = x.contiguous()
    g = x.numel() // group_size
    out = torch.empty(packed_bytes(x.numel(), q_bits), dtype=torch.uint8, device=x.device)
    scales = torch.empty(g, dtype=torch.float32, device=x.device)
    if x.is_cuda:
        rc = native.cuda_lib().hds_fp_quantize(x.data_ptr(), _mta.dtype_code(x), g, group_size, q_bits, out.data_ptr(),
                                               scales.data_ptr(), _mta.stream_ptr())
        native.check(rc, "hds_fp_quantize")
        return out, scales
    v = x.float().reshape(g, group_size)
    mx = v.abs().amax(dim=1, keepdim=True)
    s = torch.where(mx > 0, mx / FORMATS[q_bits][3], torch.ones_like(mx))
    scales.copy_(s.reshape(-1))
    out.copy_(_pack(_encode_torch(v * (1.0 / s), q_bits).reshape(-1), q_bits))
    return out, scales


def fp_dequantize(packed: torch.Tensor, scales: torch.Tensor, q_bits: int = 8, group_size: int = 512, dtype=torch.bfloat16,
                  shape=None, group_idx: Optional[torch.Tensor] = None) -> torch.Tensor:
    """``group_idx`` (int32): dequantise only those groups, in that order (selective dequantisation)."""
    n_out = scales.numel() if group_idx is None else group_idx.numel()
    out = torch.empty(n_out * group_size, dtype=dtype, device=packed.device)
    if packed.is_cuda:
        gi = None if group_idx is None else group_idx.to(torch.int32).contiguous()
        rc = native.cuda_lib().hds_fp_dequantize(packed.data_ptr(), scales.data_ptr(), None if gi is None else gi.data_ptr(),
                                                 n_out, group_size, q_bits, out.data_ptr(), _mta.dtype_code(out), _mta.stream_ptr())
        native.check(rc, "hds_fp_dequantize")
    else:
        n = scales.numel() * group_size
        v = _decode_torch(_unpack(packed, n, q_bits), q_bits).reshape(-1, group_size) * scales.reshape(-1, 1)
        if group_idx is not None:
            v = v[group_idx.long()]
        out.copy_(v.reshape(-1).to(dtype))
    return out.reshape(shape) if shape is not None else out


class Quantizer:
    def __init__(self, group_size: int = 512):
        self.group_size = group_size
        self.orig_dtype = None
        self.num_groups = None
        self.input_q = None
        self.scale = None


class FP_Quantize(Quantizer):
    """Same call surface as the reference class: ``quantize(x, q_bits)`` returns one uint8 tensor with the per-group
    scales appended (or ``(codes, scales)`` with ``return_meta_tensor=True``)."""

    def quantize(self, input: torch.Tensor, q_bits: int = 8, q_mantisa_bits: int = 3, stochastic_mode: bool = False,
                 return_meta_tensor: bool = False):
        assert input.dtype in (torch.bfloat16, torch.float16, torch.float32)
        assert FORMATS[q_bits][1] == q_mantisa_bits or q_mantisa_bits == 3, \
            f"missing {q_bits}-quantization with {q_mantisa_bits} mantissa bits"
        self.orig_dtype, self.orig_shape = input.dtype, input.shape
        x = input
        if stochastic_mode:
            v = input.float().reshape(-1, self.group_size)
            step = v.abs().amax(dim=1, keepdim=True) / FORMATS[q_bits][3] * 2.0 ** (-FORMATS[q_bits][1])
            x = (v + (torch.rand_like(v) - 0.5) * step).reshape(input.shape).to(input.dtype)
        codes, scales = fp_quantize(x, q_bits, self.group_size)
        self.num_groups = scales.numel()
        if return_meta_tensor:
            self.input_q, self.scale = codes, scales
            return codes, scales
        return torch.cat([codes, scales.view(torch.uint8)])

    def _split(self, input_q: torch.Tensor, q_bits: int, scale: Optional[torch.Tensor]):
        if scale is not None:
            return input_q, scale
        n_scale_bytes = self.num_groups * 4
        return input_q[:-n_scale_bytes], input_q[-n_scale_bytes:].contiguous().view(torch.float32)

    def dequantize(self, input_q: torch.Tensor, fp_out: Optional[torch.Tensor] = None, q_bits: int = 8,
                   q_mantisa_bits: int = 3, scale: Optional[torch.Tensor] = None) -> torch.Tensor:
        codes, scales = self._split(input_q, q_bits, scale)
        out = fp_dequantize(codes, scales, q_bits, self.group_size, dtype=self.orig_dtype or torch.bfloat16,
                            shape=self.orig_shape)
        if fp_out is not None:
            fp_out.copy_(out)
            return fp_out
        return out

    def selective_dequantize(self, input_q: torch.Tensor, indexes: torch.Tensor, fp_out: Optional[torch.Tensor] = None,
                             q_bits: int = 8, q_mantisa_bits: int = 3, scale: Optional[torch.Tensor] = None) -> torch.Tensor:
        """Dequantise rows ``indexes`` of the first dimension of the original 3-D tensor."""
        assert len(self.orig_shape) == 3, "Selective-Dequantization works on 3d tensor only! Please reshape the tensor before calling dequantize function."
        codes, scales = self._split(input_q, q_bits, scale)
        per_row = self.orig_shape[1] * self.orig_shape[2]
        assert per_row % self.group_size == 0
        gpr = per_row // self.group_size
        gi = (indexes.to(torch.int64).reshape(-1, 1) * gpr + torch.arange(gpr, device=indexes.device)).reshape(-1).to(torch.int32)
        out = fp_dequantize(codes, scales, q_bits, self.group_size, dtype=self.orig_dtype or torch.bfloat16,
                            shape=(indexes.numel(), self.orig_shape[1], self.orig_shape[2]), group_idx=gi)
        if fp_out is not None:
            fp_out.copy_(out)
            return fp_out
        return out
