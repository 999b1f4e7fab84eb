"""Coalesced collectives for ZeRO: one flat reduce-scatter for a list of tensors, and the ZeRO++ quantised
gradient reduction (qgZ: int4/int8 codes, intra-node then inter-node all-to-all, dequant-reduce-requant between).
Reference: ``deepspeed/runtime/comm/coalesced_collectives.py:31-155``."""
from __future__ import annotations

from typing import List

import torch

from ... import comm as dist
from ...ops.quantizer import dequantize_int, quantize_int


def reduce_scatter_coalesced(tensors: List[torch.Tensor], group=None) -> List[torch.Tensor]:
    """Each tensor is (zero-padded and) split into ``world`` parts; rank r gets the averaged r-th part of every tensor.
    All tensors travel in ONE reduce-scatter."""
    ws = dist.get_world_size(group) if dist.is_initialized() else 1
    rank = dist.get_rank(group) if dist.is_initialized() else 0
    psizes = [(t.numel() + ws - 1) // ws for t in tensors]
    if ws == 1:
        return [t.reshape(-1).clone() for t in tensors]
    total = sum(psizes)
    send = torch.zeros(ws * total, dtype=tensors[0].dtype, device=tensors[0].device)
    off = 0
    for t, ps in zip(tensors, psizes):
        flat = t.reshape(-1)
        for r in range(ws):
            lo, hi = r * ps, min(flat.numel(), (r + 1) * ps)
            if lo < hi:
                send[r * total + off:r * total + off + (hi - lo)] = flat[lo:hi]
        off += ps
    out = torch.empty(total, dtype=send.dtype, device=send.device)
    dist.reduce_scatter_tensor(out, send, group=group)
    out /= ws
    res, off = [], 0
    for ps in psizes:
        res.append(out[off:off + ps])
        off += ps
    return res


def all_to_all_quant_reduce(tensors: List[torch.Tensor], groups: dict, bits: int = 4) -> List[torch.Tensor]:
    """ZeRO++ qgZ for a single node (or any flat group): quantise, all-to-all, dequantise + average locally.
    ``groups`` may hold 'local_0' (intra-node group); the inter-node hop degenerates when there is one node."""
    group = groups.get("local_0") if isinstance(groups, dict) else groups
    ws = dist.get_world_size(group) if dist.is_initialized() else 1
    out = []
    for t in tensors:
        flat = t.reshape(-1).float()
        pad = (-flat.numel()) % (ws * 256)
        if pad:
            flat = torch.cat([flat, flat.new_zeros(pad)])
        if ws == 1:
            out.append(flat[:t.numel()])
            continue
        n_groups = flat.numel() // 256
        codes, params = quantize_int(flat, n_groups, bits, symmetric=True)
        rc, rp = torch.empty_like(codes), torch.empty_like(params)
        dist.all_to_all_single(rc, codes, group=group)
        dist.all_to_all_single(rp.view(-1), params.view(-1).contiguous(), group=group)
        deq = dequantize_int(rc, rp, bits, torch.float32).reshape(ws, -1)
        out.append(deq.mean(dim=0))
    return out


def all_to_all_loco_quant_reduce(params, groups: dict, loco_param=None, bits: int = 4) -> List[torch.Tensor]:
    """ZeRO++ qgZ with LoCo error feedback (ref: coalesced_collectives.py:81): like ``all_to_all_quant_reduce`` over the
    gradients of ``params``, but what quantisation lost on the previous call (kept per parameter in ``p.intra_ef_buf``,
    scaled by ``err_beta``, reset every ``reset_T`` calls) is added back before quantising. 1-D gradients and sizes the
    exchange cannot split evenly go through the plain reduce-scatter."""
    loco_param = loco_param or {}
    beta, reset_t = float(loco_param.get("err_beta", 0.8)), int(loco_param.get("reset_T", 1024))
    group = groups.get("local_0") if isinstance(groups, dict) else groups
    ws = dist.get_world_size(group) if dist.is_initialized() else 1
    out: List[torch.Tensor] = []
    for p in params:
        g = p.grad
        if g.dim() == 1 or g.numel() % (2 * ws) != 0 or ws == 1:
            out.append(reduce_scatter_coalesced([g], group=group)[0] if ws > 1 else g.reshape(-1).float())
            continue
        flat = g.reshape(-1).float()
        pad = (-flat.numel()) % (ws * 256)
        if pad:
            flat = torch.cat([flat, flat.new_zeros(pad)])
        err, age = getattr(p, "intra_ef_buf", (None, 0))
        if err is None or age >= reset_t or err.shape != flat.shape:
            err, age = torch.zeros_like(flat), 0
        flat = flat + beta * err
        codes, qp = quantize_int(flat, flat.numel() // 256, bits, symmetric=True)
        p.intra_ef_buf = (flat - dequantize_int(codes, qp, bits, torch.float32).reshape(-1), age + 1)
        rc, rp = torch.empty_like(codes), torch.empty_like(qp)
        dist.all_to_all_single(rc, codes, group=group)
        dist.all_to_all_single(rp.view(-1), qp.view(-1).contiguous(), group=group)
        out.append(dequantize_int(rc, rp, bits, torch.float32).reshape(ws, -1).mean(dim=0))
    return out
