"""Profiles the passes are calibrated with (reference ``deepspeed/compile/profilers/{comm_profile,graph_profile}.py``)."""
from __future__ import annotations

import time
from typing import Dict, List, Tuple

import torch

from .. import comm as dist


def profile_all_gather(sizes_numel: List[int], dtype=torch.bfloat16, device=None, group=None, iters: int = 5) -> List[Tuple[int, float]]:
    """``(numel, seconds)`` of an all-gather whose OUTPUT has ``numel`` elements, max over ranks."""
    ws = dist.get_world_size(group=group) if dist.is_initialized() else 1
    device = device or ("cuda" if torch.cuda.is_available() else "cpu")
    out = []
    for n in sizes_numel:
        shard = max(1, n // ws)
        src = torch.zeros(shard, dtype=dtype, device=device)
        dst = torch.empty(shard * ws, dtype=dtype, device=device)
        best = float("inf")
        for it in range(iters + 1):
            if torch.cuda.is_available():
                torch.cuda.synchronize()
            t0 = time.perf_counter()
            if ws > 1:
                dist.all_gather_into_tensor(dst, src, group=group)
            else:
                dst.copy_(src)
            if torch.cuda.is_available():
                torch.cuda.synchronize()
            if it:
                best = min(best, time.perf_counter() - t0)
        t = torch.tensor([best], dtype=torch.float64, device=device if ws > 1 and torch.cuda.is_available() else "cpu")
        if ws > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX, group=group)
        out.append((n, float(t)))
    return out


def fit_alpha_beta(samples: List[Tuple[int, float]], element_size: int = 2) -> Dict[str, float]:
    """Least-squares ``t = alpha + bytes / bandwidth`` over the samples."""
    if len(samples) < 2:
        return {"alpha_s": samples[0][1] if samples else 0.0, "bandwidth_Bps": float("inf")}
    xs = [n * element_size for n, _ in samples]
    ys = [t for _, t in samples]
    mx, my = sum(xs) / len(xs), sum(ys) / len(ys)
    den = sum((x - mx) ** 2 for x in xs)
    slope = sum((x - mx) * (y - my) for x, y in zip(xs, ys)) / den if den else 0.0
    alpha = max(my - slope * mx, 0.0)
    return {"alpha_s": alpha, "bandwidth_Bps": 1.0 / slope if slope > 0 else float("inf")}
