"""Bubble-free restore planner: which layers come back as hidden states, which as KV pages, which are recomputed.

Restoring layer ``l`` from hidden states moves ``T*H*e`` bytes and costs one K/V projection; restoring it from
offloaded KV moves ``T*2*kv_dim*e`` bytes and costs (almost) no compute; recomputing it moves nothing and costs a
full layer forward — but only a *prefix* of layers can be recomputed (layer ``l`` needs layer ``l-1``'s output).
Transfers (copy engine / NVMe) and compute run on different streams, so restore time is
``max(total_io, total_compute)`` plus pipeline fill; the plan balances the two (HCache paper §4: mix hidden-state
restore with KV offload so neither the link nor the GPU idles).

On B200 the host link is the scarce resource (PCIe Gen5 x16 ≈ 55 GB/s vs > 1 PFLOP/s of projection throughput):
for MHA models every layer is restored from hidden states (half the bytes of KV); for GQA models whose KV is smaller
than the hidden state (Llama-2-70B: 4x smaller) every layer is restored from KV. Mixed plans appear when storage is
fast relative to compute (e.g. many sessions restored concurrently, or small T).
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import List, Optional

from ..models.arch import ArchSpec


@dataclass
class LinkModel:
    """Measured rates the planner works from (bytes/s, FLOP/s). Defaults are B200 measurements of this repo."""
    io_bytes_per_s: float = 54e9          # pinned host -> device over PCIe Gen5 x16 (bench: 54 GB/s)
    nvme_bytes_per_s: float = 6e9         # NVMe -> host staging (per drive; raise for RAID / GDS)
    proj_flops_per_s: float = 1.25e15     # restore kernel (tcgen05 K/V projection + RoPE + paged store)
    fwd_flops_per_s: float = 1.0e15       # full-layer forward at prefill sizes
    io_latency_s: float = 30e-6           # per-transfer fixed cost
    tp_size: int = 1                      # hidden states can be fetched 1/tp per rank and all-gathered over NVLink


@dataclass
class RestorePlan:
    kinds: List[str]                      # per layer: "hidden" | "kv" | "none" (recompute)
    io_s: float
    compute_s: float
    layer_fwd_s: float = 0.0              # modelled time to recompute one layer
    layer_hidden_io_s: float = 0.0        # modelled time to transfer one layer's hidden states

    @property
    def interleave(self) -> float:
        """Saved layers whose transfer completes while one prefix layer is recomputed."""
        return self.layer_fwd_s / self.layer_hidden_io_s if self.layer_hidden_io_s > 0 else 0.0

    @property
    def est_seconds(self) -> float:
        return max(self.io_s, self.compute_s)

    def layers_of(self, kind: str) -> List[int]:
        return [i for i, k in enumerate(self.kinds) if k == kind]


def _costs(spec: ArchSpec, n_tokens: int, link: LinkModel, from_nvme: bool, elem: int):
    bw = link.io_bytes_per_s if not from_nvme else min(link.io_bytes_per_s, link.nvme_bytes_per_s)
    tp = max(1, link.tp_size)
    a = n_tokens * spec.hidden * elem / tp / bw + link.io_latency_s            # hidden: 1/tp of the rows per rank
    b = n_tokens * 2 * (spec.kv_dim / tp) * elem / bw + link.io_latency_s      # kv: this rank's heads only
    c = 2.0 * n_tokens * spec.hidden * 2 * (spec.kv_dim / tp) / link.proj_flops_per_s
    inter = spec.moe_intermediate * spec.top_k if spec.is_moe else spec.intermediate
    layer_flops = 2.0 * n_tokens * (spec.hidden * (spec.q_dim + 2 * spec.kv_dim) + spec.q_dim * spec.hidden +
                                    (3 if spec.gated_mlp else 2) * spec.hidden * inter) / tp
    layer_flops += 4.0 * n_tokens * n_tokens / 2 * spec.q_dim / tp             # causal attention
    f = layer_flops / link.fwd_flops_per_s
    return a, b, c, f


def plan_restore(spec: ArchSpec, n_tokens: int, link: Optional[LinkModel] = None, from_nvme: bool = False,
                 elem: int = 2, allow_recompute: bool = True, allow_kv: bool = True,
                 extra_compute_s: float = 0.0) -> RestorePlan:
    """Choose per-layer restore methods minimising ``max(io, compute)``.

    Layout of the plan: the first ``r`` layers are recomputed (optional), then ``n_h`` layers come from hidden
    states, the remaining ``n_k`` from KV pages. All ``(r, n_h)`` splits are enumerated (L is at most a few
    hundred), so the result is optimal for the cost model above.

    ``extra_compute_s``: GPU time that shares the restore window — with deferred restore the new turn's prefill is
    interleaved with the projections, so its duration belongs on the compute side of the balance.
    """
    link = link or LinkModel()
    L = spec.n_layers
    a, b, c, f = _costs(spec, n_tokens, link, from_nvme, elem)
    best = None
    for r in range(0, (L if allow_recompute else 0) + 1):
        for n_h in range(0, L - r + 1):
            n_k = L - r - n_h
            if n_k and not allow_kv:
                continue
            io = n_h * a + n_k * b
            comp = r * f + n_h * c + extra_compute_s
            t = max(io, comp)
            # tie-break: fewer stored bytes (prefer whichever of hidden/kv is smaller), then less compute
            stored = n_h * n_tokens * spec.hidden + n_k * n_tokens * 2 * spec.kv_dim
            key = (round(t, 9), stored, comp)
            if best is None or key < best[0]:
                best = (key, r, n_h, n_k, io, comp)
    _, r, n_h, n_k, io, comp = best
    kinds = ["none"] * r + ["hidden"] * n_h + ["kv"] * n_k
    return RestorePlan(kinds, io, comp, layer_fwd_s=f, layer_hidden_io_s=a)


def storage_bytes(spec: ArchSpec, n_tokens: int, plan: RestorePlan, elem: int = 2) -> int:
    per = {"hidden": spec.hidden * elem, "kv": 2 * spec.kv_dim * elem, "none": 0}
    return sum(per[k] for k in plan.kinds) * n_tokens


def measure_link(device, nbytes: int = 128 << 20, repeats: int = 4, group=None, tp_size: int = 1) -> LinkModel:
    """A :class:`LinkModel` whose host-to-device rate is MEASURED: every rank of ``group`` (default: all ranks) copies a
    pinned buffer at the same time, as restores do in steady state, and the slowest rank's rate is taken. On a node
    whose GPUs share host memory bandwidth the contended rate can be less than half of the single-GPU one, which moves
    the planner's optimum from "transfer everything" to "recompute a prefix" (``profiles/bench_runs.md``)."""
    import torch
    if getattr(device, "type", str(device)) != "cuda" or not torch.cuda.is_available():
        return LinkModel(tp_size=tp_size)
    import torch.distributed as dist
    src = torch.empty(nbytes, dtype=torch.uint8, pin_memory=True)
    dst = torch.empty(nbytes, dtype=torch.uint8, device=device)
    dst.copy_(src, non_blocking=True)
    torch.cuda.synchronize(device)
    multi = dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1
    if multi:
        dist.barrier(group=group)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(repeats):
        dst.copy_(src, non_blocking=True)
    e1.record()
    torch.cuda.synchronize(device)
    bw = torch.tensor([repeats * nbytes / (e0.elapsed_time(e1) / 1e3)], device=device)
    if multi:
        dist.all_reduce(bw, op=dist.ReduceOp.MIN, group=group)
    return LinkModel(io_bytes_per_s=float(bw.item()), tp_size=tp_size)
