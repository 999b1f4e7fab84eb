"""GShard-style sharded mixture of experts: top-1 / top-2 / top-k gating with capacity, auxiliary load-balancing
loss, and the dispatch -> all-to-all -> experts -> all-to-all -> combine layer.

Dispatch and combine are dense einsum contractions with a [tokens, experts, capacity] mask, which keeps the layer
differentiable end to end; token exchange between expert-parallel ranks is ``all_to_all_single`` (NCCL over NVSwitch
on GPUs). Reference: ``deepspeed/moe/sharded_moe.py`` (_AllToAll :96, top1gating :183, top2gating :290,
topkgating :374, TopKGate :449, MOELayer :533).
"""
from __future__ import annotations

import math
import time
from typing import Callable, Optional, Tuple

import torch
import torch.nn.functional as F
from torch import Tensor

from .. import comm as dist
from ..utils import groups
from .mappings import drop_tokens, gather_tokens

uniform_map, gumbel_map, exp_selection_uniform_map = {}, {}, {}


def multiplicative_jitter(x: Tensor, device, epsilon: float = 1e-2) -> Tensor:
    if epsilon == 0:
        return x
    u = uniform_map.get(device)
    if u is None:
        u = torch.distributions.uniform.Uniform(low=torch.tensor(1.0 - epsilon, device=device),
                                                high=torch.tensor(1.0 + epsilon, device=device)).rsample
        uniform_map[device] = u
    return x * u(x.shape)


def gumbel_rsample(shape, device) -> Tensor:
    g = gumbel_map.get(device)
    if g is None:
        g = torch.distributions.gumbel.Gumbel(torch.tensor(0.0, device=device), torch.tensor(1.0, device=device)).rsample
        gumbel_map[device] = g
    return g(shape)


class _AllToAll(torch.autograd.Function):
    @staticmethod
    def forward(ctx, group, x: Tensor) -> Tensor:
        ctx.group = group
        x = x.contiguous()
        out = torch.empty_like(x)
        dist.all_to_all_single(out, x, group=group)
        return out

    @staticmethod
    def backward(ctx, grad: Tensor):
        return None, _AllToAll.apply(ctx.group, grad)


def _capacity(gates: Tensor, capacity_factor: float, min_capacity: int) -> int:
    num_tokens, num_experts = gates.shape[0], gates.shape[1]
    cap = math.ceil((num_tokens / num_experts) * capacity_factor)
    return max(cap, min_capacity)


def _one_hot(idx: Tensor, n: int) -> Tensor:
    return F.one_hot(idx, num_classes=n)


def _keep_within_capacity(mask: Tensor, capacity: int, prior: Optional[Tensor] = None) -> Tuple[Tensor, Tensor]:
    """Position of each token inside its expert (exclusive cumsum over tokens), dropping tokens past ``capacity``.
    ``prior``: per-expert counts already taken by higher-priority choices."""
    loc = torch.cumsum(mask, dim=0) - 1
    if prior is not None:
        loc = loc + prior
    keep = mask * (loc < capacity)
    return keep, loc


def top1gating(logits: Tensor, capacity_factor: float, min_capacity: int, used_token: Optional[Tensor] = None,
               noisy_gate_policy: Optional[str] = None, drop_tokens: bool = True, use_rts: bool = True, ep_group=None,
               use_tutel: bool = False):
    """Returns (l_aux, combine_weights [S,E,C], dispatch_mask [S,E,C] bool, exp_counts [E])."""
    if noisy_gate_policy == "RSample":
        logits_w_noise = logits + gumbel_rsample(logits.shape, device=logits.device)
    gates = F.softmax(logits, dim=1)
    capacity = _capacity(gates, capacity_factor, min_capacity)
    idx = torch.argmax(logits_w_noise if noisy_gate_policy == "RSample" else gates, dim=1)
    E = gates.shape[1]
    mask1 = _one_hot(idx, E)
    if used_token is not None:
        mask1 = torch.einsum("s,se->se", used_token, mask1)
    exp_counts = torch.sum(mask1, dim=0).detach()
    if not drop_tokens:
        new_cap = torch.max(exp_counts)
        if ep_group is not None and dist.is_initialized():
            dist.all_reduce(new_cap, op=dist.ReduceOp.MAX, group=ep_group)
        capacity = max(int(new_cap), min_capacity)
        tp = groups._get_model_parallel_world_size() if groups.mpu is not None else 1
        capacity = math.ceil(capacity / tp) * tp
    me = torch.mean(gates, dim=0)
    ce = torch.mean(mask1.float(), dim=0)
    l_aux = torch.sum(me * ce) * E
    if use_rts:
        # random token selection: keep a random subset (not the first arrivals) when an expert overflows
        u = exp_selection_uniform_map.get(logits.device)
        if u is None:
            u = torch.distributions.uniform.Uniform(low=torch.tensor(0.0, device=logits.device),
                                                    high=torch.tensor(1.0, device=logits.device)).rsample
            exp_selection_uniform_map[logits.device] = u
        mask1_rand = mask1 * u(mask1.shape)
    else:
        mask1_rand = mask1
    assert logits.shape[0] >= min_capacity, "No. of tokens (batch-size) should be greater than min_capacity."
    top_idx = torch.topk(mask1_rand, k=min(capacity, mask1_rand.shape[0]), dim=0)[1]
    mask1 = mask1 * torch.zeros_like(mask1).scatter_(0, top_idx, 1)
    locations1 = torch.cumsum(mask1, dim=0) - 1
    locations1_s = torch.sum(locations1 * mask1, dim=1)
    gates = gates * mask1.float()
    locations1_sc = _one_hot(locations1_s, capacity).float()
    combine = torch.einsum("se,sc->sec", gates, locations1_sc)
    return l_aux, combine, combine.bool(), exp_counts


def top2gating(logits: Tensor, capacity_factor: float, min_capacity: int, drop_tokens: bool = True, ep_group=None,
               top2_2nd_expert_sampling: bool = True):
    gates = F.softmax(logits, dim=1)
    E = gates.shape[1]
    idx1 = torch.argmax(gates, dim=1)
    mask1 = _one_hot(idx1, E)
    noisy = logits + gumbel_rsample(logits.shape, device=logits.device) if top2_2nd_expert_sampling else logits
    idx2 = torch.argmax(noisy.masked_fill(mask1.bool(), float("-inf")), dim=1)
    mask2 = _one_hot(idx2, E)
    loc1 = torch.cumsum(mask1, dim=0) - 1
    loc2 = torch.cumsum(mask2, dim=0) - 1 + torch.sum(mask1, dim=0, keepdim=True)
    exp_counts = torch.sum(mask1 + mask2, dim=0).detach()
    me = torch.mean(gates, dim=0)
    ce = torch.mean(mask1.float(), dim=0)
    l_aux = torch.mean(me * ce) * E * E
    if drop_tokens:
        capacity = _capacity(gates, capacity_factor * 2, min_capacity)
        mask1 = mask1 * (loc1 < capacity)
        mask2 = mask2 * (loc2 < capacity)
    else:
        new_cap = torch.max(exp_counts)
        if ep_group is not None and dist.is_initialized():
            dist.all_reduce(new_cap, op=dist.ReduceOp.MAX, group=ep_group)
        capacity = max(int(new_cap), min_capacity)
        tp = groups._get_model_parallel_world_size() if groups.mpu is not None else 1
        capacity = math.ceil(capacity / tp) * tp
    loc1_s = torch.sum(loc1 * mask1, dim=1)
    loc2_s = torch.sum(loc2 * mask2, dim=1)
    g1 = torch.einsum("se,se->s", gates, mask1.float())
    g2 = torch.einsum("se,se->s", gates, mask2.float())
    denom = torch.clamp(g1 + g2, min=torch.finfo(g1.dtype).eps)
    g1, g2 = g1 / denom, g2 / denom
    gates1 = torch.einsum("s,se->se", g1, mask1.float())
    gates2 = torch.einsum("s,se->se", g2, mask2.float())
    combine = torch.einsum("se,sc->sec", gates1, _one_hot(loc1_s, capacity).float()) + \
        torch.einsum("se,sc->sec", gates2, _one_hot(loc2_s, capacity).float())
    return l_aux, combine, combine.bool(), exp_counts


def topkgating(logits: Tensor, k: int, capacity_factor: float, min_capacity: int, drop_tokens: bool = True, ep_group=None,
               drop_policy: str = "probs"):
    gates = F.softmax(logits, dim=1)
    S, E = gates.shape
    topv, topi = torch.topk(gates, k=k, dim=1)
    mask = torch.zeros_like(gates, dtype=torch.bool).scatter_(1, topi, 1)
    exp_counts = torch.sum(mask, dim=0).detach()
    me = torch.mean(gates, dim=0)
    ce = torch.mean(mask.float(), dim=0)
    l_aux = torch.mean(me * ce) * E * E / k
    if drop_tokens:
        capacity = _capacity(gates, capacity_factor * k, min_capacity)
        if drop_policy == "probs":
            keep_idx = torch.topk(gates.masked_fill(~mask, -1.0), k=min(capacity, S), dim=0)[1]
            cap_mask = torch.zeros_like(mask).scatter_(0, keep_idx, 1)
            mask = mask & cap_mask
            loc = torch.cumsum(mask.int(), dim=0) - 1
        elif drop_policy == "position":
            loc = torch.cumsum(mask.int(), dim=0) - 1
            mask = mask & (loc < capacity)
        else:
            raise ValueError(f"Invalid drop_policy: {drop_policy}")
    else:
        new_cap = torch.max(exp_counts)
        if ep_group is not None and dist.is_initialized():
            dist.all_reduce(new_cap, op=dist.ReduceOp.MAX, group=ep_group)
        capacity = max(int(new_cap), min_capacity)
        loc = torch.cumsum(mask.int(), dim=0) - 1
    g = gates * mask.float()
    g = g / torch.clamp(g.sum(dim=1, keepdim=True), min=torch.finfo(g.dtype).eps)
    loc = (loc * mask.int()).clamp(max=capacity - 1)
    loc_oh = F.one_hot(loc, num_classes=capacity).float() * mask.unsqueeze(-1).float()
    combine = g.unsqueeze(-1) * loc_oh
    return l_aux, combine, combine.bool(), exp_counts


class TopKGate(torch.nn.Module):
    """Linear router (fp32) + the gating function for k = 1, 2 or k > 2."""

    def __init__(self, model_dim: int, num_experts: int, k: int = 1, capacity_factor: float = 1.0,
                 eval_capacity_factor: float = 1.0, min_capacity: int = 8, noisy_gate_policy: Optional[str] = None,
                 drop_tokens: bool = True, use_rts: bool = True, ep_group=None, top2_2nd_expert_sampling: bool = True):
        super().__init__()
        self.wg = torch.nn.Linear(model_dim, num_experts, bias=False)
        self.ep_group = ep_group
        self.k, self.capacity_factor, self.eval_capacity_factor = k, capacity_factor, eval_capacity_factor
        self.min_capacity, self.noisy_gate_policy = min_capacity, noisy_gate_policy
        self.drop_tokens, self.use_rts, self.top2_2nd_expert_sampling = drop_tokens, use_rts, top2_2nd_expert_sampling
        self.gate_time = 0.0
        self.wall_clock_breakdown = False

    def _set_ep_group(self, ep_group):
        assert self.ep_group is None, "Attempting to override an existing ep_group"
        self.ep_group = ep_group

    def forward(self, x: Tensor, used_token: Optional[Tensor] = None, use_tutel: bool = False):
        xf = x.float()
        if self.noisy_gate_policy == "Jitter" and self.training:
            xf = multiplicative_jitter(xf, device=x.device)
        logits = F.linear(xf, self.wg.weight.float())
        cf = self.capacity_factor if self.training else self.eval_capacity_factor
        if self.k == 1:
            return top1gating(logits, cf, self.min_capacity, used_token, self.noisy_gate_policy if self.training else None,
                              self.drop_tokens, self.use_rts, self.ep_group, use_tutel)
        if self.k == 2:
            return top2gating(logits, cf, self.min_capacity, self.drop_tokens, self.ep_group, self.top2_2nd_expert_sampling)
        return topkgating(logits, self.k, cf, self.min_capacity, self.drop_tokens, self.ep_group)


class MOELayer(torch.nn.Module):
    """dispatch -> all-to-all -> local experts -> all-to-all -> combine."""

    def __init__(self, gate: torch.nn.Module, experts: torch.nn.Module, ep_group_name: str, ep_size: int,
                 num_local_experts: int, use_tutel: bool = False):
        super().__init__()
        self.gate, self.experts = gate, experts
        self.ep_group = None
        self.ep_size, self.ep_group_name, self.num_local_experts = ep_size, ep_group_name, num_local_experts
        self.l_aux, self.exp_counts = None, None
        self.wall_clock_breakdown = False
        self.gate_time = self.time_falltoall = self.time_salltoall = self.time_moe = 0.0   # ms, of the last forward

    def _set_ep_group(self, ep_group):
        self.ep_group = ep_group
        self.gate._set_ep_group(ep_group)

    def _now(self, x: Tensor) -> float:
        """Wall clock in ms for the ``wall_clock_breakdown`` counters (the device is drained first so spans are real)."""
        if x.is_cuda:
            torch.cuda.synchronize(x.device)
        return time.perf_counter() * 1e3

    def forward(self, *inputs: Tensor, **kwargs) -> Tensor:
        x = inputs[0]
        timed = self.wall_clock_breakdown
        d_model = x.shape[-1]
        tokens = x.reshape(-1, d_model)
        t0 = self._now(x) if timed else 0.0
        self.l_aux, combine, dispatch, self.exp_counts = self.gate(tokens, inputs[1] if len(inputs) > 1 else None)
        if timed:
            t1 = self._now(x)
            self.gate_time = t1 - t0
        dispatched = torch.einsum("sec,sm->ecm", dispatch.type_as(x), tokens)
        tp = groups._get_model_parallel_world_size() if groups.mpu is not None else 1
        if tp > 1:
            dispatched = drop_tokens(dispatched, dim=1)
        ta = self._now(x) if timed else 0.0
        if self.ep_size > 1:
            dispatched = _AllToAll.apply(self.ep_group, dispatched)
        if timed:
            self.time_falltoall = self._now(x) - ta
        dispatched = dispatched.reshape(self.ep_size, self.num_local_experts, -1, d_model)
        out = self.experts(dispatched)
        out = out.reshape(self.ep_size * self.num_local_experts, -1, d_model)
        ta = self._now(x) if timed else 0.0
        if self.ep_size > 1:
            out = _AllToAll.apply(self.ep_group, out)
        if timed:
            self.time_salltoall = self._now(x) - ta
        if tp > 1:
            out = gather_tokens(out, dim=1)
        combined = torch.einsum("sec,ecm->sm", combine.type_as(x), out)
        if timed:
            self.time_moe = self._now(x) - t0
        return combined.reshape(x.shape)


def einsum(rule: str, a: Tensor, b: Tensor) -> Tensor:
    """The handful of contractions the MoE layer uses, as explicit matmuls / broadcasts where that is faster than
    ``torch.einsum`` (dispatch / combine are batched matmuls); anything else falls through to ``torch.einsum``."""
    if rule == "s,se->se":
        return a.unsqueeze(1) * b
    if rule == "se,sc->sec":
        return a.unsqueeze(2) * b.unsqueeze(1)
    if rule == "se,se->s":
        return (a * b).sum(dim=1)
    if rule == "sec,sm->ecm":
        s = a.shape[0]
        return torch.matmul(a.reshape(s, -1).t(), b).reshape(a.shape[1], a.shape[2], b.shape[1])
    if rule == "sec,ecm->sm":
        return torch.matmul(a.reshape(a.shape[0], -1), b.reshape(-1, b.shape[-1]))
    return torch.einsum(rule, a, b)
