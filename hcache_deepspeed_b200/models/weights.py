"""Model weights in the layout the sm_100a kernels consume, and how to get there.

Layouts (all ``[out_features, in_features]`` row-major, i.e. the HF / DeepSpeed orientation):

* ``w_qkv``  rows = local Q heads, local K heads, local V heads (``qkv_parameters.py:29-45`` order);
             ``w_kv`` — the operand of the HCache restore kernel — is a zero-copy row slice of it.
* ``w_up``   gated MLPs interleave rows gate_0, up_0, gate_1, up_1, ... so one GEMM epilogue can
             compute ``act(gate) * up`` from adjacent accumulator columns (same convention as the
             reference's ``GatedMLPParameter``, ``common_parameters/mlp_parameters.py:43-66``).
* ``w_o`` / ``w_down`` are column (input-dim) shards under tensor parallelism; their biases live on
  TP rank 0 only.

Checkpoints arrive as ``(hf_name, tensor)`` pairs from a checkpoint engine; a per-family mapper turns
each into canonical pieces, pieces are sharded immediately (host memory stays bounded) and a layer
is finalised once all of its pieces are present.
"""
from __future__ import annotations

import re
from dataclasses import dataclass, field
from typing import Callable, Dict, Iterable, List, Optional, Tuple

import torch

from ..parallel import sharding as sh
from .arch import ArchSpec


@dataclass
class ExpertWeights:
    w_up: torch.Tensor
    w_down: torch.Tensor


@dataclass
class LayerWeights:
    norm1_g: Optional[torch.Tensor] = None
    norm1_b: Optional[torch.Tensor] = None
    norm2_g: Optional[torch.Tensor] = None
    norm2_b: Optional[torch.Tensor] = None
    w_qkv: Optional[torch.Tensor] = None
    b_qkv: Optional[torch.Tensor] = None
    w_o: Optional[torch.Tensor] = None
    b_o: Optional[torch.Tensor] = None
    w_up: Optional[torch.Tensor] = None
    b_up: Optional[torch.Tensor] = None
    w_down: Optional[torch.Tensor] = None
    b_down: Optional[torch.Tensor] = None
    # mixture of experts
    w_router: Optional[torch.Tensor] = None
    experts: List[ExpertWeights] = field(default_factory=list)
    shared_expert: Optional[ExpertWeights] = None
    w_shared_gate: Optional[torch.Tensor] = None

    def tensors(self) -> Dict[str, torch.Tensor]:
        out = {}
        for k, v in self.__dict__.items():
            if isinstance(v, torch.Tensor):
                out[k] = v
        for i, e in enumerate(self.experts):
            out[f"experts.{i}.w_up"] = e.w_up
            out[f"experts.{i}.w_down"] = e.w_down
        if self.shared_expert is not None:
            out["shared_expert.w_up"] = self.shared_expert.w_up
            out["shared_expert.w_down"] = self.shared_expert.w_down
        return out


@dataclass
class ModelWeights:
    embed: Optional[torch.Tensor] = None
    pos_embed: Optional[torch.Tensor] = None
    embed_norm_g: Optional[torch.Tensor] = None     # LayerNorm applied to the token embeddings (Bloom)
    embed_norm_b: Optional[torch.Tensor] = None
    final_norm_g: Optional[torch.Tensor] = None
    final_norm_b: Optional[torch.Tensor] = None
    lm_head: Optional[torch.Tensor] = None
    lm_head_b: Optional[torch.Tensor] = None
    layers: List[LayerWeights] = field(default_factory=list)

    def named_tensors(self) -> Dict[str, torch.Tensor]:
        out = {}
        for k in ("embed", "pos_embed", "embed_norm_g", "embed_norm_b", "final_norm_g", "final_norm_b", "lm_head", "lm_head_b"):
            v = getattr(self, k)
            if v is not None:
                out[k] = v
        for i, layer in enumerate(self.layers):
            for k, v in layer.tensors().items():
                out[f"layers.{i}.{k}"] = v
        return out

    def n_bytes(self) -> int:
        seen, total = set(), 0
        for t in self.named_tensors().values():
            key = (t.untyped_storage().data_ptr(), t.storage_offset(), t.numel())
            if key not in seen:
                seen.add(key)
                total += t.numel() * t.element_size()
        return total


# ---------------------------------------------------------------------------------------------------
# tensor-parallel geometry of one rank
# ---------------------------------------------------------------------------------------------------
@dataclass
class ShardInfo:
    tp_rank: int
    tp_size: int
    q_heads: List[int]
    kv_heads: List[int]
    inter_range: Tuple[int, int]
    moe_inter_range: Tuple[int, int]
    shared_inter_range: Tuple[int, int]
    vocab_range: Tuple[int, int]

    @property
    def n_q(self) -> int:
        return len(self.q_heads)

    @property
    def n_kv(self) -> int:
        return len(self.kv_heads)

    @property
    def inter(self) -> int:
        return self.inter_range[1] - self.inter_range[0]

    @property
    def vocab(self) -> int:
        return self.vocab_range[1] - self.vocab_range[0]


def shard_info(spec: ArchSpec, tp_rank: int = 0, tp_size: int = 1) -> ShardInfo:
    q_heads, kv_heads = sh.local_head_indices(tp_rank, tp_size, spec.n_heads, spec.n_kv_heads)
    gran = sh.DEFAULT_SHARD_GRANULARITY

    def rng(n):
        if n == 0:
            return (0, 0)
        g = gran if n % gran == 0 else 1
        return sh.get_shard_endpoints(n, tp_rank, tp_size, granularity=g)

    return ShardInfo(tp_rank, tp_size, q_heads, kv_heads, rng(spec.intermediate), rng(spec.moe_intermediate),
                     rng(spec.shared_expert_intermediate),
                     sh.get_shard_endpoints(spec.vocab, tp_rank, tp_size, granularity=1))


def _head_rows(heads: List[int], d: int) -> torch.Tensor:
    return (torch.tensor(heads, dtype=torch.long)[:, None] * d + torch.arange(d)[None, :]).reshape(-1)


def interleave_gate_up(gate: torch.Tensor, up: torch.Tensor) -> torch.Tensor:
    """rows gate_0, up_0, gate_1, up_1, ...  (works for 2-D weights and 1-D biases)."""
    return torch.stack([gate, up], dim=1).reshape((gate.shape[0] * 2, ) + tuple(gate.shape[1:])).contiguous()


# ---------------------------------------------------------------------------------------------------
# canonical-piece assembler
# ---------------------------------------------------------------------------------------------------
class WeightAssembler:
    """Consumes canonical pieces (see module docstring of :mod:`families`), shards them for one TP rank and
    emits :class:`ModelWeights` on ``device`` in ``dtype``."""

    def __init__(self, spec: ArchSpec, tp_rank: int, tp_size: int, dtype: torch.dtype, device: torch.device):
        self.spec, self.dtype, self.device = spec, dtype, device
        self.info = shard_info(spec, tp_rank, tp_size)
        self.top: Dict[str, torch.Tensor] = {}
        self.layers: Dict[int, Dict[str, torch.Tensor]] = {}
        self.out = ModelWeights(layers=[LayerWeights() for _ in range(spec.n_layers)])
        self._done_layers = set()

    # -- sharding of one piece ------------------------------------------------------------------------
    def _shard(self, name: str, t: torch.Tensor) -> Optional[torch.Tensor]:
        s, info, d = self.spec, self.info, self.spec.head_dim
        r0 = info.tp_rank == 0
        base = name.split(".")[-1] if not name.startswith("moe.") else name
        if name in ("attn.q", "attn.q_b"):
            return t[_head_rows(info.q_heads, d)]
        if name in ("attn.k", "attn.v", "attn.k_b", "attn.v_b"):
            return t[_head_rows(info.kv_heads, d)]
        if name == "attn.o":
            return t[:, _head_rows(info.q_heads, d)]
        if name in ("attn.o_b", "mlp.down_b"):
            return t if r0 else None
        if name in ("mlp.gate", "mlp.up", "mlp.gate_b", "mlp.up_b"):
            a, b = info.inter_range
            return t[a:b]
        if name == "mlp.down":
            a, b = info.inter_range
            return t[:, a:b]
        m = re.match(r"moe\.experts\.(\d+)\.(gate|up|down)$", name)
        if m:
            a, b = info.moe_inter_range
            return t[:, a:b] if m.group(2) == "down" else t[a:b]
        m = re.match(r"moe\.shared\.(gate|up|down)$", name)
        if m:
            a, b = info.shared_inter_range
            return t[:, a:b] if m.group(1) == "down" else t[a:b]
        if name in ("lm_head", "lm_head_b"):
            a, b = info.vocab_range
            return t[a:b]
        del base
        return t  # norms, router, embeddings: replicated

    def add(self, name: str, tensor: torch.Tensor, layer: Optional[int] = None) -> None:
        piece = self._shard(name, tensor)
        if piece is None:
            piece = torch.empty(0)
        piece = piece.detach()
        if layer is None:
            self.top[name] = piece
        else:
            self.layers.setdefault(layer, {})[name] = piece
            self._maybe_finalize_layer(layer)

    # -- completion tracking ----------------------------------------------------------------------------
    def _required(self) -> List[str]:
        s = self.spec
        req = ["attn.q", "attn.k", "attn.v", "attn.o", "norm1.g"]
        if s.qkv_bias:
            req += ["attn.q_b", "attn.k_b", "attn.v_b"]
        if s.attn_out_bias:
            req += ["attn.o_b"]
        if s.norm_bias:
            req += ["norm1.b"]
        two_norms = (not s.parallel_residual) or s.parallel_two_norms
        if two_norms:
            req += ["norm2.g"] + (["norm2.b"] if s.norm_bias else [])
        if s.is_moe:
            req += ["moe.router"]
            for e in range(s.n_experts):
                req += [f"moe.experts.{e}.up", f"moe.experts.{e}.down"] + ([f"moe.experts.{e}.gate"] if s.gated_mlp else [])
            if s.shared_expert_intermediate:
                req += ["moe.shared.gate", "moe.shared.up", "moe.shared.down", "moe.shared_gate"]
        else:
            req += ["mlp.up", "mlp.down"] + (["mlp.gate"] if s.gated_mlp else [])
            if s.mlp_bias:
                req += ["mlp.up_b", "mlp.down_b"] + (["mlp.gate_b"] if s.gated_mlp else [])
        return req

    def _to(self, t: Optional[torch.Tensor]) -> Optional[torch.Tensor]:
        if t is None or t.numel() == 0:
            return None
        return t.to(device=self.device, dtype=self.dtype).contiguous()

    def _maybe_finalize_layer(self, li: int) -> None:
        if li in self._done_layers:
            return
        p = self.layers[li]
        if any(r not in p for r in self._required()):
            return
        s = self.spec
        lw = self.out.layers[li]
        lw.norm1_g, lw.norm1_b = self._to(p["norm1.g"]), self._to(p.get("norm1.b"))
        lw.norm2_g, lw.norm2_b = self._to(p.get("norm2.g")), self._to(p.get("norm2.b"))
        lw.w_qkv = self._to(torch.cat([p["attn.q"], p["attn.k"], p["attn.v"]], dim=0))
        if s.qkv_bias:
            lw.b_qkv = self._to(torch.cat([p["attn.q_b"], p["attn.k_b"], p["attn.v_b"]], dim=0))
        lw.w_o, lw.b_o = self._to(p["attn.o"]), self._to(p.get("attn.o_b"))

        def up_of(prefix: str) -> torch.Tensor:
            if s.gated_mlp:
                return interleave_gate_up(p[f"{prefix}gate"], p[f"{prefix}up"])
            return p[f"{prefix}up"]

        if s.is_moe:
            lw.w_router = self._to(p["moe.router"])
            lw.experts = [ExpertWeights(self._to(up_of(f"moe.experts.{e}.")), self._to(p[f"moe.experts.{e}.down"]))
                          for e in range(s.n_experts)]
            if s.shared_expert_intermediate:
                lw.shared_expert = ExpertWeights(self._to(up_of("moe.shared.")), self._to(p["moe.shared.down"]))
                lw.w_shared_gate = self._to(p["moe.shared_gate"])
        else:
            lw.w_up = self._to(up_of("mlp."))
            lw.w_down = self._to(p["mlp.down"])
            if s.mlp_bias:
                lw.b_up = self._to(interleave_gate_up(p["mlp.gate_b"], p["mlp.up_b"]) if s.gated_mlp else p["mlp.up_b"])
                lw.b_down = self._to(p.get("mlp.down_b"))
        self._done_layers.add(li)
        del self.layers[li]

    def finish(self) -> ModelWeights:
        s = self.spec
        missing = [i for i in range(s.n_layers) if i not in self._done_layers]
        if missing:
            have = {i: sorted(self.layers.get(i, {})) for i in missing[:2]}
            raise ValueError(f"checkpoint is missing parameters for layers {missing[:8]}; got {have}; "
                             f"need {self._required()}")
        if "embed" not in self.top:
            raise ValueError("checkpoint has no token embedding")
        o = self.out
        o.embed = self._to(self.top["embed"])
        o.pos_embed = self._to(self.top.get("pos_embed"))
        o.final_norm_g, o.final_norm_b = self._to(self.top.get("final_norm.g")), self._to(self.top.get("final_norm.b"))
        o.embed_norm_g, o.embed_norm_b = self._to(self.top.get("embed_norm.g")), self._to(self.top.get("embed_norm.b"))
        if self.spec.embed_norm and o.embed_norm_g is None:
            raise ValueError("checkpoint has no embedding LayerNorm but the architecture needs one")
        if "lm_head" in self.top:
            o.lm_head = self._to(self.top["lm_head"])
        else:  # tied embeddings: the head is this rank's vocabulary slice of the embedding
            a, b = self.info.vocab_range
            o.lm_head = o.embed[a:b] if self.info.tp_size > 1 else o.embed
        o.lm_head_b = self._to(self.top.get("lm_head_b"))
        return o


def load_weights(spec: ArchSpec, named_params: Iterable[Tuple[str, torch.Tensor]],
                 mapper: Callable[[str, torch.Tensor, ArchSpec], List[Tuple[Optional[int], str, torch.Tensor]]],
                 tp_rank: int = 0, tp_size: int = 1, dtype: torch.dtype = torch.bfloat16,
                 device: Optional[torch.device] = None) -> ModelWeights:
    """Stream ``(hf_name, tensor)`` pairs through a family mapper into kernel-ready weights."""
    device = device if device is not None else torch.device("cpu")
    asm = WeightAssembler(spec, tp_rank, tp_size, dtype, device)
    for name, tensor in named_params:
        for layer, cname, piece in mapper(name, tensor, spec):
            asm.add(cname, piece, layer)
    return asm.finish()


def random_weights(spec: ArchSpec, tp_rank: int = 0, tp_size: int = 1, dtype: torch.dtype = torch.bfloat16,
                   device: Optional[torch.device] = None, seed: int = 0, std: float = 0.02) -> ModelWeights:
    """Random-init weights of the right shapes, generated directly on ``device`` in kernel layout (benchmarks
    have no checkpoints). Replicated tensors use a rank-independent seed so TP ranks agree on them."""
    device = device if device is not None else torch.device("cpu")
    info = shard_info(spec, tp_rank, tp_size)
    g_shared = torch.Generator(device=device).manual_seed(seed)
    g_rank = torch.Generator(device=device).manual_seed(seed * 1000003 + 17 + tp_rank)
    H, d = spec.hidden, spec.head_dim

    def rnd(shape, gen, scale=std):
        return (torch.randn(shape, generator=gen, device=device, dtype=torch.float32) * scale).to(dtype)

    def ones(n):
        return (1.0 + 0.05 * torch.randn(n, generator=g_shared, device=device)).to(dtype)

    mw = ModelWeights()
    mw.embed = rnd((spec.vocab, H), g_shared)
    if spec.pos_emb == "learned":
        mw.pos_embed = rnd((spec.max_positions + spec.learned_pos_offset, H), g_shared)
    if spec.embed_norm:
        mw.embed_norm_g, mw.embed_norm_b = ones(H), rnd((H, ), g_shared)
    if spec.final_norm:
        mw.final_norm_g = ones(H)
        mw.final_norm_b = rnd((H, ), g_shared) if spec.norm_bias else None
    a, b = info.vocab_range
    mw.lm_head = (mw.embed[a:b] if tp_size > 1 else mw.embed) if spec.tie_word_embeddings else rnd((b - a, H), g_rank)
    mw.lm_head_b = rnd((b - a, ), g_rank) if spec.lm_head_bias else None
    r0 = tp_rank == 0
    two_norms = (not spec.parallel_residual) or spec.parallel_two_norms
    out_scale = std / max(1.0, (2.0 * spec.n_layers) ** 0.5)
    for _ in range(spec.n_layers):
        lw = LayerWeights()
        lw.norm1_g = ones(H)
        lw.norm1_b = rnd((H, ), g_shared) if spec.norm_bias else None
        if two_norms:
            lw.norm2_g = ones(H)
            lw.norm2_b = rnd((H, ), g_shared) if spec.norm_bias else None
        n_rows = (info.n_q + 2 * info.n_kv) * d
        lw.w_qkv = rnd((n_rows, H), g_rank)
        lw.b_qkv = rnd((n_rows, ), g_rank) if spec.qkv_bias else None
        lw.w_o = rnd((H, info.n_q * d), g_rank, out_scale)
        lw.b_o = rnd((H, ), g_shared) if (spec.attn_out_bias and r0) else None
        mult = 2 if spec.gated_mlp else 1
        if spec.is_moe:
            lw.w_router = rnd((spec.n_experts, H), g_shared)
            mi = info.moe_inter_range[1] - info.moe_inter_range[0]
            lw.experts = [ExpertWeights(rnd((mult * mi, H), g_rank), rnd((H, mi), g_rank, out_scale))
                          for _ in range(spec.n_experts)]
            if spec.shared_expert_intermediate:
                si = info.shared_inter_range[1] - info.shared_inter_range[0]
                lw.shared_expert = ExpertWeights(rnd((mult * si, H), g_rank), rnd((H, si), g_rank, out_scale))
                lw.w_shared_gate = rnd((1, H), g_shared)
        else:
            lw.w_up = rnd((mult * info.inter, H), g_rank)
            lw.w_down = rnd((H, info.inter), g_rank, out_scale)
            if spec.mlp_bias:
                lw.b_up = rnd((mult * info.inter, ), g_rank)
                lw.b_down = rnd((H, ), g_shared) if r0 else None
        mw.layers.append(lw)
    return mw
