"""HuggingFace checkpoint name -> canonical weight pieces, one mapper per model family.

Canonical piece names (per layer unless noted):
  ``attn.q|k|v|o`` (+ ``_b`` biases), ``norm1.g|b``, ``norm2.g|b``,
  ``mlp.gate|up|down`` (+ ``_b``), ``moe.router``, ``moe.experts.{e}.gate|up|down``,
  ``moe.shared.gate|up|down``, ``moe.shared_gate``;
  top level: ``embed``, ``pos_embed``, ``final_norm.g|b``, ``lm_head``, ``lm_head_b``.

A mapper returns ``[(layer_index_or_None, canonical_name, tensor), ...]`` (empty list = ignore).
Replaces the ``PARAM_MAPPING`` tables of the reference's per-family containers
(e.g. ``deepspeed/inference/v2/model_implementations/llama_v2/container.py``,
``opt/container.py``, ``falcon/container.py``, ``mixtral/container.py``).
"""
from __future__ import annotations

import re
from typing import Callable, Dict, List, Optional, Tuple

import torch

from .arch import ArchSpec

Piece = Tuple[Optional[int], str, torch.Tensor]
Mapper = Callable[[str, torch.Tensor, ArchSpec], List[Piece]]

MAPPERS: Dict[str, Mapper] = {}


def register_mapper(*model_types: str):

    def deco(fn):
        for mt in model_types:
            MAPPERS[mt] = fn
        return fn

    return deco


def get_mapper(model_type: str) -> Mapper:
    if model_type not in MAPPERS:
        raise ValueError(f"no checkpoint mapper for model type {model_type}")
    return MAPPERS[model_type]


_LAYER_RE = re.compile(r"(?:^|\.)(?:layers|h)\.(\d+)\.(.*)$")


def _split_layer(name: str):
    m = _LAYER_RE.search(name)
    if not m:
        return None, name
    return int(m.group(1)), m.group(2)


def _wb(suffix: str) -> str:
    return "_b" if suffix == "bias" else ""


# ---------------------------------------------------------------------------------------------------
@register_mapper("llama", "mistral", "qwen2", "internlm", "mixtral", "qwen2_moe", "phi3")
def map_llama_like(name: str, t: torch.Tensor, spec: ArchSpec) -> List[Piece]:
    layer, rest = _split_layer(name)
    if layer is None:
        if name.endswith("embed_tokens.weight"):
            return [(None, "embed", t)]
        if name in ("model.norm.weight", "norm.weight"):
            return [(None, "final_norm.g", t)]
        if name == "lm_head.weight":
            return [] if spec.tie_word_embeddings else [(None, "lm_head", t)]
        return []
    parts = rest.split(".")
    kind = parts[-1]
    if kind.startswith("lora_"):
        return []   # adapter tensors of LoRA-wrapped projections (folded into the base weight by the hybrid engine)
    if rest.startswith("input_layernorm."):
        return [(layer, "norm1.g", t)]
    if rest.startswith("post_attention_layernorm."):
        return [(layer, "norm2.g", t)]
    if rest.startswith("self_attn."):
        proj = parts[1]
        if proj in ("q_proj", "k_proj", "v_proj"):
            return [(layer, f"attn.{proj[0]}{_wb(kind)}", t)]
        if proj == "o_proj":
            return [(layer, f"attn.o{_wb(kind)}", t)]
        if proj == "qkv_proj":  # phi3: fused rows [Q; K; V]
            q, k, v = t.split([spec.q_dim, spec.kv_dim, spec.kv_dim], dim=0)
            sfx = _wb(kind)
            return [(layer, f"attn.q{sfx}", q), (layer, f"attn.k{sfx}", k), (layer, f"attn.v{sfx}", v)]
        return []
    if rest.startswith("mlp.") or rest.startswith("block_sparse_moe."):
        sub = parts[1]
        if sub in ("gate_proj", "up_proj", "down_proj"):
            return [(layer, f"mlp.{sub.split('_')[0]}{_wb(kind)}", t)]
        if sub == "gate_up_proj":  # phi3: fused rows [gate; up]
            g, u = t.chunk(2, dim=0)
            return [(layer, "mlp.gate", g), (layer, "mlp.up", u)]
        if sub == "gate" and kind == "weight":  # MoE router
            return [(layer, "moe.router", t)]
        if sub == "experts" and len(parts) == 3 and t.dim() == 3:
            # transformers >= 5 keeps experts fused: gate_up_proj [E, 2I, H] (rows [gate; up]), down_proj [E, H, I]
            out: List[Piece] = []
            for e in range(t.shape[0]):
                if parts[2] == "gate_up_proj":
                    g, u = t[e].chunk(2, dim=0)
                    out += [(layer, f"moe.experts.{e}.gate", g), (layer, f"moe.experts.{e}.up", u)]
                elif parts[2] == "down_proj":
                    out.append((layer, f"moe.experts.{e}.down", t[e]))
            return out
        if sub == "experts" and len(parts) >= 5:
            e, proj = int(parts[2]), parts[3]
            table = {"w1": "gate", "w3": "up", "w2": "down", "gate_proj": "gate", "up_proj": "up", "down_proj": "down"}
            if proj in table:
                return [(layer, f"moe.experts.{e}.{table[proj]}", t)]
            return []
        if sub == "shared_expert":
            return [(layer, f"moe.shared.{parts[2].split('_')[0]}", t)]
        if sub == "shared_expert_gate":
            return [(layer, "moe.shared_gate", t)]
    return []


# ---------------------------------------------------------------------------------------------------
@register_mapper("opt")
def map_opt(name: str, t: torch.Tensor, spec: ArchSpec) -> List[Piece]:
    layer, rest = _split_layer(name)
    kind = name.split(".")[-1]
    if layer is None:
        if name.endswith("decoder.embed_tokens.weight"):
            return [(None, "embed", t)]
        if name.endswith("decoder.embed_positions.weight"):
            return [(None, "pos_embed", t)]
        if "decoder.final_layer_norm." in name:
            return [(None, "final_norm.g" if kind == "weight" else "final_norm.b", t)]
        return []  # lm_head is tied
    if name.split(".")[-1].startswith("lora_"):
        return []
    parts = rest.split(".")
    if parts[0] == "self_attn":
        proj = parts[1]
        if proj in ("q_proj", "k_proj", "v_proj"):
            return [(layer, f"attn.{proj[0]}{_wb(kind)}", t)]
        if proj == "out_proj":
            return [(layer, f"attn.o{_wb(kind)}", t)]
    if parts[0] == "self_attn_layer_norm":
        return [(layer, "norm1.g" if kind == "weight" else "norm1.b", t)]
    if parts[0] == "final_layer_norm":
        return [(layer, "norm2.g" if kind == "weight" else "norm2.b", t)]
    if parts[0] == "fc1":
        return [(layer, f"mlp.up{_wb(kind)}", t)]
    if parts[0] == "fc2":
        return [(layer, f"mlp.down{_wb(kind)}", t)]
    return []


# ---------------------------------------------------------------------------------------------------
def _split_falcon_qkv(t: torch.Tensor, spec: ArchSpec):
    d, nq, nkv = spec.head_dim, spec.n_heads, spec.n_kv_heads
    tail = tuple(t.shape[1:])
    if spec.extras.get("new_decoder_architecture", False):
        ratio = nq // nkv
        g = t.reshape((nkv, ratio + 2, d) + tail)
        q = g[:, :ratio].reshape((nq * d, ) + tail)
        k = g[:, ratio].reshape((nkv * d, ) + tail)
        v = g[:, ratio + 1].reshape((nkv * d, ) + tail)
        return q, k, v
    if nkv == 1:  # multi-query: [all Q heads; K; V]
        return t[:nq * d], t[nq * d:(nq + 1) * d], t[(nq + 1) * d:]
    g = t.reshape((nq, 3, d) + tail)  # per-head interleaved q, k, v
    return (g[:, 0].reshape((nq * d, ) + tail), g[:, 1].reshape((nq * d, ) + tail), g[:, 2].reshape((nq * d, ) + tail))


@register_mapper("falcon")
def map_falcon(name: str, t: torch.Tensor, spec: ArchSpec) -> List[Piece]:
    layer, rest = _split_layer(name)
    kind = name.split(".")[-1]
    if layer is None:
        if name.endswith("word_embeddings.weight"):
            return [(None, "embed", t)]
        if ".ln_f." in name or name.startswith("ln_f."):
            return [(None, "final_norm.g" if kind == "weight" else "final_norm.b", t)]
        return []
    if name.split(".")[-1].startswith("lora_"):
        return []
    parts = rest.split(".")
    gb = "g" if kind == "weight" else "b"
    if parts[0] == "self_attention":
        if parts[1] == "query_key_value":
            q, k, v = _split_falcon_qkv(t, spec)
            sfx = _wb(kind)
            return [(layer, f"attn.q{sfx}", q), (layer, f"attn.k{sfx}", k), (layer, f"attn.v{sfx}", v)]
        if parts[1] == "dense":
            return [(layer, f"attn.o{_wb(kind)}", t)]
    if parts[0] == "mlp":
        if parts[1] == "dense_h_to_4h":
            return [(layer, f"mlp.up{_wb(kind)}", t)]
        if parts[1] == "dense_4h_to_h":
            return [(layer, f"mlp.down{_wb(kind)}", t)]
    if parts[0] in ("input_layernorm", "ln_attn"):
        return [(layer, f"norm1.{gb}", t)]
    if parts[0] in ("post_attention_layernorm", "ln_mlp"):
        return [(layer, f"norm2.{gb}", t)]
    return []


# ---------------------------------------------------------------------------------------------------
@register_mapper("phi")
def map_phi(name: str, t: torch.Tensor, spec: ArchSpec) -> List[Piece]:
    layer, rest = _split_layer(name)
    kind = name.split(".")[-1]
    if layer is None:
        if name.endswith("embed_tokens.weight"):
            return [(None, "embed", t)]
        if "final_layernorm." in name:
            return [(None, "final_norm.g" if kind == "weight" else "final_norm.b", t)]
        if name == "lm_head.weight":
            return [(None, "lm_head", t)]
        if name == "lm_head.bias":
            return [(None, "lm_head_b", t)]
        return []
    if name.split(".")[-1].startswith("lora_"):
        return []
    parts = rest.split(".")
    if parts[0] == "self_attn":
        if parts[1] in ("q_proj", "k_proj", "v_proj"):
            return [(layer, f"attn.{parts[1][0]}{_wb(kind)}", t)]
        if parts[1] == "dense":
            return [(layer, f"attn.o{_wb(kind)}", t)]
    if parts[0] == "mlp":
        if parts[1] == "fc1":
            return [(layer, f"mlp.up{_wb(kind)}", t)]
        if parts[1] == "fc2":
            return [(layer, f"mlp.down{_wb(kind)}", t)]
    if parts[0] == "input_layernorm":
        return [(layer, "norm1.g" if kind == "weight" else "norm1.b", t)]
    return []


# ---------------------------------------------------------------------------------------------------
@register_mapper("qwen")
def map_qwen(name: str, t: torch.Tensor, spec: ArchSpec) -> List[Piece]:
    layer, rest = _split_layer(name)
    kind = name.split(".")[-1]
    if layer is None:
        if name.endswith("wte.weight"):
            return [(None, "embed", t)]
        if name.endswith("ln_f.weight"):
            return [(None, "final_norm.g", t)]
        if name == "lm_head.weight":
            return [(None, "lm_head", t)]
        return []
    if name.split(".")[-1].startswith("lora_"):
        return []
    parts = rest.split(".")
    if parts[0] == "ln_1":
        return [(layer, "norm1.g", t)]
    if parts[0] == "ln_2":
        return [(layer, "norm2.g", t)]
    if parts[0] == "attn":
        if parts[1] == "c_attn":
            q, k, v = t.chunk(3, dim=0)
            sfx = _wb(kind)
            return [(layer, f"attn.q{sfx}", q), (layer, f"attn.k{sfx}", k), (layer, f"attn.v{sfx}", v)]
        if parts[1] == "c_proj":
            return [(layer, f"attn.o{_wb(kind)}", t)]
    if parts[0] == "mlp":
        # Qwen-1: out = c_proj(w1(x) * silu(w2(x)))  ->  w2 is the gate, w1 the up projection
        table = {"w1": "up", "w2": "gate", "c_proj": "down"}
        if parts[1] in table:
            return [(layer, f"mlp.{table[parts[1]]}", t)]
    return []


# ---------------------------------------------------------------------------------------------------
# GPT-2 / GPT-J / GPT-NeoX: served by the reference through v1 kernel injection only
# (module_inject/containers/{gpt2,gptj,gptneox}.py); here they load into the same executor as everything else.
# ---------------------------------------------------------------------------------------------------
@register_mapper("gpt2")
def map_gpt2(name: str, t: torch.Tensor, spec: ArchSpec) -> List[Piece]:
    layer, rest = _split_layer(name)
    kind = name.split(".")[-1]
    if layer is None:
        if name.endswith("wte.weight"):
            return [(None, "embed", t)]
        if name.endswith("wpe.weight"):
            return [(None, "pos_embed", t)]
        if "ln_f." in name:
            return [(None, "final_norm.g" if kind == "weight" else "final_norm.b", t)]
        return []  # lm_head is tied to wte
    parts = rest.split(".")
    if kind not in ("weight", "bias") or len(parts) < 2:
        return []  # causal-mask buffers (attn.bias / attn.masked_bias)
    gb = "g" if kind == "weight" else "b"
    if parts[0] == "ln_1":
        return [(layer, f"norm1.{gb}", t)]
    if parts[0] == "ln_2":
        return [(layer, f"norm2.{gb}", t)]
    # HF's Conv1D keeps weights [in, out]: transpose to the [out, in] layout every GEMM here takes
    w = t.t() if (kind == "weight" and t.dim() == 2) else t
    sfx = _wb(kind)
    if parts[0] == "attn" and len(parts) == 3:
        if parts[1] == "c_attn":
            q, k, v = w.chunk(3, dim=0)
            return [(layer, f"attn.q{sfx}", q), (layer, f"attn.k{sfx}", k), (layer, f"attn.v{sfx}", v)]
        if parts[1] == "c_proj":
            return [(layer, f"attn.o{sfx}", w)]
    if parts[0] == "mlp" and len(parts) == 3:
        if parts[1] == "c_fc":
            return [(layer, f"mlp.up{sfx}", w)]
        if parts[1] == "c_proj":
            return [(layer, f"mlp.down{sfx}", w)]
    return []


def _pairs_to_halves(t: torch.Tensor, spec: ArchSpec) -> torch.Tensor:
    """GPT-J rotates adjacent pairs (2i, 2i+1); the kernels rotate (i, i + rot/2). Re-ordering the first ``rot`` output
    rows of every Q and K head the same way leaves every q.k dot product unchanged, so the checkpoint is converted once
    here instead of carrying a second RoPE variant through the kernels."""
    d, rot = spec.head_dim, spec.rot_dim
    perm = torch.cat([torch.arange(0, rot, 2), torch.arange(1, rot, 2), torch.arange(rot, d)])
    heads = t.shape[0] // d
    return t.reshape((heads, d) + tuple(t.shape[1:]))[:, perm].reshape(t.shape)


@register_mapper("gptj")
def map_gptj(name: str, t: torch.Tensor, spec: ArchSpec) -> List[Piece]:
    layer, rest = _split_layer(name)
    kind = name.split(".")[-1]
    if layer is None:
        if name.endswith("wte.weight"):
            return [(None, "embed", t)]
        if "ln_f." in name:
            return [(None, "final_norm.g" if kind == "weight" else "final_norm.b", t)]
        if name == "lm_head.weight":
            return [] if spec.tie_word_embeddings else [(None, "lm_head", t)]
        if name == "lm_head.bias":
            return [(None, "lm_head_b", t)]
        return []
    parts = rest.split(".")
    if kind not in ("weight", "bias"):
        return []
    sfx = _wb(kind)
    if parts[0] == "ln_1":
        return [(layer, "norm1.g" if kind == "weight" else "norm1.b", t)]
    if parts[0] == "attn" and len(parts) == 3:
        if parts[1] in ("q_proj", "k_proj"):
            return [(layer, f"attn.{parts[1][0]}{sfx}", _pairs_to_halves(t, spec))]
        if parts[1] == "v_proj":
            return [(layer, f"attn.v{sfx}", t)]
        if parts[1] == "out_proj":
            return [(layer, f"attn.o{sfx}", t)]
    if parts[0] == "mlp" and len(parts) == 3:
        if parts[1] == "fc_in":
            return [(layer, f"mlp.up{sfx}", t)]
        if parts[1] == "fc_out":
            return [(layer, f"mlp.down{sfx}", t)]
    return []


@register_mapper("gpt_neox")
def map_gpt_neox(name: str, t: torch.Tensor, spec: ArchSpec) -> List[Piece]:
    layer, rest = _split_layer(name)
    kind = name.split(".")[-1]
    if layer is None:
        if name.endswith("embed_in.weight"):
            return [(None, "embed", t)]
        if "final_layer_norm." in name:
            return [(None, "final_norm.g" if kind == "weight" else "final_norm.b", t)]
        if name == "embed_out.weight":
            return [] if spec.tie_word_embeddings else [(None, "lm_head", t)]
        return []
    parts = rest.split(".")
    if kind not in ("weight", "bias"):
        return []  # rotary inv_freq / mask buffers of older checkpoints
    gb = "g" if kind == "weight" else "b"
    sfx = _wb(kind)
    if parts[0] == "input_layernorm":
        return [(layer, f"norm1.{gb}", t)]
    if parts[0] == "post_attention_layernorm":
        return [(layer, f"norm2.{gb}", t)]
    if parts[0] == "attention" and len(parts) == 3:
        if parts[1] == "query_key_value":   # rows grouped per head: [head][q | k | v][head_dim]
            g = t.reshape((spec.n_heads, 3, spec.head_dim) + tuple(t.shape[1:]))
            flat = (spec.q_dim, ) + tuple(t.shape[1:])
            return [(layer, f"attn.q{sfx}", g[:, 0].reshape(flat)), (layer, f"attn.k{sfx}", g[:, 1].reshape(flat)),
                    (layer, f"attn.v{sfx}", g[:, 2].reshape(flat))]
        if parts[1] == "dense":
            return [(layer, f"attn.o{sfx}", t)]
    if parts[0] == "mlp" and len(parts) == 3:
        if parts[1] == "dense_h_to_4h":
            return [(layer, f"mlp.up{sfx}", t)]
        if parts[1] == "dense_4h_to_h":
            return [(layer, f"mlp.down{sfx}", t)]
    return []


# ---------------------------------------------------------------------------------------------------
@register_mapper("bloom")
def map_bloom(name: str, t: torch.Tensor, spec: ArchSpec) -> List[Piece]:
    """Bloom (ref: module_inject/containers/bloom.py): ALiBi attention, LayerNorm on the embeddings, fused QKV with rows
    grouped per head [head][q | k | v][head_dim]."""
    layer, rest = _split_layer(name)
    kind = name.split(".")[-1]
    if kind not in ("weight", "bias"):
        return []
    gb = "g" if kind == "weight" else "b"
    if layer is None:
        if name.endswith("word_embeddings.weight"):
            return [(None, "embed", t)]
        if "word_embeddings_layernorm." in name:
            return [(None, f"embed_norm.{gb}", t)]
        if ".ln_f." in name or name.startswith("ln_f."):
            return [(None, f"final_norm.{gb}", t)]
        return []  # lm_head is tied
    parts = rest.split(".")
    sfx = _wb(kind)
    if parts[0] == "input_layernorm":
        return [(layer, f"norm1.{gb}", t)]
    if parts[0] == "post_attention_layernorm":
        return [(layer, f"norm2.{gb}", t)]
    if parts[0] == "self_attention" and len(parts) == 3:
        if parts[1] == "query_key_value":
            g = t.reshape((spec.n_heads, 3, spec.head_dim) + tuple(t.shape[1:]))
            flat = (spec.q_dim, ) + tuple(t.shape[1:])
            return [(layer, f"attn.q{sfx}", g[:, 0].reshape(flat)), (layer, f"attn.k{sfx}", g[:, 1].reshape(flat)),
                    (layer, f"attn.v{sfx}", g[:, 2].reshape(flat))]
        if parts[1] == "dense":
            return [(layer, f"attn.o{sfx}", t)]
    if parts[0] == "mlp" and len(parts) == 3:
        if parts[1] == "dense_h_to_4h":
            return [(layer, f"mlp.up{sfx}", t)]
        if parts[1] == "dense_4h_to_h":
            return [(layer, f"mlp.down{sfx}", t)]
    return []


# ---------------------------------------------------------------------------------------------------
@register_mapper("gpt_neo")
def map_gpt_neo(name: str, t: torch.Tensor, spec: ArchSpec) -> List[Piece]:
    """GPT-Neo (ref: module_inject/containers/gptneo.py): nn.Linear projections under ``attn.attention``, no Q/K/V bias."""
    layer, rest = _split_layer(name)
    kind = name.split(".")[-1]
    if kind not in ("weight", "bias"):
        return []
    gb = "g" if kind == "weight" else "b"
    if layer is None:
        if name.endswith("wte.weight"):
            return [(None, "embed", t)]
        if name.endswith("wpe.weight"):
            return [(None, "pos_embed", t)]
        if "ln_f." in name:
            return [(None, f"final_norm.{gb}", t)]
        return []
    parts = rest.split(".")
    sfx = _wb(kind)
    if parts[0] == "ln_1":
        return [(layer, f"norm1.{gb}", t)]
    if parts[0] == "ln_2":
        return [(layer, f"norm2.{gb}", t)]
    if parts[0] == "attn" and len(parts) == 4 and parts[1] == "attention":
        if parts[2] in ("q_proj", "k_proj", "v_proj"):
            return [(layer, f"attn.{parts[2][0]}{sfx}", t)]
        if parts[2] == "out_proj":
            return [(layer, f"attn.o{sfx}", t)]
    if parts[0] == "mlp" and len(parts) == 3:
        if parts[1] == "c_fc":
            return [(layer, f"mlp.up{sfx}", t)]
        if parts[1] == "c_proj":
            return [(layer, f"mlp.down{sfx}", t)]
    return []
