"""Splitting fused QKV (and gate/up) weights across tensor-parallel ranks for the layouts HF models use.
Reference: ``deepspeed/module_inject/fusedqkv_utils.py:29-150``."""
from __future__ import annotations

import torch

from .tp_shard import get_shard_size_list

FUSED_QKV_TYPES = {
    "CodeGenBlock": "codegentype", "BloomBlock": "bloomtype", "GLMBlock": "glmtype", "MPTBlock": "glmtype",
    "MptBlock": "glmtype", "BaichuanLayer": "glmtype", "QWenBlock": "qwentype", "FalconDecoderLayer": "bloomtype",
    "GPTBigCodeBlock": "bigcodetype", "DecoderLayer": "glmtype", "Phi3DecoderLayer": "phi3type", "GPTNeoXLayer": "bloomtype",
    "GPT2Block": "glmtype",
}


def fused_qkv_layout(block) -> str:
    """Layout name of ``block``'s fused QKV weight. Falcon has three, chosen by its attention module's flags: grouped
    [kv group][ratio x q | k | v] (new decoder architecture), [q heads | k | v] (multi-query) or per-head [q k v]."""
    name = type(block).__name__
    if "Falcon" in name:
        for m in block.modules():
            if hasattr(m, "new_decoder_architecture") and hasattr(m, "multi_query"):
                if m.new_decoder_architecture:
                    return "falcon_grouped"
                return "bigcodetype" if m.multi_query else "bloomtype"
    for k in sorted(FUSED_QKV_TYPES, key=len, reverse=True):     # most specific class name first
        if k in name:
            return FUSED_QKV_TYPES[k]
    return "glmtype"


def _falcon_grouped(w: torch.Tensor, mp_size: int, rank: int, n_q: int, n_kv: int) -> torch.Tensor:
    """[kv group][(n_q / n_kv) query heads | k | v][head_dim]: whole groups per rank."""
    per = n_q // n_kv + 2
    d = w.shape[0] // (n_kv * per)
    g = w.reshape(n_kv, per * d, *w.shape[1:])
    sizes = get_shard_size_list(n_kv, mp_size)
    assert min(sizes) > 0, "grouped fused QKV needs at least one KV group per rank"
    start = sum(sizes[:rank])
    return g[start:start + sizes[rank]].reshape(-1, *w.shape[1:])


def require_tp_fused_qkvw(name: str, mp_size: int) -> bool:
    if mp_size == 1:
        return False
    return any(k in name for k in ("qkv_proj", "query_key_value", "attn.Wqkv", "self_attn.W_pack", "c_attn"))


def _glm_type(w: torch.Tensor, mp_size: int, rank: int) -> torch.Tensor:
    """[q_all | k_all | v_all] stacked along dim 0: slice each third."""
    thirds = torch.chunk(w, 3, dim=0)
    outs = []
    for t in thirds:
        sizes = get_shard_size_list(t.shape[0], mp_size)
        start = sum(sizes[:rank])
        outs.append(t[start:start + sizes[rank]])
    return torch.cat(outs, dim=0)


def _bloom_type(w: torch.Tensor, mp_size: int, rank: int, num_heads: int) -> torch.Tensor:
    """Per-head interleaved [h, (q k v), d]: slice whole heads."""
    hd3 = w.shape[0] // num_heads
    v = w.reshape(num_heads, hd3, *w.shape[1:])
    sizes = get_shard_size_list(num_heads, mp_size)
    start = sum(sizes[:rank])
    return v[start:start + sizes[rank]].reshape(-1, *w.shape[1:])


def _phi3_type(w: torch.Tensor, mp_size: int, rank: int, n_q: int, n_kv: int) -> torch.Tensor:
    """[q (n_q heads) | k (n_kv) | v (n_kv)] with GQA."""
    d = w.shape[0] // (n_q + 2 * n_kv)
    q, k, v = torch.split(w, [n_q * d, n_kv * d, n_kv * d], dim=0)
    outs = []
    for t, heads in ((q, n_q), (k, n_kv), (v, n_kv)):
        sizes = get_shard_size_list(heads, mp_size)
        start = sum(sizes[:rank]) * d
        outs.append(t[start:start + sizes[rank] * d])
    return torch.cat(outs, dim=0)


def prepare_tp_fused_qkvw(module, src: torch.Tensor, mp_size: int, gpu_index: int, num_heads: int = None,
                          num_kv_heads: int = None) -> torch.Tensor:
    """Return rank ``gpu_index``'s slice of a fused QKV weight / bias of ``module``'s block type."""
    kind = None
    name = module if isinstance(module, str) else type(module).__name__
    if name in ("glmtype", "qwentype", "codegentype", "bloomtype", "phi3type", "bigcodetype", "falcon_grouped"):
        kind = name
    else:
        for k in sorted(FUSED_QKV_TYPES, key=len, reverse=True):
            if k in name:
                kind = FUSED_QKV_TYPES[k]
                break
    if kind is None:
        kind = "glmtype"
    if kind == "falcon_grouped":
        assert num_heads is not None and num_kv_heads is not None
        return _falcon_grouped(src, mp_size, gpu_index, num_heads, num_kv_heads)
    if kind in ("glmtype", "qwentype", "codegentype"):
        return _glm_type(src, mp_size, gpu_index)
    if kind == "bloomtype":
        assert num_heads is not None, "bloom-style fused QKV needs the head count"
        return _bloom_type(src, mp_size, gpu_index, num_heads)
    if kind == "phi3type":
        assert num_heads is not None
        return _phi3_type(src, mp_size, gpu_index, num_heads, num_kv_heads or num_heads)
    if kind == "bigcodetype":  # multi-query: q sharded, the single k/v head replicated
        d = src.shape[0] - 2 * (src.shape[0] // (num_heads + 2)) if num_heads else src.shape[0]
        q, kv = src[:d], src[d:]
        sizes = get_shard_size_list(q.shape[0], mp_size)
        start = sum(sizes[:gpu_index])
        return torch.cat([q[start:start + sizes[gpu_index]], kv], dim=0)
    raise ValueError(f"unsupported fused qkv type {kind}")


def shard_chunk_mlp(weight: torch.Tensor, bias, rank: int, world_size: int):
    """Fused gate_up weight [2*I, H] -> this rank's [2*I/tp, H] keeping gate and up halves aligned."""
    gate, up = torch.chunk(weight, 2, dim=0)
    sizes = get_shard_size_list(gate.shape[0], world_size, "mlp")
    start = sum(sizes[:rank])
    w = torch.cat([gate[start:start + sizes[rank]], up[start:start + sizes[rank]]], dim=0)
    b = None
    if bias is not None:
        bg, bu = torch.chunk(bias, 2, dim=0)
        b = torch.cat([bg[start:start + sizes[rank]], bu[start:start + sizes[rank]]], dim=0)
    return w, b
