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


def split_by_qkvlist_and_refuse(qkv_list, split_size, split_dim: int = 0, cat_dim: int = 0):
    """Each of Q, K, V is cut into pieces of ``split_size`` along ``split_dim``; piece ``i`` of all three is joined along
    ``cat_dim`` - one fused QKV block per tensor-parallel rank (ref: fusedqkv_utils.py:10)."""
    pieces = [torch.split(m, split_size, dim=split_dim) for m in qkv_list]
    return [torch.cat([p[i] for p in pieces], dim=cat_dim) for i in range(len(pieces[0]))]


def shard_value_with_share_qk(weight: torch.Tensor, bias, rank: int, world_size: int, shard_value: bool = True):
    """Value (``shard_value``) or output (otherwise) projection of attention blocks whose V heads are shared by pairs of
    neighbouring Q heads and again by the matching K heads: rank ``r`` owns Q heads ``[r*h, (r+1)*h)`` and needs V head
    ``q // 2`` for every second one of them, plus the same heads offset by half the head count for the K side. V is cut
    along its rows, the output projection along its columns (its bias is divided by the world size, to survive the
    all-reduce); ref: fusedqkv_utils.py:162."""
    from .tp_shard import get_num_kv_heads
    dim = 0 if shard_value else 1
    num_heads = get_num_kv_heads()
    head_dim = weight.shape[dim] // num_heads
    assert num_heads % world_size == 0, f"{num_heads} heads do not divide over {world_size} ranks"
    per_rank = num_heads // world_size
    first = rank * per_rank
    v_heads = [(first + i) // 2 for i in range(0, per_rank, 2)]
    v_heads += [h + num_heads // 2 for h in v_heads]
    w = torch.cat([weight.narrow(dim, h * head_dim, head_dim) for h in v_heads], dim=dim)
    if bias is None:
        return torch.nn.Parameter(w), None
    if shard_value:
        b = torch.cat([bias.data[h * head_dim:(h + 1) * head_dim] for h in v_heads], dim=0)
    else:
        b = bias.data / float(world_size)
    return torch.nn.Parameter(w), torch.nn.Parameter(b)
