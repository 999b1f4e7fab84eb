"""Shard a full (tp=1) kernel-layout ModelWeights for one TP rank (test helper: lets every rank of a TP test derive
its weights from the same model so logits can be compared with the single-rank run)."""
import torch

from hcache_deepspeed_b200.models.weights import ExpertWeights, LayerWeights, ModelWeights, interleave_gate_up, shard_info


def _rows(heads, d, base=0):
    return (torch.tensor(heads)[:, None] * d + torch.arange(d)[None, :] + base).reshape(-1)


def shard_full_weights(spec, full: ModelWeights, rank: int, world: int) -> ModelWeights:
    info = shard_info(spec, rank, world)
    d, nq, nkv = spec.head_dim, spec.n_heads, spec.n_kv_heads
    qr = _rows(info.q_heads, d)
    kr = _rows(info.kv_heads, d, nq * d)
    vr = _rows(info.kv_heads, d, (nq + nkv) * d)
    a, b = info.inter_range
    out = ModelWeights(embed=full.embed, pos_embed=full.pos_embed, final_norm_g=full.final_norm_g,
                       final_norm_b=full.final_norm_b, embed_norm_g=full.embed_norm_g, embed_norm_b=full.embed_norm_b)
    va, vb = info.vocab_range
    out.lm_head = full.lm_head[va:vb].contiguous()
    out.lm_head_b = full.lm_head_b[va:vb].contiguous() if full.lm_head_b is not None else None
    for lw in full.layers:
        n = LayerWeights(norm1_g=lw.norm1_g, norm1_b=lw.norm1_b, norm2_g=lw.norm2_g, norm2_b=lw.norm2_b)
        rows = torch.cat([qr, kr, vr])
        n.w_qkv = lw.w_qkv[rows].contiguous()
        n.b_qkv = lw.b_qkv[rows].contiguous() if lw.b_qkv is not None else None
        n.w_o = lw.w_o[:, qr].contiguous()
        n.b_o = lw.b_o if rank == 0 else None
        if spec.gated_mlp:
            gate, up = lw.w_up[0::2], lw.w_up[1::2]
            n.w_up = interleave_gate_up(gate[a:b], up[a:b])
            if lw.b_up is not None:
                n.b_up = interleave_gate_up(lw.b_up[0::2][a:b], lw.b_up[1::2][a:b])
        else:
            n.w_up = lw.w_up[a:b].contiguous()
            n.b_up = lw.b_up[a:b].contiguous() if lw.b_up is not None else None
        n.w_down = lw.w_down[:, a:b].contiguous()
        n.b_down = lw.b_down if rank == 0 else None
        out.layers.append(n)
    return out
