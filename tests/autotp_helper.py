import torch
from torch import nn


class Attn(nn.Module):
    def __init__(self, h, heads):
        super().__init__()
        self.num_heads, self.head_dim = heads, h // heads
        self.q_proj, self.k_proj, self.v_proj = nn.Linear(h, h), nn.Linear(h, h), nn.Linear(h, h)
        self.o_proj = nn.Linear(h, h)

    def forward(self, x):
        b, s, _ = x.shape
        q = self.q_proj(x).view(b, s, self.num_heads, self.head_dim).transpose(1, 2)
        k = self.k_proj(x).view(b, s, self.num_heads, self.head_dim).transpose(1, 2)
        v = self.v_proj(x).view(b, s, self.num_heads, self.head_dim).transpose(1, 2)
        o = torch.nn.functional.scaled_dot_product_attention(q, k, v, is_causal=True)
        return self.o_proj(o.transpose(1, 2).reshape(b, s, self.num_heads * self.head_dim))


class MLP(nn.Module):
    def __init__(self, h):
        super().__init__()
        self.gate_proj, self.up_proj, self.down_proj = nn.Linear(h, 3 * h, bias=False), nn.Linear(h, 3 * h, bias=False), nn.Linear(3 * h, h, bias=False)

    def forward(self, x):
        return self.down_proj(torch.nn.functional.silu(self.gate_proj(x)) * self.up_proj(x))


class Block(nn.Module):
    def __init__(self, h, heads):
        super().__init__()
        self.ln1, self.ln2 = nn.LayerNorm(h), nn.LayerNorm(h)
        self.self_attn, self.mlp = Attn(h, heads), MLP(h)

    def forward(self, x):
        x = x + self.self_attn(self.ln1(x))
        return x + self.mlp(self.ln2(x))


class TinyLM(nn.Module):
    def __init__(self, vocab=50, h=32, heads=4, layers=2):
        super().__init__()
        self.embed = nn.Embedding(vocab, h)
        self.layers = nn.ModuleList([Block(h, heads) for _ in range(layers)])
        self.norm = nn.LayerNorm(h)
        self.lm_head = nn.Linear(h, vocab, bias=False)

    def forward(self, ids):
        x = self.embed(ids)
        for l in self.layers:
            x = l(x)
        return self.lm_head(self.norm(x))


def autotp_worker(rank, world):
    import hcache_deepspeed_b200 as ds
    from hcache_deepspeed_b200 import comm as dist
    from hcache_deepspeed_b200.module_inject import AutoTP, LinearAllreduce, LinearLayer, LmHeadLinearAllreduce
    ds.init_distributed(dist_backend="gloo")
    torch.manual_seed(0)
    model = TinyLM().eval()
    ids = torch.randint(0, 50, (2, 9))
    with torch.no_grad():
        want = model(ids)
    policies = AutoTP.tp_parser(model)
    AutoTP.apply(model, world, rank, dist.get_world_group())
    with torch.no_grad():
        got = model(ids)
    blk = model.layers[0]
    return {"err": float((got - want).abs().max()), "gems": sorted(policies[0][1]),
            "types": [type(blk.self_attn.q_proj).__name__, type(blk.self_attn.o_proj).__name__, type(blk.mlp.down_proj).__name__,
                      type(model.lm_head).__name__], "heads": blk.self_attn.num_heads,
            "q_rows": blk.self_attn.q_proj.weight.shape[0], "o_cols": blk.self_attn.o_proj.weight.shape[1]}
