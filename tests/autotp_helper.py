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


HF_AUTOTP_CASES = ("llama_gqa", "opt", "gpt2", "gptj", "gpt_neox", "bloom", "falcon_mqa", "falcon_new", "falcon_mha", "phi", "phi3",
                   "qwen2", "mixtral", "gpt_neo")


def _hf_tiny(kind):
    import transformers as T
    common = dict(hidden_size=64, num_hidden_layers=2, num_attention_heads=4, vocab_size=120, max_position_embeddings=128)
    if kind == "llama_gqa":
        return T.LlamaForCausalLM(T.LlamaConfig(intermediate_size=128, num_key_value_heads=2, **common))
    if kind == "qwen2":
        return T.Qwen2ForCausalLM(T.Qwen2Config(intermediate_size=128, num_key_value_heads=2, **common))
    if kind == "opt":
        return T.OPTForCausalLM(T.OPTConfig(ffn_dim=128, word_embed_proj_dim=64, **common))
    if kind == "gpt2":
        return T.GPT2LMHeadModel(T.GPT2Config(n_embd=64, n_layer=2, n_head=4, vocab_size=120, n_positions=128))
    if kind == "gptj":
        return T.GPTJForCausalLM(T.GPTJConfig(n_embd=64, n_layer=2, n_head=4, vocab_size=120, n_positions=128, rotary_dim=8))
    if kind == "gpt_neox":
        return T.GPTNeoXForCausalLM(T.GPTNeoXConfig(intermediate_size=128, **common))
    if kind == "gpt_neo":
        return T.GPTNeoForCausalLM(T.GPTNeoConfig(hidden_size=64, num_layers=2, num_heads=4, vocab_size=120, window_size=8,
                                                  attention_types=[[["global", "local"], 1]], max_position_embeddings=128))
    if kind == "bloom":
        return T.BloomForCausalLM(T.BloomConfig(hidden_size=64, n_layer=2, n_head=4, vocab_size=120))
    if kind == "falcon_mqa":
        return T.FalconForCausalLM(T.FalconConfig(hidden_size=64, num_hidden_layers=2, num_attention_heads=4, vocab_size=120,
                                                  multi_query=True, parallel_attn=True, new_decoder_architecture=False,
                                                  alibi=False, bias=False))
    if kind == "falcon_new":
        return T.FalconForCausalLM(T.FalconConfig(hidden_size=64, num_hidden_layers=2, num_attention_heads=4, num_kv_heads=2,
                                                  vocab_size=120, new_decoder_architecture=True, alibi=False, bias=False))
    if kind == "falcon_mha":
        return T.FalconForCausalLM(T.FalconConfig(hidden_size=64, num_hidden_layers=2, num_attention_heads=4, vocab_size=120,
                                                  multi_query=False, parallel_attn=False, new_decoder_architecture=False,
                                                  alibi=False, bias=True))
    if kind == "phi3":
        return T.Phi3ForCausalLM(T.Phi3Config(intermediate_size=128, num_key_value_heads=2, pad_token_id=0, **common))
    if kind == "phi":
        return T.PhiForCausalLM(T.PhiConfig(intermediate_size=128, partial_rotary_factor=0.5, **common))
    if kind == "mixtral":
        return T.MixtralForCausalLM(T.MixtralConfig(intermediate_size=96, num_key_value_heads=2, num_local_experts=4,
                                                    num_experts_per_tok=2, sliding_window=None, **common))
    raise KeyError(kind)


def hf_autotp_worker(rank, world, kinds):
    """AutoTP (no kernel injection) on real HuggingFace architectures: logits of the sharded model == unsharded logits."""
    import hcache_deepspeed_b200 as ds
    ds.init_distributed(dist_backend="gloo")
    out = {}
    for kind in kinds:
        torch.manual_seed(0)
        try:
            model = _hf_tiny(kind).eval().float()
        except Exception as e:  # config drift between transformers versions
            out[kind] = f"skip: {type(e).__name__}: {e}"
            continue
        ids = torch.randint(1, 120, (2, 11))
        with torch.no_grad():
            want = model(ids).logits
        try:
            eng = ds.init_inference(model, dtype=torch.float32, tensor_parallel={"tp_size": world})
            with torch.no_grad():
                got = eng(ids).logits
            out[kind] = float((got - want).abs().max())
        except Exception as e:
            import traceback
            out[kind] = f"error: {type(e).__name__}: {e} | {traceback.format_exc().splitlines()[-3].strip()}"
    return out


def v1_inject_tp_worker(rank, world, kinds):
    """``init_inference(replace_with_kernel_inject=True, tensor_parallel={tp_size: world})`` on real HF models: the stack
    moves onto the ragged executor, sharded over the ranks; logits match the unsharded HF model and ``generate`` runs."""
    import hcache_deepspeed_b200 as ds
    ds.init_distributed(dist_backend="gloo")
    out = {}
    for kind in kinds:
        torch.manual_seed(0)
        model = _hf_tiny(kind).eval().float()
        ids = torch.randint(1, 120, (2, 11))
        with torch.no_grad():
            want = model(ids).logits
            want_gen = model.generate(ids, max_new_tokens=4, do_sample=False, pad_token_id=0)
        eng = ds.init_inference(model, dtype=torch.float32, replace_with_kernel_inject=True, tensor_parallel={"tp_size": world})
        got = eng(ids).logits
        gen = eng.generate(ids, max_new_tokens=4, do_sample=False)
        out[kind] = {"err": float((got - want).abs().max()), "injected": bool(eng.injected), "gen_equal": bool(torch.equal(gen, want_gen))}
    return out


def injection_policy_worker(rank, world):
    """``init_inference(model, tensor_parallel=..., injection_policy={Block: (row-parallel linears)})`` on a model whose
    layer names the parser could not guess."""
    import hcache_deepspeed_b200 as ds
    from hcache_deepspeed_b200.module_inject import LinearAllreduce, LinearLayer
    ds.init_distributed(dist_backend="gloo")

    class OddBlock(nn.Module):
        def __init__(self, h):
            super().__init__()
            self.first, self.merge = nn.Linear(h, 2 * h), nn.Linear(2 * h, h)     # "merge" is the row-parallel one

        def forward(self, x):
            return x + self.merge(torch.relu(self.first(x)))

    class OddLM(nn.Module):
        def __init__(self):
            super().__init__()
            self.layers = nn.ModuleList([OddBlock(16) for _ in range(2)])

        def forward(self, x):
            for l in self.layers:
                x = l(x)
            return x

    torch.manual_seed(0)
    model = OddLM().eval()
    x = torch.randn(3, 16)
    with torch.no_grad():
        want = model(x)
    bad = None
    try:
        ds.init_inference(OddLM(), dtype=torch.float32, tensor_parallel={"tp_size": world}, injection_policy={"NoSuchBlock": ("merge", )})
    except ValueError as e:
        bad = str(e)
    eng = ds.init_inference(model, dtype=torch.float32, tensor_parallel={"tp_size": world}, injection_policy={OddBlock: ("merge", )})
    with torch.no_grad():
        got = eng(x)
    blk = model.layers[0]
    return {"err": float((got - want).abs().max()), "types": [type(blk.first).__name__, type(blk.merge).__name__],
            "bad": bad}


def tp_training_worker(rank, world, tp):
    """Training-time tensor parallelism (``tp_model_init`` + ``"tensor_parallel": {"autotp_size": tp}``) on a HuggingFace
    Llama: losses equal the unsharded model's, step by step; with world > tp the remaining ranks are data parallel."""
    import transformers as T
    import hcache_deepspeed_b200 as ds
    from hcache_deepspeed_b200.runtime.tensor_parallel.tp_manager import TpTrainingManager
    ds.init_distributed(dist_backend="gloo")
    dp = world // tp
    torch.manual_seed(0)
    cfg_m = T.LlamaConfig(hidden_size=64, num_hidden_layers=2, num_attention_heads=4, vocab_size=120, max_position_embeddings=128,
                          intermediate_size=128, num_key_value_heads=2)
    model = T.LlamaForCausalLM(cfg_m)
    g = torch.Generator().manual_seed(5)
    batches = [torch.randint(1, 120, (2, 16), generator=g) for _ in range(dp)]       # one micro-batch per data-parallel replica
    ref = T.LlamaForCausalLM(cfg_m)
    ref.load_state_dict(model.state_dict())
    opt = torch.optim.SGD(ref.parameters(), lr=0.1)
    want = []
    for _ in range(4):
        losses = [ref(input_ids=b, labels=b).loss for b in batches]
        (sum(losses) / dp).backward()
        opt.step()
        opt.zero_grad()
        want.append([float(l.detach()) for l in losses])
    model = ds.tp_model_init(model, tp_size=tp, dtype=torch.float32)
    engine, _, _, _ = ds.initialize(model=model, config={
        "train_micro_batch_size_per_gpu": 2, "optimizer": {"type": "SGD", "params": {"lr": 0.1}}, "steps_per_print": 1000,
        "tensor_parallel": {"autotp_size": tp}, "zero_optimization": {"stage": 1}})
    mine = TpTrainingManager.broadcast_batch(batches[rank // tp].clone())           # the TP group's ranks share one batch
    got = []
    for _ in range(4):
        loss = engine(input_ids=mine, labels=mine).loss
        engine.backward(loss)
        engine.step()
        got.append(float(loss.detach()))
    return {"want": [w[rank // tp] for w in want], "got": got, "dp": engine.dp_world_size,
            "q_rows": model.model.layers[0].self_attn.q_proj.weight.shape[0]}


def tp_layers_worker(rank, world):
    """The layers that shard themselves from a full module (``Layer(module, mp_group)``; ref: module_inject/layers.py:142-720):
    column / row parallel linears, their Conv1D variants, packed gate-up, sharded Conv2d; frozen in inference mode, trainable
    with complete gradients in training mode, and reassembled by ``gather_params``."""
    import copy
    import torch
    import hcache_deepspeed_b200 as ds
    from hcache_deepspeed_b200.module_inject import layers as L
    from transformers.pytorch_utils import Conv1D
    ds.init_distributed(dist_backend="gloo")
    group = ds.comm.new_group(list(range(world)))
    torch.manual_seed(0)
    out = {}
    x = torch.randn(2, 5, 6)
    up, down = torch.nn.Linear(6, 6), torch.nn.Linear(6, 4)
    want = down(torch.relu(up(x)))
    col = L.LinearLayer(copy.deepcopy(up), mp_group=group)
    row = L.LinearAllreduce(copy.deepcopy(down), group)
    col.eval(), row.eval()
    got = row(torch.relu(col(x)))
    out["col_err"] = float((col(x) - up(x).chunk(world, dim=-1)[rank]).abs().max())
    out["row_err"] = float((got - want).abs().max())
    out["col_shape"], out["row_shape"] = tuple(col.weight.shape), tuple(row.weight.shape)
    out["frozen"] = not col.weight.requires_grad and not row.weight.requires_grad and col.weight.tensor_model_parallel
    # Conv1D-born variants
    c_up, c_down = Conv1D(6, 6), Conv1D(4, 6)
    torch.nn.init.normal_(c_up.weight), torch.nn.init.normal_(c_down.weight)
    want_c = c_down(torch.relu(c_up(x)))
    ccol = L.conv_LinearLayer(copy.deepcopy(c_up), mp_group=group).eval()
    crow = L.Conv_LinearALlreduce(copy.deepcopy(c_down), group).eval()
    out["conv_err"] = float((ccol(x) - c_up(x).chunk(world, dim=-1)[rank]).abs().max())
    out["convrow_err"] = float((crow(torch.relu(ccol(x))) - want_c).abs().max())
    # packed gate+up: every rank holds matching slices of both halves
    packed = torch.nn.Linear(6, 8)
    gu = L.GateUpPack_LinearLayer(copy.deepcopy(packed), mp_group=group).eval()
    g_full, u_full = packed(x).chunk(2, dim=-1)
    g, u = gu(x).chunk(2, dim=-1)
    out["gateup_err"] = float(max((g - g_full.chunk(world, dim=-1)[rank]).abs().max(), (u - u_full.chunk(world, dim=-1)[rank]).abs().max()))
    # Conv2d sharded over output channels, then over input channels with a reduced sum
    conv_a, conv_b = torch.nn.Conv2d(3, 4, 3, padding=1), torch.nn.Conv2d(4, 2, 1)
    img = torch.randn(1, 3, 5, 5)
    want_img = conv_b(conv_a(img))
    oc = L.TensorParallelOcShardConv2d(copy.deepcopy(conv_a), rank, world)
    ic = L.TensorParallelIcShardConv2d(copy.deepcopy(conv_b), rank, world)
    out["conv2d_err"] = float((ic(oc(img)) - want_img).abs().max())
    # training mode: shards get gradients, the input gradient is complete, gather_params rebuilds the full weight
    L.set_autotp_mode(training=True)
    out["mode_roundtrip"] = L.is_autotp_training_mode()
    tcol = L.LinearLayer(copy.deepcopy(up), mp_group=group).train()
    trow = L.LinearAllreduce(copy.deepcopy(down), group).train()
    out["trainable"] = tcol.weight.requires_grad and trow.weight.requires_grad
    xi = x.clone().requires_grad_(True)
    trow(torch.relu(tcol(xi))).pow(2).sum().backward()
    xr = x.clone().requires_grad_(True)
    down(torch.relu(up(xr))).pow(2).sum().backward()
    out["train_grad_err"] = float(max((xi.grad - xr.grad).abs().max(),
                                      (tcol.weight.grad - up.weight.grad.chunk(world, dim=0)[rank]).abs().max()))
    tcol.gather_params([tcol.weight, tcol.bias])
    trow.gather_params([trow.weight])
    out["gathered"] = bool(torch.equal(tcol.weight.data, up.weight.data) and torch.equal(trow.weight.data, down.weight.data))
    L.set_autotp_mode(training=False)
    out["mode_roundtrip"] = out["mode_roundtrip"] and not L.is_autotp_training_mode()
    return out
