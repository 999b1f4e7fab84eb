"""v1 kernel injection through the per-family container policies: logits match HuggingFace, ``generate`` decodes one
token per step through the fused layers' KV cache (reference tests: tests/unit/inference/test_inference.py)."""
import copy

import pytest
import torch

transformers = pytest.importorskip("transformers")
from transformers import (AutoModel, AutoModelForCausalLM, BertConfig, BloomConfig, DistilBertConfig, GPT2Config,  # noqa: E402
                          GPTJConfig, GPTNeoConfig, GPTNeoXConfig, LlamaConfig, MistralConfig, OPTConfig, Qwen2Config)

import hcache_deepspeed_b200 as ds  # noqa: E402
from hcache_deepspeed_b200.module_inject.containers import (POLICIES, InjectedLayer, LayerTensors, build_fused_layer,  # noqa: E402
                                                            inject_fused_layers)

CAUSAL = {
    "gpt2": lambda: GPT2Config(n_embd=32, n_head=4, n_layer=2, vocab_size=50, n_positions=64),
    "gpt2_scaled": lambda: GPT2Config(n_embd=32, n_head=4, n_layer=2, vocab_size=50, n_positions=64,
                                      scale_attn_by_inverse_layer_idx=True),
    "gptneox": lambda: GPTNeoXConfig(hidden_size=32, num_attention_heads=4, num_hidden_layers=2, vocab_size=50,
                                     intermediate_size=64, rotary_pct=0.5, max_position_embeddings=64),
    "gptneox_sequential": lambda: GPTNeoXConfig(hidden_size=32, num_attention_heads=4, num_hidden_layers=2, vocab_size=50,
                                                intermediate_size=64, use_parallel_residual=False, max_position_embeddings=64),
    "gptj": lambda: GPTJConfig(n_embd=32, n_head=4, n_layer=2, vocab_size=50, rotary_dim=4, n_positions=64),
    "gptneo": lambda: GPTNeoConfig(hidden_size=32, num_heads=4, num_layers=2, vocab_size=50, attention_types=[[["global"], 2]],
                                   max_position_embeddings=64),
    "bloom": lambda: BloomConfig(hidden_size=32, n_head=4, n_layer=2, vocab_size=50),
    "llama_gqa": lambda: LlamaConfig(hidden_size=32, num_attention_heads=4, num_key_value_heads=2, num_hidden_layers=2, vocab_size=50,
                                     intermediate_size=64, max_position_embeddings=64),
    "qwen2_qkv_bias": lambda: Qwen2Config(hidden_size=32, num_attention_heads=4, num_key_value_heads=2, num_hidden_layers=2,
                                          vocab_size=50, intermediate_size=64, max_position_embeddings=64),
    "mistral": lambda: MistralConfig(hidden_size=32, num_attention_heads=4, num_key_value_heads=4, num_hidden_layers=2, vocab_size=50,
                                     intermediate_size=64, max_position_embeddings=64, sliding_window=None),
}
ENCODERS = {
    "bert": lambda: BertConfig(hidden_size=32, num_attention_heads=4, num_hidden_layers=2, vocab_size=50, intermediate_size=64),
    "distil_bert": lambda: DistilBertConfig(dim=32, n_heads=4, n_layers=2, vocab_size=50, hidden_dim=64),
}


@pytest.mark.parametrize("name", sorted(CAUSAL))
def test_injected_causal_matches_hf_and_generates(name):
    torch.manual_seed(0)
    cfg = CAUSAL[name]()
    model = AutoModelForCausalLM.from_config(cfg).eval().float()
    model.generation_config.pad_token_id = 0
    ids = torch.randint(1, 50, (2, 9))
    with torch.no_grad():
        want = model(ids).logits
        want_gen = model.generate(ids, max_new_tokens=5, do_sample=False)
    injected = copy.deepcopy(model)
    assert inject_fused_layers(injected) == 2
    layers = [m for m in injected.modules() if isinstance(m, InjectedLayer)]
    assert len(layers) == 2
    steps = []
    layers[0].register_forward_pre_hook(lambda mod, args: steps.append(args[0].shape[1]))
    with torch.no_grad():
        torch.testing.assert_close(injected(ids).logits, want, rtol=1e-4, atol=1e-5)
        steps.clear()
        got_gen = injected.generate(ids, max_new_tokens=5, do_sample=False)
    assert torch.equal(got_gen, want_gen)
    assert steps == [9, 1, 1, 1, 1]                 # prompt once, then one token per step out of the fused KV cache


@pytest.mark.parametrize("name", sorted(ENCODERS))
def test_injected_encoder_matches_hf(name):
    torch.manual_seed(0)
    model = AutoModel.from_config(ENCODERS[name]()).eval().float()
    ids = torch.randint(1, 50, (2, 9))
    mask = torch.ones(2, 9, dtype=torch.long)
    mask[1, 6:] = 0
    with torch.no_grad():
        want = model(ids, attention_mask=mask).last_hidden_state
        assert inject_fused_layers(model) == 2
        got = model(ids, attention_mask=mask).last_hidden_state
    torch.testing.assert_close(got[0], want[0], rtol=1e-4, atol=1e-5)
    torch.testing.assert_close(got[1, :6], want[1, :6], rtol=1e-4, atol=1e-5)    # padded positions are don't-care


def test_init_inference_kernel_inject_uses_containers():
    torch.manual_seed(0)
    ids = torch.randint(1, 50, (1, 6))
    # families / variants outside the ragged executor: block-by-block injection through the container policies
    for name in ("gpt2_scaled", ):
        model = AutoModelForCausalLM.from_config(CAUSAL[name]()).eval().float()
        with torch.no_grad():
            want = model(ids).logits
        engine = ds.init_inference(model, dtype=torch.float32, replace_with_kernel_inject=True)
        assert engine.fused_layers == 2 and not engine.injected, name
        with torch.no_grad():
            torch.testing.assert_close(engine(ids).logits, want, rtol=1e-4, atol=1e-5)
    # GPT-2 / GPT-J / GPT-NeoX / Bloom are lifted whole onto the ragged executor (paged KV, the sm_100a kernels on a GPU)
    for name in ("gpt2", "gptj", "gptneox", "gptneox_sequential", "bloom", "gptneo"):
        model = AutoModelForCausalLM.from_config(CAUSAL[name]()).eval().float()
        with torch.no_grad():
            want = model(ids).logits
        engine = ds.init_inference(model, dtype=torch.float32, replace_with_kernel_inject=True)
        assert engine.injected and engine.fused_layers == 0 and engine.ragged_engine is not None, name
        with torch.no_grad():
            torch.testing.assert_close(engine(ids).logits, want, rtol=2e-4, atol=2e-5)
    with pytest.raises(ValueError, match="kernel injection is not available"):
        ds.init_inference(torch.nn.Linear(2, 2), dtype=torch.float32, replace_with_kernel_inject=True)


def test_tensor_parallel_slices_sum_to_the_full_layer():
    """Two TP slices of one block (column split QKV / MLP-1, row split out-proj / MLP-2), outputs summed by hand."""
    torch.manual_seed(0)
    cfg = CAUSAL["gpt2"]()
    model = AutoModelForCausalLM.from_config(cfg).eval().float()
    block = model.transformer.h[0]
    x = torch.randn(2, 5, 32)
    lt = POLICIES["GPT2Block"](block, cfg)
    assert isinstance(lt, LayerTensors)
    full = build_fused_layer(lt, 1, torch.float32)
    with torch.no_grad():
        want = full(x)
        parts = []
        for r in range(2):
            lay = build_fused_layer(lt, 1, torch.float32, mp_size=2, mp_rank=r)
            assert lay.attention.attn_qkvw.shape == (32, 48) and lay.mlp.output_w.shape == (64, 32)
            a = lay.attention
            attn_out, _, _, _, _ = a(x, None, lay.norm_w, lay.norm_b, is_prompt=True)
            parts.append((lay, attn_out))
        attn = parts[0][1] + parts[1][1]                       # what the all-reduce produces
        h = x + attn + parts[0][0].attention.attn_ob           # rank 0 holds the row-parallel bias
        mlp = sum(lay.mlp.ops.mlp_gemm(torch.zeros_like(h), h, None, lay.mlp.inter_w, lay.mlp.output_w, lay.mlp.inter_b,
                                       lay.mlp.attn_nw, lay.mlp.attn_nb, 1e-5, True, True, act="gelu_new")[0] for lay, _ in parts)
        got = h + mlp + parts[0][0].mlp.output_b
    torch.testing.assert_close(got, want, rtol=1e-4, atol=1e-5)


def test_checkpoint_json_into_meta_model_then_inject(tmp_path):
    """init_inference(meta model, checkpoint=json): weights come from sharded files, blocks are injected afterwards;
    save_mp_checkpoint_path writes a reloadable per-rank checkpoint (ref: inference/engine.py:366-475)."""
    import json
    torch.manual_seed(0)
    cfg = CAUSAL["gpt2_scaled"]()          # a variant that stays on the block-injection path
    src = AutoModelForCausalLM.from_config(cfg).eval().float()
    ids = torch.randint(1, 50, (1, 6))
    with torch.no_grad():
        want = src(ids).logits
    sd = {k: v for k, v in src.state_dict().items() if k != "lm_head.weight"}        # tied to wte
    keys = sorted(sd)
    torch.save({k: sd[k] for k in keys[::2]}, tmp_path / "part0.pt")
    torch.save({k: sd[k] for k in keys[1::2]}, tmp_path / "part1.pt")
    (tmp_path / "ckpt.json").write_text(json.dumps({"type": "ds_model", "checkpoints": ["part0.pt", "part1.pt"], "version": 1.0}))
    with torch.device("meta"):
        meta_model = AutoModelForCausalLM.from_config(cfg).eval()
    assert next(meta_model.parameters()).is_meta
    out_dir = tmp_path / "mp"
    engine = ds.init_inference(meta_model, dtype=torch.float32, replace_with_kernel_inject=True, checkpoint=str(tmp_path / "ckpt.json"),
                               base_dir=str(tmp_path), save_mp_checkpoint_path=str(out_dir))
    assert engine.fused_layers == 2
    with torch.no_grad():
        torch.testing.assert_close(engine(ids).logits, want, rtol=1e-4, atol=1e-5)
    saved = json.loads((out_dir / "ds_inference_config.json").read_text())
    assert saved["checkpoints"]["tp"] == ["tp_00_00.pt"] and (out_dir / "tp_00_00.pt").exists()

    # a plain folder of weight files loads too
    plain = AutoModelForCausalLM.from_config(cfg).eval().float()
    folder = tmp_path / "w"
    folder.mkdir()
    torch.save(sd, folder / "pytorch_model.bin")
    eng2 = ds.init_inference(plain, dtype=torch.float32, checkpoint=str(folder))
    with torch.no_grad():
        torch.testing.assert_close(eng2(ids).logits, want, rtol=1e-4, atol=1e-5)


def test_megatron_gpt_policy_against_a_plain_block():
    """A stand-in with Megatron's ParallelTransformerLayer attribute layout (per-head [h, 3, d] fused QKV)."""
    from types import SimpleNamespace

    from torch import nn
    torch.manual_seed(0)
    H, heads, d = 32, 4, 8

    class ParallelTransformerLayer(nn.Module):
        def __init__(self):
            super().__init__()
            self.input_layernorm, self.post_attention_layernorm = nn.LayerNorm(H), nn.LayerNorm(H)
            self.attention = nn.Module()
            self.attention.query_key_value, self.attention.dense = nn.Linear(H, 3 * H), nn.Linear(H, H)
            self.mlp = nn.Module()
            self.mlp.dense_h_to_4h, self.mlp.dense_4h_to_h = nn.Linear(H, 4 * H), nn.Linear(4 * H, H)

        def forward(self, x):
            b, s, _ = x.shape
            qkv = self.attention.query_key_value(self.input_layernorm(x)).view(b, s, heads, 3, d)
            q, k, v = (qkv[:, :, :, i].transpose(1, 2) for i in range(3))
            ctx = torch.nn.functional.scaled_dot_product_attention(q, k, v, is_causal=True).transpose(1, 2).reshape(b, s, H)
            x = x + self.attention.dense(ctx)
            return x + self.mlp.dense_4h_to_h(torch.nn.functional.gelu(self.mlp.dense_h_to_4h(self.post_attention_layernorm(x))))

    block = ParallelTransformerLayer().eval()
    cfg = SimpleNamespace(hidden_size=H, num_attention_heads=heads, layernorm_epsilon=1e-5)
    lt = POLICIES["ParallelTransformerLayer"](block, cfg)
    fused = build_fused_layer(lt, 1, torch.float32)
    x = torch.randn(2, 6, H)
    with torch.no_grad():
        got = fused(x)
        got = got[0] if isinstance(got, tuple) else got
        torch.testing.assert_close(got, block(x), atol=1e-5, rtol=1e-5)
