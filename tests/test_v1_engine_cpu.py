"""``init_inference`` (v1 API) on CPU — BASELINE.json config 1: an OPT-shaped model, one rank, hidden states saved for a
512-token context, KV rebuilt from them, logits compared with recompute. Also the config rules of
``deepspeed/__init__.py:341-362`` and ``inference/config.py`` (conflicts, deprecated aliases), generate(), beams."""
import pytest
import torch

pytestmark = pytest.mark.timeout(600)


def _tiny_opt():
    from transformers import OPTConfig, OPTForCausalLM
    torch.manual_seed(0)
    cfg = OPTConfig(vocab_size=512, hidden_size=64, num_hidden_layers=3, ffn_dim=128, num_attention_heads=4,
                    max_position_embeddings=1024, word_embed_proj_dim=64, do_layer_norm_before=True, dropout=0.0)
    return OPTForCausalLM(cfg).eval()


def _tiny_llama():
    from transformers import LlamaConfig, LlamaForCausalLM
    torch.manual_seed(0)
    cfg = LlamaConfig(vocab_size=300, hidden_size=64, intermediate_size=128, num_hidden_layers=2, num_attention_heads=4,
                      num_key_value_heads=2, max_position_embeddings=512)
    return LlamaForCausalLM(cfg).eval()


def test_config_merge_rules(tmp_path):
    import json

    import hcache_deepspeed_b200 as ds
    m = _tiny_llama()
    with pytest.raises(ValueError):
        ds.init_inference(m, config={"dtype": "fp32"}, dtype="bf16")
    p = tmp_path / "cfg.json"
    p.write_text(json.dumps({"dtype": "fp32", "mp_size": 1}))
    eng = ds.init_inference(m, config=str(p))      # deprecated mp_size alias accepted
    assert eng._config.tensor_parallel.tp_size == 1 and eng._config.dtype == torch.float32
    with pytest.raises(ValueError):
        ds.init_inference(m, config=3)
    assert "replace_with_kernel_inject" in ds.default_inference_config()


def test_uninjected_engine_is_a_passthrough():
    import hcache_deepspeed_b200 as ds
    m = _tiny_llama()
    eng = ds.init_inference(m, dtype=torch.float32)
    ids = torch.randint(0, 300, (2, 9))
    torch.testing.assert_close(eng(input_ids=ids).logits, m(input_ids=ids).logits)
    eng.profile_model_time()
    eng(input_ids=ids)
    assert len(eng.model_times()) == 1
    with pytest.raises(NotImplementedError):
        eng.generate(input_ids=ids, num_beams=2)
    with pytest.raises(KeyError):
        eng.hidden_states(0)
    with pytest.raises(RuntimeError):
        eng.flush(0)


@pytest.mark.parametrize("make", [_tiny_opt, _tiny_llama])
def test_injected_engine_matches_hf_and_restores_from_hidden_states(make):
    import hcache_deepspeed_b200 as ds
    hf = make()
    vocab = hf.config.vocab_size
    eng = ds.init_inference(hf, dtype=torch.float32, replace_with_kernel_inject=True,
                            hcache={"enabled": True, "max_context": 1024, "max_batch_tokens": 768, "kv_memory_bytes": 64 << 20})
    g = torch.Generator().manual_seed(1)
    ctx = torch.randint(0, vocab, (1, 512), generator=g)
    turn = torch.randint(0, vocab, (1, 16), generator=g)
    with torch.no_grad():
        want_ctx = hf(input_ids=ctx).logits
        want_turn = hf(input_ids=torch.cat([ctx, turn], dim=1)).logits[:, 512:]
    out = eng(input_ids=ctx, session_ids=[7])
    torch.testing.assert_close(out.logits, want_ctx, rtol=2e-3, atol=2e-3)
    lat = eng.hidden_states(7)
    assert lat.shape == (hf.config.num_hidden_layers, 512, hf.config.hidden_size)
    live = eng(input_ids=turn, session_ids=[7]).logits            # continuation over the live KV cache
    torch.testing.assert_close(live, want_turn, rtol=2e-3, atol=2e-3)
    # end of turn: drop the KV cache, come back later, rebuild it from the saved hidden states instead of recomputing
    eng.flush(7)
    eng.restore_hidden_states(7, ctx, lat)
    restored = eng(input_ids=turn, session_ids=[7]).logits
    assert float((restored - want_turn).abs().max()) < 1e-3          # restore-vs-recompute (BASELINE config 1)
    torch.testing.assert_close(restored, live, rtol=1e-4, atol=1e-4)


def test_injected_generate_greedy_matches_hf():
    import hcache_deepspeed_b200 as ds
    hf = _tiny_llama()
    eng = ds.init_inference(hf, dtype=torch.float32, replace_with_kernel_inject=True,
                            hcache={"max_context": 256, "max_batch_tokens": 256, "kv_memory_bytes": 16 << 20})
    ids = torch.randint(0, 300, (2, 12), generator=torch.Generator().manual_seed(3))
    got = eng.generate(input_ids=ids, max_new_tokens=8, do_sample=False)
    want = hf.generate(input_ids=ids, max_new_tokens=8, do_sample=False, pad_token_id=0)
    assert torch.equal(got[:, :20], want[:, :20])


def test_v1_fused_layer_incremental_decode_matches_full_forward():
    """DeepSpeedTransformerInference with its contiguous KV cache: prompt + token-by-token decode equals one full pass."""
    from hcache_deepspeed_b200.model_implementations import DeepSpeedTransformerInference
    from hcache_deepspeed_b200.ops.transformer.inference import DeepSpeedInferenceConfig
    DeepSpeedTransformerInference.release()
    torch.manual_seed(0)
    cfg = DeepSpeedInferenceConfig(hidden_size=32, heads=4, num_hidden_layers=1, dtype=torch.float32, pre_layer_norm=True,
                                   rotary_dim=8, rotate_half=True, rotate_every_two=False, mlp_act_func_type="gelu",
                                   max_out_tokens=32, layer_norm_eps=1e-5)
    layer = DeepSpeedTransformerInference(cfg).eval()
    for p in layer.parameters():
        if p.dim() > 1:
            p.data.normal_(0, 0.1)
    x = torch.randn(2, 7, 32)
    full = layer(x)[0]
    DeepSpeedTransformerInference.reset_cache()
    out_p, present = layer(x[:, :5], use_cache=True)
    assert present[0].shape == (2, 4, 5, 8)
    outs = [out_p]
    for t in (5, 6):
        outs.append(layer(x[:, t:t + 1], layer_past=present, use_cache=True)[0])
    torch.testing.assert_close(torch.cat(outs, dim=1), full, rtol=1e-4, atol=1e-5)
    # reference formula for the first token (no attention context beyond itself)
    ops = layer.mlp.ops
    q_k = ops.apply_rotary_pos_emb(torch.randn(1, 2, 3, 8), torch.randn(1, 2, 3, 8), 8, 0, 2, True)
    assert q_k[0].shape == (1, 2, 3, 8)
    DeepSpeedTransformerInference.release()


def test_moe_inference_layer_matches_dense_evaluation():
    """DeepSpeedMoEInference: top-k routed experts (+ residual PR-MoE blend) equal evaluating every expert densely."""
    import torch.nn.functional as F
    from hcache_deepspeed_b200.ops.transformer.inference.moe_inference import DeepSpeedMoEInference, DeepSpeedMoEInferenceConfig
    torch.manual_seed(0)
    for mlp_type, k in (("standard", 1), ("standard", 2), ("residual", 2)):
        cfg = DeepSpeedMoEInferenceConfig(hidden_size=32, intermediate_size=64, heads=4, num_hidden_layers=1, layer_norm_eps=1e-5,
                                          moe_experts=4, global_experts=4, k=k, mlp_type=mlp_type, return_tuple=False)
        layer = DeepSpeedMoEInference(cfg)
        for p in layer.parameters():
            if p.dim() > 1:
                torch.nn.init.normal_(p, std=0.2)
        x = torch.randn(2, 6, 32)
        out = layer(x)
        # oracle
        attn_out, _, _, _, _ = layer.attention(x, None, layer.norm_w, layer.norm_b, is_prompt=True)
        res = x + attn_out + layer.attention.attn_ob
        flat = F.layer_norm(res, (32, ), layer.attn_nw, layer.attn_nb, 1e-5).reshape(-1, 32)
        probs = torch.softmax(flat @ layer.gate_w.t(), dim=-1)
        top_p, top_e = probs.topk(k, dim=-1)
        if k > 1:
            top_p = top_p / top_p.sum(-1, keepdim=True)
        dense = torch.stack([e(flat) for e in layer.mlp], dim=1)                      # [tokens, experts, hidden]
        weights = torch.zeros_like(probs).scatter_(1, top_e, top_p)
        moe = (dense * weights[:, :, None]).sum(dim=1)
        if mlp_type == "residual":
            coef = torch.softmax(flat @ layer.res_coef, dim=-1)
            moe = layer.res_mlp(flat) * coef[:, 0:1] + moe * coef[:, 1:2]
        torch.testing.assert_close(out, res + moe.view_as(res), rtol=1e-4, atol=1e-5)
    y, (kk, vv) = layer(x, use_cache=True)
    assert kk.shape[2] == 6


def test_v1_op_binding_objects():
    import math

    import torch

    from hcache_deepspeed_b200.ops.transformer.inference.config import DeepSpeedInferenceConfig
    from hcache_deepspeed_b200.ops.transformer.inference.op_binding import (BiasGeluOp, EinsumSecSmEcmOp, GatedActivationOp,
                                                                           LayerNormOp, LinearOp, MLPGemmOp, QKVGemmOp,
                                                                           RMSNormOp, SoftmaxContextOp, SoftmaxOp, WorkspaceOp)
    torch.manual_seed(0)
    cfg = DeepSpeedInferenceConfig(hidden_size=16, heads=2, intermediate_size=32, num_hidden_layers=1, dtype=torch.float32,
                                   pre_layer_norm=True, triangular_masking=True, max_out_tokens=32, rotate_every_two=False)
    x, g, b = torch.randn(2, 5, 16), torch.rand(16) + 0.5, torch.randn(16)
    torch.testing.assert_close(LayerNormOp(cfg)(x, g, b, 1e-5), torch.nn.functional.layer_norm(x, (16,), g, b, 1e-5))
    torch.testing.assert_close(RMSNormOp(cfg)(x, g, 1e-6), x * torch.rsqrt(x.pow(2).mean(-1, keepdim=True) + 1e-6) * g)
    torch.testing.assert_close(BiasGeluOp(cfg)(x, b), torch.nn.functional.gelu(x + b))
    h = torch.randn(2, 5, 32)
    a_, g_ = h.chunk(2, -1)
    torch.testing.assert_close(GatedActivationOp(cfg)(h, None, "silu"), a_ * torch.nn.functional.silu(g_))
    wqkv, bqkv = torch.randn(16, 48) / 4, torch.randn(48)
    qkv = QKVGemmOp(cfg)(x, wqkv, bqkv, g, b)
    qkv = qkv[0] if isinstance(qkv, tuple) else qkv
    torch.testing.assert_close(qkv, torch.nn.functional.layer_norm(x, (16,), g, b, cfg.layer_norm_eps) @ wqkv + bqkv, atol=1e-5, rtol=1e-5)
    ctx, kc, vc = SoftmaxContextOp(cfg)(qkv, None, 2, 2, 8 ** -0.25, False, 0, 1, None)   # the factor is applied to q and to k
    q, k, v = (t.view(2, 5, 2, 8).transpose(1, 2) for t in qkv.split(16, -1))
    want = torch.nn.functional.scaled_dot_product_attention(q, k, v, is_causal=True).transpose(1, 2).reshape(2, 5, 16)
    torch.testing.assert_close(ctx, want, atol=1e-5, rtol=1e-5)
    scores = torch.randn(2, 2, 5, 5)
    p = SoftmaxOp(cfg)(scores, None, None, True, False, False, 1, False, 1.0, 0)
    assert torch.allclose(p.sum(-1), torch.ones(2, 2, 5)) and float(p[0, 0, 0, 1:].sum()) == 0.0
    w = torch.randn(16, 8)
    torch.testing.assert_close(LinearOp(cfg)(x, w, torch.ones(8), add_bias=True), x @ w + 1)
    assert EinsumSecSmEcmOp(cfg)(torch.randn(3, 4, 5), torch.randn(3, 6)).shape == (4, 5, 6)
    out = MLPGemmOp(cfg)(x, x, torch.randn(16, 32) / 4, torch.randn(32, 16) / 4, None, torch.zeros(32), g, b)
    assert out[0].shape == (2, 5, 16)
    ws = WorkspaceOp(cfg)
    ws.reset_cache()
    assert ws.is_allocated() and ws.retake_workspace()


def test_diffusers_transformer_block_matches_equivalent_module():
    """A stand-in with the attribute layout of diffusers' BasicTransformerBlock -> the fused block gives the same output."""
    import torch
    from torch import nn

    from hcache_deepspeed_b200.ops.transformer.inference import (DeepSpeedDiffusersTransformerBlock,
                                                                 Diffusers2DTransformerConfig)
    from hcache_deepspeed_b200.ops.transformer.inference.bias_add import nhwc_bias_add
    from hcache_deepspeed_b200.ops.transformer.inference.ds_attention import BloomSelfAttention, DeepSpeedSelfAttention
    assert BloomSelfAttention is DeepSpeedSelfAttention and callable(nhwc_bias_add)

    class Attn(nn.Module):
        def __init__(self, dim, ctx_dim, heads):
            super().__init__()
            self.heads = heads
            self.to_q, self.to_k, self.to_v = nn.Linear(dim, dim, bias=False), nn.Linear(ctx_dim, dim, bias=False), nn.Linear(ctx_dim, dim, bias=False)
            self.to_out = nn.ModuleList([nn.Linear(dim, dim), nn.Dropout(0.0)])

        def forward(self, x, context=None):
            c = x if context is None else context
            sp = lambda t: t.view(t.shape[0], t.shape[1], self.heads, -1).transpose(1, 2)  # noqa: E731
            o = torch.nn.functional.scaled_dot_product_attention(sp(self.to_q(x)), sp(self.to_k(c)), sp(self.to_v(c)))
            return self.to_out[0](o.transpose(1, 2).reshape(x.shape))

    class GEGLU(nn.Module):
        def __init__(self, dim, inner):
            super().__init__()
            self.proj = nn.Linear(dim, 2 * inner)

        def forward(self, x):
            a, g = self.proj(x).chunk(2, -1)
            return a * torch.nn.functional.gelu(g)

    class Block(nn.Module):
        def __init__(self, dim=32, ctx_dim=24, heads=4):
            super().__init__()
            self.norm1, self.norm2, self.norm3 = nn.LayerNorm(dim), nn.LayerNorm(dim), nn.LayerNorm(dim)
            self.attn1, self.attn2 = Attn(dim, dim, heads), Attn(dim, ctx_dim, heads)
            self.ff = nn.Module()
            self.ff.net = nn.ModuleList([GEGLU(dim, 64), nn.Dropout(0.0), nn.Linear(64, dim)])

        def forward(self, x, context):
            x = x + self.attn1(self.norm1(x))
            x = x + self.attn2(self.norm2(x), context)
            return x + self.ff.net[2](self.ff.net[0](self.norm3(x)))

    torch.manual_seed(0)
    blk = Block().eval()
    x, ctx = torch.randn(2, 10, 32), torch.randn(2, 7, 24)
    fused = DeepSpeedDiffusersTransformerBlock(blk, Diffusers2DTransformerConfig())
    with torch.no_grad():
        torch.testing.assert_close(fused(x, context=ctx), blk(x, ctx), atol=1e-5, rtol=1e-5)
        torch.testing.assert_close(fused(x, encoder_hidden_states=ctx), blk(x, ctx), atol=1e-5, rtol=1e-5)


def test_int8_dtype_selects_8bit_serving():
    """``init_inference(dtype=torch.int8, replace_with_kernel_inject=True)``: the reference's 8-bit serving request maps to
    the executor's 8-bit tensor-core mode (E4M3 weights + per-token activations); logits track the fp32 model."""
    transformers = pytest.importorskip("transformers")
    import copy
    import hcache_deepspeed_b200 as ds
    from hcache_deepspeed_b200.ops.fp8 import Fp8Weight
    torch.manual_seed(0)
    model = transformers.LlamaForCausalLM(transformers.LlamaConfig(
        hidden_size=128, num_hidden_layers=2, num_attention_heads=4, vocab_size=120, max_position_embeddings=128,
        intermediate_size=256, num_key_value_heads=2)).eval().float()
    ids = torch.randint(1, 120, (1, 20))
    with torch.no_grad():
        want = model(ids).logits
    eng = ds.init_inference(copy.deepcopy(model), dtype=torch.int8, replace_with_kernel_inject=True)
    got = eng(ids).logits
    assert isinstance(eng.ragged_engine.model().w.layers[0].w_qkv, Fp8Weight)
    assert torch.nn.functional.cosine_similarity(got.flatten(), want.flatten(), dim=0) > 0.99


def test_v1_engine_model_fixups_and_compile_flag():
    """``build_alibi_tensor`` / ``build_attn_bias`` / ``remove_mask_prepare_for_bloom`` / ``compile`` of the v1 engine
    (ref: inference/engine.py:207-228, 608-625), on stand-in modules with the attribute layout of BLOOM, MPT and Baichuan."""
    import types
    import hcache_deepspeed_b200 as ds
    from hcache_deepspeed_b200.module_inject.auto_tp_model_utils import build_bloom_alibi_tensor, slice_heads

    class Inner(torch.nn.Module):
        def __init__(self):
            super().__init__()
            self.fc = torch.nn.Linear(4, 4)

        def build_alibi_tensor(self, attention_mask, num_heads, dtype):
            raise AssertionError("replaced by the engine")

        def build_mpt_alibi_tensor(self, heads, seq):
            return torch.arange(heads * seq, dtype=torch.float32).reshape(1, heads, 1, seq)

        def _attn_bias(self, device, dtype, attention_mask=None, prefix_mask=None, sequence_id=None):
            return torch.arange(8 * 3, dtype=dtype).reshape(1, 8, 1, 3), attention_mask

        def _prepare_attn_mask(self, attention_mask, *a, **k):
            return attention_mask * 0

        def get_alibi_mask(self, tensor, seq):
            return torch.arange(8 * seq * seq, dtype=torch.float32).reshape(8, seq, seq)

    class Model(torch.nn.Module):
        def __init__(self):
            super().__init__()
            self.transformer = Inner()
            self.model = self.transformer

        def forward(self, x):
            return self.transformer.fc(x)

    eng = ds.init_inference(Model(), dtype=torch.float32)
    eng.build_alibi_tensor()
    eng.build_attn_bias()
    eng.remove_mask_prepare_for_bloom()
    tr = eng.module.transformer
    mask = torch.ones(2, 5, dtype=torch.long)
    torch.testing.assert_close(tr.build_alibi_tensor(mask, 8, torch.float32), build_bloom_alibi_tensor(mask, 8, torch.float32, 1, 0))
    assert tr.build_mpt_alibi_tensor(8, 3).shape == (1, 8, 1, 3)              # tp = 1: all heads
    bias, m = tr._attn_bias("cpu", torch.float32, attention_mask=mask)
    assert bias.shape == (1, 8, 1, 3) and m is mask
    assert torch.equal(tr._prepare_attn_mask(mask), mask)
    assert tr.get_alibi_mask(None, 4).shape == (8, 4, 4)
    # the slicing helper the patched builders use under tp > 1: uneven head counts follow the attention sharding
    t = torch.arange(7 * 2).reshape(1, 7, 2)
    parts = [slice_heads(t, 1, r, 3) for r in range(3)]
    assert [p.shape[1] for p in parts] == [3, 2, 2] and torch.equal(torch.cat(parts, 1), t)
    assert eng.is_compiled is False
    eng.compile()
    assert eng.is_compiled is True and eng._config.enable_cuda_graph is True
    x = torch.randn(2, 4)
    torch.testing.assert_close(eng(x), eng.module(x))                           # CPU: graph flag is inert
    import pytest
    with pytest.raises(ValueError):
        eng.load_model_with_checkpoint()
