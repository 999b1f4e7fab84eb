"""v2 module registry (reference: tests/unit/inference/v2/modules/*): configs, registries, implementations on the CPU ops."""
import pytest
import torch


def test_registry_and_implementations():
    from hcache_deepspeed_b200.inference.v2.modules import (ConfigBundle, DSEmbeddingsConfig, DSLinearConfig, DSLinearRegistry,
                                                            DSMoEConfig, DSNormConfig, DSUnembedConfig, NormTypeEnum,
                                                            instantiate_embed, instantiate_linear, instantiate_moe,
                                                            instantiate_post_norm, instantiate_pre_norm, instantiate_unembed)
    torch.manual_seed(0)
    f32 = dict(input_dtype=torch.float32, output_dtype=torch.float32)
    lin = instantiate_linear(DSLinearConfig(in_channels=16, out_channels=8, activation="siglu", **f32))
    w, x = torch.randn(16, 16), torch.randn(3, 16)
    want = torch.nn.functional.silu(x @ w[:8].T) * (x @ w[8:].T)
    torch.testing.assert_close(lin(x, lin.transform_param(w)), want, rtol=1e-5, atol=1e-5)
    plain = instantiate_linear(DSLinearConfig(in_channels=16, out_channels=16, activation="gelu", **f32))
    b = torch.randn(16)
    torch.testing.assert_close(plain(x, w, b), torch.nn.functional.gelu(x @ w.T + b), rtol=1e-4, atol=1e-4)
    pre = instantiate_pre_norm(DSNormConfig(type=NormTypeEnum.RMSNorm, channels=16, eps=1e-5))
    res, h, g = torch.randn(3, 16), torch.randn(3, 16), torch.rand(16) + 0.5
    s = res + h
    r, o = pre(res, h, g)
    torch.testing.assert_close(r, s)
    torch.testing.assert_close(o, s * torch.rsqrt(s.pow(2).mean(-1, keepdim=True) + 1e-5) * g, rtol=1e-5, atol=1e-5)
    assert type(instantiate_pre_norm(DSNormConfig(type=NormTypeEnum.LayerNorm, channels=16))).name() == "cuda_pre_ln"
    post = instantiate_post_norm(DSNormConfig(type=NormTypeEnum.LayerNorm, channels=16))
    res2 = torch.randn(3, 16)
    want_ln = torch.nn.functional.layer_norm(res2 + h, (16, ), g, b, 1e-5)
    torch.testing.assert_close(post(res2, h, g, b), want_ln, rtol=1e-5, atol=1e-5)
    emb = instantiate_embed(DSEmbeddingsConfig(embedding_dim=16, positional_embedding=True, positional_offset=2,
                                               residual_dtype=torch.float32))
    W, P = torch.randn(10, 16), torch.randn(12, 16)
    ids, pos = torch.tensor([1, 5, 9], dtype=torch.int32), torch.tensor([0, 1, 2], dtype=torch.int32)
    torch.testing.assert_close(emb(ids, W, P, pos), W[ids.long()] + P[pos.long() + 2])
    un = instantiate_unembed(DSUnembedConfig(vocab_size=10, norm_type=NormTypeEnum.RMSNorm, dtype=torch.float32, model_dim=16))
    hid = torch.randn(7, 16)
    last = torch.tensor([2, 6], dtype=torch.int32)
    pk = hid[last.long()]
    want_lg = (pk * torch.rsqrt(pk.pow(2).mean(-1, keepdim=True) + 1e-5) * g) @ W.T
    torch.testing.assert_close(un(hid, W, last, gamma=g), want_lg, rtol=1e-4, atol=1e-4)
    moe = instantiate_moe(DSMoEConfig(model_dim=16, intermediate_features=8, n_experts=4, top_k=2, activation="relu",
                                      normalize_scores=True, **f32))
    assert moe(x, torch.randn(4, 16), torch.randn(4, 8, 16), torch.randn(4, 16, 8)).shape == (3, 16)
    with pytest.raises(KeyError):
        DSLinearRegistry.instantiate_config(ConfigBundle("nope", DSLinearConfig(in_channels=8, out_channels=8)))
    with pytest.raises(ValueError):
        DSLinearRegistry.instantiate_config(ConfigBundle("blas_fp_linear", DSLinearConfig(in_channels=8, out_channels=8,
                                                                                          activation="tanh")))


def test_fp8_linear_module_cpu_path():
    """wf8af8 picks the fp8 linear; its CPU restatement (same quantisation rule as the CUDA path) tracks the fp linear."""
    from types import SimpleNamespace
    from hcache_deepspeed_b200.inference.v2.modules import DSLinearConfig, instantiate_linear
    torch.manual_seed(0)
    cfg = DSLinearConfig(max_tokens=64, in_channels=128, out_channels=96, activation="identity",
                         input_dtype=torch.bfloat16, output_dtype=torch.bfloat16)
    eng = SimpleNamespace(quantization=SimpleNamespace(quantization_mode="wf8af8"))
    lin = instantiate_linear(cfg, eng)
    assert lin.name() == "fp8_tcgen05_linear"
    w = (torch.randn(96, 128) * 128 ** -0.5).to(torch.bfloat16)
    x = torch.randn(10, 128).to(torch.bfloat16)
    wq = lin.transform_param(w)
    assert wq.dtype == torch.uint8 and wq.scales.shape == (96, )
    y = lin(x, wq)
    want = x.float() @ w.float().t()
    assert (y.float() - want).abs().mean() < 0.05 * want.abs().mean() + 0.02
