"""inference.v2.kernels wrapper objects on host tensors (reference formulations with the native buffer contract)."""
import pytest
import torch

from hcache_deepspeed_b200.inference.v2.kernels import (AtomBuilder, BlasLibLinear, CUDABiasActivation, CUDAFPLN, CUDAFPPreLN,
                                                        CUDAGatedActivation, CUDARMSNorm, CUDARMSPreNorm, MixedGEMM, MoEGather,
                                                        MoEGEMM, MoEScatter, RaggedEmbeddingKernel, RaggedLogitsGather,
                                                        RaggedTopKGating)


def test_core_op_wrappers():
    torch.manual_seed(0)
    x, w = torch.randn(5, 32), torch.randn(16, 32)
    out = torch.empty(5, 16)
    torch.testing.assert_close(BlasLibLinear(torch.float32)(out, x, w), x @ w.t(), atol=1e-5, rtol=1e-5)
    g, b = torch.rand(32) + 0.5, torch.randn(32)
    y = torch.empty_like(x)
    CUDARMSNorm(32, torch.float32, 1e-6)(y, x, g)
    torch.testing.assert_close(y, x * torch.rsqrt(x.pow(2).mean(-1, keepdim=True) + 1e-6) * g, atol=1e-5, rtol=1e-5)
    CUDAFPLN(32, torch.float32, 1e-5)(y, x, g, b)
    torch.testing.assert_close(y, torch.nn.functional.layer_norm(x, (32,), g, b, 1e-5), atol=1e-5, rtol=1e-5)
    res, hid = torch.randn(5, 32), torch.randn(5, 32)
    zr, zh = torch.empty(5, 32), torch.empty(5, 32)
    CUDARMSPreNorm(32, torch.float32, 1e-6)(zr, zh, res, hid, g)
    s = res + hid
    torch.testing.assert_close(zr, s)
    torch.testing.assert_close(zh, s * torch.rsqrt(s.pow(2).mean(-1, keepdim=True) + 1e-6) * g, atol=1e-5, rtol=1e-5)
    CUDAFPPreLN(32, torch.float32, 1e-5)(zr, zh, res, hid, g, b)
    torch.testing.assert_close(zh, torch.nn.functional.layer_norm(s, (32,), g, b, 1e-5), atol=1e-5, rtol=1e-5)
    a = x.clone()
    CUDABiasActivation(32, torch.float32, "relu")(a, b)
    torch.testing.assert_close(a, torch.relu(x + b))
    go = torch.empty(5, 16)
    CUDAGatedActivation(32, torch.float32, "silu")(go, x)
    torch.testing.assert_close(go, torch.nn.functional.silu(x[:, 0::2]) * x[:, 1::2])
    with pytest.raises(ValueError):
        CUDARMSNorm(3, torch.float16)
    with pytest.raises(ValueError):
        BlasLibLinear(torch.int8)


def test_moe_pipeline_wrappers_match_dense_reference():
    torch.manual_seed(1)
    T, H, I, E, k = 9, 16, 24, 4, 2
    x = torch.randn(T, H)
    gate_w, w1, w2 = torch.randn(E, H), torch.randn(E, 2 * I, H) / 4, torch.randn(E, H, I) / 4
    logits = x @ gate_w.t()
    counts = torch.zeros(E, dtype=torch.int32)
    scores, ids, offs, slots = torch.empty(T, k), *(torch.empty(T, k, dtype=torch.int32) for _ in range(3))
    RaggedTopKGating(torch.float32)(counts, scores, ids, offs, logits)
    assert int(counts.sum()) == T * k and torch.allclose(scores.sum(-1), torch.ones(T))
    rows = int(((counts + 127) // 128 * 128).sum())
    xs, cum = torch.zeros(rows, H), torch.zeros(E + 1, dtype=torch.int32)
    MoEScatter(torch.float32)(xs, cum, slots, x, counts, ids, offs)
    assert int(counts.sum()) == 0 and cum[-1] == rows and len(set(slots.reshape(-1).tolist())) == T * k
    hs, ys = torch.zeros(rows, I), torch.zeros(rows, H)
    MoEGEMM(torch.float32, "silu", gated=True)(hs, xs, w1, cum)
    MoEGEMM(torch.float32)(ys, hs, w2, cum)
    out = torch.zeros(T, H)
    MoEGather(torch.float32)(out, ys, scores, slots)
    want = torch.zeros(T, H)
    probs = torch.softmax(logits, -1)
    tv, ti = probs.topk(k, -1)
    tv = tv / tv.sum(-1, keepdim=True)
    for t in range(T):
        for j in range(k):
            e = int(ti[t, j])
            h = x[t] @ w1[e].t()
            want[t] += tv[t, j] * ((torch.nn.functional.silu(h[0::2]) * h[1::2]) @ w2[e].t())
    torch.testing.assert_close(out, want, atol=1e-4, rtol=1e-4)


def test_ragged_embed_gather_atoms_and_mixed_gemm():
    torch.manual_seed(2)
    emb = torch.randn(20, 8)
    ids = torch.tensor([3, 7, 7, 1, 19], dtype=torch.int32)
    out = torch.empty(5, 8)
    torch.testing.assert_close(RaggedEmbeddingKernel(torch.float32)(out, ids, emb), emb[ids.long()])
    last = torch.empty(2, 8)
    RaggedLogitsGather(8, torch.float32)(last, out, torch.tensor([2, 4], dtype=torch.int32))
    torch.testing.assert_close(last, out[[2, 4]])
    items, n = AtomBuilder()(torch.zeros(2, 4, dtype=torch.int32), [3, 2], [10, 0])
    assert n == 2 and items.tolist() == [[0, 0, 3, 10], [1, 3, 2, 0]]
    w = torch.randn(6, 8)
    scale = w.abs().amax(1) / 127
    q8 = torch.round(w / scale[:, None]).to(torch.int8)
    x = torch.randn(4, 8)
    o = torch.empty(4, 6)
    MixedGEMM(torch.float32, num_bits=8)(o, x, q8, scale)
    torch.testing.assert_close(o, x @ (q8.float() * scale[:, None]).t(), atol=1e-5, rtol=1e-5)
    s4 = w.abs().amax(1) / 7
    q4 = (torch.round(w / s4[:, None]).clamp(-8, 7) + 8).to(torch.uint8)
    packed = (q4[:, 0::2] | (q4[:, 1::2] << 4)).to(torch.uint8)
    MixedGEMM(torch.float32, num_bits=4)(o, x, packed, s4)
    torch.testing.assert_close(o, x @ ((q4.float() - 8) * s4[:, None]).t(), atol=1e-5, rtol=1e-5)


def test_v2_reference_import_paths():
    from hcache_deepspeed_b200.inference.v2.model_implementations.flat_model_helpers import (flatten_inference_model,  # noqa: F401
                                                                                             make_param_filename)
    from hcache_deepspeed_b200.inference.v2.model_implementations.inference_policy_base import ContainerMap
    from hcache_deepspeed_b200.inference.v2.model_implementations.inference_transformer_base import DSTransformerModelBase  # noqa: F401
    from hcache_deepspeed_b200.inference.v2.model_implementations.sharding import ShardingType, shard_mlp_1_param
    from hcache_deepspeed_b200.inference.v2.modules.interfaces import DSLinearBase, DSLinearRegistry  # noqa: F401
    from hcache_deepspeed_b200.inference.v2.modules.module_registry import DSModuleRegistryBase
    assert issubclass(DSLinearRegistry, DSModuleRegistryBase) and ShardingType.OUTER_DIMENSION.value == 0
    w = torch.arange(128.0 * 4).view(128, 4)
    assert shard_mlp_1_param(w, 1, 2).shape[0] == 64 and make_param_filename("/x", 0, 2).startswith("/x")

    class C:
        def __init__(self):
            self.got = {}

        def set_dependency(self, n, t):
            self.got[n] = t

    m, layers, nt = ContainerMap(), [C(), C()], C()
    m.set_transformer_params(["model.layers"], layers)
    m.set_non_transformer_params(nt)
    m.set_unmapped_params(["rotary"])
    m.map_param("model.layers.1.self_attn.q_proj.weight", 1)
    m.map_param("model.embed_tokens.weight", 2)
    m.map_param("rotary.inv_freq", 3)
    assert layers[1].got == {"self_attn.q_proj.weight": 1} and nt.got == {"model.embed_tokens.weight": 2} and layers[0].got == {}
