"""Declarative parameters / layer containers (reference tests: tests/unit/inference/v2/model_implementations/parameters/*,
.../sharding/*)."""
import pytest
import torch

from hcache_deepspeed_b200.inference.v2.inference_parameter import InferenceParameter
from hcache_deepspeed_b200.inference.v2.model_implementations import common_parameters as cp
from hcache_deepspeed_b200.inference.v2.model_implementations.inference_model_base import DSTransformerModelBase
from hcache_deepspeed_b200.inference.v2.model_implementations.layer_container_base import LayerContainer
from hcache_deepspeed_b200.inference.v2.model_implementations.parameter_base import ParameterBase, ParamList
from hcache_deepspeed_b200.models.arch import ArchSpec


class _Model(DSTransformerModelBase):
    def __init__(self, arch, rank=0, size=1):
        super().__init__(arch, None, arch=arch, dtype=torch.float32, device=torch.device("cpu"))
        self._r, self._s = rank, size

    tp_rank = property(lambda self: self._r)
    tp_size = property(lambda self: self._s)

    def get_kv_requirements(self, *a): ...
    def get_remaining_block_capacity(self, *a): ...
    def maybe_allocate_kv(self, *a): ...
    def kv_cache_config(self): ...
    def prepare_batch(self, *a): ...
    def forward(self, *a): ...


def _arch(**kw):
    base = dict(model_type="llama", n_layers=1, hidden=16, n_heads=4, n_kv_heads=2, head_dim=4, intermediate=256, vocab=32,
                n_experts=2, top_k=1)
    base.update(kw)
    return ArchSpec(**base)


class _Layer(LayerContainer):
    qkv_w: cp.UnfusedQKVParameter
    attn_out_w: cp.AttentionOutputParameter
    mlp_1_w: cp.GatedMLPParameter
    mlp_2_w: cp.MLP2Parameter
    norm: cp.NormParameter
    moe_1: cp.UnfusedMoEGatedMLPParameter
    PARAM_MAPPING = {
        "self_attn.q_proj.weight": "qkv_w.q_params", "self_attn.k_proj.weight": "qkv_w.k_params",
        "self_attn.v_proj.weight": "qkv_w.v_params", "self_attn.o_proj.weight": "attn_out_w.params",
        "mlp.gate_proj.weight": "mlp_1_w.gate_params", "mlp.up_proj.weight": "mlp_1_w.up_params",
        "mlp.down_proj.weight": "mlp_2_w.params", "input_layernorm.weight": "norm.params",
        "moe.experts.*.w1.weight": "moe_1.gating_experts", "moe.experts.*.w3.weight": "moe_1.up_experts",
    }


def _tensors():
    g = torch.Generator().manual_seed(0)
    r = lambda *s: torch.randn(*s, generator=g)
    t = {"self_attn.q_proj.weight": r(16, 16), "self_attn.k_proj.weight": r(8, 16), "self_attn.v_proj.weight": r(8, 16),
         "self_attn.o_proj.weight": r(16, 16), "mlp.gate_proj.weight": r(256, 16), "mlp.up_proj.weight": r(256, 16),
         "mlp.down_proj.weight": r(16, 256), "input_layernorm.weight": r(16)}
    for e in range(2):
        t[f"moe.experts.{e}.w1.weight"] = r(256, 16)
        t[f"moe.experts.{e}.w3.weight"] = r(256, 16)
    return t


def test_container_finalizes_as_dependencies_arrive():
    layer = _Layer(_Model(_arch()))
    t = _tensors()
    assert layer.n_params == 6 and not layer.is_populated
    names = list(t)
    for i, n in enumerate(reversed(names)):          # arrival order does not matter
        layer.set_dependency(n, t[n])
        if n == "self_attn.k_proj.weight":
            assert isinstance(layer.qkv_w, cp.UnfusedQKVParameter)     # q still missing
    assert layer.is_populated and layer.is_initialized or not torch.cuda.is_available()
    assert isinstance(layer.qkv_w, InferenceParameter)
    torch.testing.assert_close(torch.Tensor(layer.qkv_w), torch.cat([t["self_attn.q_proj.weight"], t["self_attn.k_proj.weight"],
                                                                     t["self_attn.v_proj.weight"]]))
    m1 = torch.Tensor(layer.mlp_1_w)
    torch.testing.assert_close(m1[0::2], t["mlp.gate_proj.weight"])
    torch.testing.assert_close(m1[1::2], t["mlp.up_proj.weight"])
    moe = torch.Tensor(layer.moe_1)
    assert moe.shape == (2, 512, 16)
    torch.testing.assert_close(moe[1, 1::2], t["moe.experts.1.w3.weight"])
    with pytest.raises(ValueError):
        layer.set_dependency("not.a.weight", torch.zeros(1))


def test_tp_sharding_through_transforms():
    t = _tensors()
    shards = []
    for rank in range(2):
        layer = _Layer(_Model(_arch(), rank, 2))
        for n, v in t.items():
            layer.set_dependency(n, v)
        shards.append(layer)
        assert layer.qkv_w.shape == (2 * 4 + 2 * 1 * 4, 16) and layer.attn_out_w.shape == (16, 8)
        assert layer.mlp_1_w.shape == (256, 16) and layer.mlp_2_w.shape == (16, 128) and layer.moe_1.shape == (2, 256, 16)
    # the two shards of the interleaved gated weight are the two halves of gate AND of up
    torch.testing.assert_close(torch.Tensor(shards[1].mlp_1_w)[0::2], t["mlp.gate_proj.weight"][128:])
    torch.testing.assert_close(torch.Tensor(shards[0].mlp_1_w)[1::2], t["mlp.up_proj.weight"][:128])
    # x @ W2^T summed over shards equals the unsharded product
    x = torch.randn(3, 256)
    full = x @ t["mlp.down_proj.weight"].t()
    part = sum(x[:, r * 128:(r + 1) * 128] @ torch.Tensor(shards[r].mlp_2_w).t() for r in range(2))
    torch.testing.assert_close(part, full, rtol=1e-4, atol=1e-4)


def test_megatron_layouts_and_fused_gated():
    m = _Model(_arch(n_kv_heads=4))
    w = torch.arange(3 * 4 * 4 * 2, dtype=torch.float32).reshape(48, 2)      # [n_heads, 3, head, in]
    got = cp.megatron_qkv_reshape(w, 4, 4).reshape(3, 4, 4, 2)
    want = w.reshape(4, 3, 4, 2).transpose(0, 1)
    torch.testing.assert_close(got, want)
    g = torch.arange((2 * 2 + 4) * 4 * 2, dtype=torch.float32).reshape(32, 2)  # 2 kv groups x (2 q + k + v)
    out = cp.transform_gqa_megatron(g, 4, 4, 2)
    grp = g.reshape(2, 4, 4, 2)
    torch.testing.assert_close(out[:16], grp[:, :2].reshape(16, 2))
    torch.testing.assert_close(out[16:24], grp[:, 2].reshape(8, 2))
    bias = torch.arange(32.0)
    assert cp.transform_gqa_megatron(bias, 4, 4, 2).shape == (32, )

    class L(LayerContainer):
        mlp: cp.FusedGatedMLPParameter
        PARAM_MAPPING = {"mlp.gate_up.weight": "mlp.params"}

    layer = L(m)
    w = torch.randn(512, 16)
    layer.set_dependency("mlp.gate_up.weight", w)
    torch.testing.assert_close(torch.Tensor(layer.mlp)[0::2], w[:256])
    torch.testing.assert_close(torch.Tensor(layer.mlp)[1::2], w[256:])


def test_mapping_validation_and_double_set():
    with pytest.raises(ValueError, match="not found in this layer"):
        class A(LayerContainer):
            p: cp.NormParameter
            PARAM_MAPPING = {"w": "q.params"}
    with pytest.raises(ValueError, match="not mapped"):
        class B(LayerContainer):
            p: cp.UnfusedQKVParameter
            PARAM_MAPPING = {"w": "p.q_params"}
    with pytest.raises(ValueError, match="multiple mapping rules"):
        class C(LayerContainer):
            p: cp.NormParameter
            PARAM_MAPPING = {"w": "p.params", "v": "p.params"}
    with pytest.raises(ValueError, match="wildcard"):
        class D(LayerContainer):
            p: cp.UnfusedMoEMLP2Parameter
            PARAM_MAPPING = {"experts.w2": "p.experts"}

    class E(LayerContainer):
        p: cp.UnfusedQKVParameter
        PARAM_MAPPING = {"q": "p.q_params", "k": "p.k_params", "v": "p.v_params"}

    class F(E):                       # mappings and fields are inherited
        n: cp.NormParameter
        PARAM_MAPPING = {"ln": "n.params"}

    layer = F(_Model(_arch()))
    assert layer.annotation_attrs == ["p", "n"] and set(layer.mapping_params) == {"q", "k", "v", "ln"}
    layer.set_dependency("q", torch.zeros(16, 16))
    with pytest.raises(ValueError, match="twice"):
        layer.set_dependency("q", torch.zeros(16, 16))
    layer.direct_injection("n", InferenceParameter.initialize(torch.ones(16)))
    assert not layer.is_populated


def test_inference_parameter_aux():
    p = InferenceParameter.initialize(torch.ones(4, dtype=torch.float32), scales=torch.full((2, ), 3.0))
    assert p.scales.tolist() == [3.0, 3.0] and set(p.aux_attrs) == {"scales"}
    q = p.to(torch.float16)
    assert isinstance(q, InferenceParameter) and q.dtype == torch.float16 and q.scales.dtype == torch.float16
    with pytest.raises(ValueError):
        InferenceParameter.initialize(torch.ones(2), scales=3.0)
    r = InferenceParameter.initialize_raw(_ds_core_param_key=torch.zeros(3), zp=torch.ones(1))
    assert r.shape == (3, ) and r.zp.item() == 1


def test_custom_parameter_with_list():
    class Stack(ParameterBase):
        parts: ParamList("n_parts")
        scale: torch.Tensor

        def finalize(self):
            return InferenceParameter.initialize(torch.stack(tuple(self.parts)) * self.scale)

    class L(LayerContainer):
        s: Stack
        PARAM_MAPPING = {"part.*.w": "s.parts", "scale": "s.scale"}

    model = type("M", (), {"n_parts": 3})()
    layer = L(model)
    for i in (2, 0):
        layer.set_dependency(f"part.{i}.w", torch.full((2, ), float(i)))
    layer.set_dependency("scale", torch.tensor(2.0))
    assert isinstance(layer.s, Stack)
    layer.set_dependency("part.1.w", torch.ones(2))
    assert torch.Tensor(layer.s).tolist() == [[0, 0], [2, 2], [4, 4]]


def test_training_kernel_injection_bert_and_policy_tables():
    """module_inject.inject: HF BERT layers -> fused training transformer layers with the same output."""
    import torch
    from transformers import BertConfig
    from transformers.models.bert.modeling_bert import BertEncoder, BertLayer

    from hcache_deepspeed_b200.model_implementations.transformers.ds_bert import DeepSpeedBERTInference  # noqa: F401
    from hcache_deepspeed_b200.module_inject import replace_policy
    from hcache_deepspeed_b200.module_inject.auto_tp_model_utils import build_bloom_alibi_tensor
    from hcache_deepspeed_b200.module_inject.inject import module_inject
    from hcache_deepspeed_b200.module_inject.utils import policy_to_ds_container
    torch.manual_seed(0)
    cfg = BertConfig(hidden_size=32, num_attention_heads=4, intermediate_size=64, num_hidden_layers=2, hidden_dropout_prob=0.0,
                     attention_probs_dropout_prob=0.0, hidden_act="gelu")
    cfg._attn_implementation = "eager"
    enc = BertEncoder(cfg).eval()
    x = torch.randn(2, 8, 32)
    want = enc(x).last_hidden_state
    module_inject(BertLayer, enc, cfg, 2, 8, 0, preln=False, fp16=False)
    h = x
    for layer in enc.layer:
        assert type(layer).__name__ == "DeepSpeedTransformerLayer"
        out = layer(h)
        h = out[0] if isinstance(out, tuple) else out
    torch.testing.assert_close(h, want, atol=1e-5, rtol=1e-5)
    assert len(replace_policy.replace_policies) >= 8 and policy_to_ds_container(block="GPT2Block").__name__ == "gpt2"
    a = build_bloom_alibi_tensor(torch.ones(2, 5, dtype=torch.long), 8, torch.float32)
    assert a.shape == (16, 1, 5) and torch.allclose(a[0, 0], torch.arange(5.0) * 0.5)
