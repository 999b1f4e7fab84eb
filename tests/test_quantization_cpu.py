"""Integer and minifloat quantizers (torch arithmetic on CPU; the CUDA kernels are checked against it in
tests/test_gpu_quant.py), quantised layers, LoRA / quantised OptimizedLinear. Mirrors tests/unit/ops/quantizer,
tests/unit/ops/fp_quantizer, tests/unit/inference/quantization, tests/unit/linear of the reference."""
import pytest
import torch


@pytest.mark.parametrize("bits,sym", [(8, True), (8, False), (4, True), (4, False)])
def test_int_quantize_roundtrip(bits, sym):
    from hcache_deepspeed_b200.ops.quantizer import dequantize_int, quantize_int
    torch.manual_seed(0)
    x = torch.randn(16, 256) * 3 + (0.0 if sym else 2.0)
    codes, params = quantize_int(x, groups=16, bits=bits, symmetric=sym)
    assert codes.numel() == x.numel() * bits // 8 and params.shape == (16, 2)
    back = dequantize_int(codes, params, bits, torch.float32, x.shape)
    step = params[:, 0].max()
    assert float((back - x).abs().max()) <= float(step) * 0.5 + 1e-5


@pytest.mark.parametrize("q_bits,rel", [(8, 0.04), (6, 0.08), (4, 0.18), (12, 0.007)])
def test_fp_quantize_roundtrip(q_bits, rel):
    from hcache_deepspeed_b200.ops.fp_quantizer import FP_Quantize
    torch.manual_seed(1)
    x = torch.randn(4, 8, 512, dtype=torch.bfloat16)
    q = FP_Quantize(group_size=512)
    packed = q.quantize(x, q_bits=q_bits)
    assert packed.dtype == torch.uint8 and packed.numel() == x.numel() * q_bits // 8 + (x.numel() // 512) * 4
    back = q.dequantize(packed, q_bits=q_bits).float()
    # worst case = half of the widest step of the format, relative to the group maximum
    err = (back - x.float()).abs().reshape(-1, 512).amax(dim=1) / x.float().abs().reshape(-1, 512).amax(dim=1)
    assert float(err.max()) < rel
    sel = q.selective_dequantize(packed, torch.tensor([2, 0]), q_bits=q_bits).float()
    torch.testing.assert_close(sel, back[[2, 0]])


def test_fp_codes_are_exact_for_representable_values():
    from hcache_deepspeed_b200.ops.fp_quantizer.quantize import _decode_torch, _encode_torch
    for q_bits in (4, 6, 8, 12):
        codes = torch.arange(1 << q_bits)
        vals = _decode_torch(codes, q_bits)
        again = _encode_torch(vals, q_bits)
        keep = vals != 0  # +0 / -0 share a value
        assert torch.equal(again[keep], codes[keep]), q_bits


def test_fake_quantizer_and_zeropp_helpers():
    from hcache_deepspeed_b200.ops.quantizer import ds_quantizer, quantize_int, quantized_reduction, swizzled_quantize, dequantize_int
    torch.manual_seed(2)
    x = torch.randn(8, 128)
    fq = ds_quantizer(x.clone(), groups=8, bit_num=8)
    assert fq.shape == x.shape and float((fq - x).abs().max()) < 0.05
    codes, params = swizzled_quantize(x, groups=8, bits=8, pipeline_size=1, nodes=2, devices_per_node=2)
    plain_c, plain_p = quantize_int(x, 8, 8)
    # swizzle = permutation of groups: (node, dev) -> (dev, node)
    perm = torch.arange(8).reshape(1, 2, 2, 2).permute(0, 2, 1, 3).reshape(-1)
    assert torch.equal(codes.reshape(8, -1), plain_c.reshape(8, -1)[perm]) and torch.equal(params, plain_p[perm])
    c2, p2 = quantized_reduction(plain_c, plain_p, in_groups=8, out_groups=4, bits=8, out_bits=8, devices_per_node=2)
    want = x.reshape(2, -1).sum(0)
    got = dequantize_int(c2, p2, 8, torch.float32)
    assert float((got - want).abs().max()) < 0.1


def test_post_init_weight_quantization():
    from hcache_deepspeed_b200.inference.quantization import QuantizedEmbedding, QuantizedLinear, _init_group_wise_weight_quantization
    torch.manual_seed(3)
    m = torch.nn.Sequential()
    m.add_module("embed", torch.nn.Embedding(32, 64))
    m.add_module("fc1", torch.nn.Linear(64, 128))
    m.add_module("head", torch.nn.Linear(128, 16))
    ids = torch.randint(0, 32, (3, 5))
    want = m(ids)
    cfg = {"weight_quantization": {"post_init_quant": {"fc": {"num_bits": 8, "group_size": 32, "group_dim": 1, "symmetric": False},
                                                       "embed": {"num_bits": 4, "group_size": 32, "group_dim": 1, "symmetric": False}}}}
    _init_group_wise_weight_quantization(m, cfg)
    assert isinstance(m.fc1, QuantizedLinear) and isinstance(m.embed, QuantizedEmbedding) and type(m.head) is torch.nn.Linear
    assert m.fc1.weight.dtype == torch.int8 and m.fc1.weight.numel() == 64 * 128 and m.embed.weight.numel() == 32 * 64 // 2
    got = m(ids)
    assert float((got - want).abs().max()) < 0.35 * float(want.abs().max())


def test_optimized_linear_variants():
    from hcache_deepspeed_b200.linear import LoRAConfig, LoRAOptimizedLinear, OptimizedLinear, QuantizationConfig, QuantizedLinear
    torch.manual_seed(4)
    assert type(OptimizedLinear(64, 32)) is torch.nn.Linear
    ql = OptimizedLinear(512, 32, quantization_config=QuantizationConfig(q_bits=8, group_size=512))
    assert isinstance(ql, QuantizedLinear) and ql.weight.data.dtype == torch.uint8
    x = torch.randn(4, 512, dtype=torch.bfloat16)
    assert ql(x).shape == (4, 32)
    lora = OptimizedLinear(512, 32, lora_config=LoRAConfig(lora_r=8, lora_alpha=16), quantization_config=QuantizationConfig(q_bits=6))
    assert isinstance(lora, LoRAOptimizedLinear)
    y = lora(x)
    y.float().sum().backward()
    assert lora.lora_weight_2.weight.grad is not None and not lora.weight.requires_grad
    # zero-initialised second LoRA matrix: the layer starts as the (quantised) base layer
    torch.testing.assert_close(y.float(), torch.nn.functional.linear(x, lora.full_weight()).float(), rtol=1e-2, atol=1e-2)
