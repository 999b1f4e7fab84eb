"""CUDA quantizer + fused-optimizer kernels against the torch restatements."""
import pytest
import torch

pytestmark = [pytest.mark.gpu, pytest.mark.timeout(300)]


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16, torch.float32])
@pytest.mark.parametrize("bits,sym", [(8, True), (8, False), (4, True), (4, False)])
def test_int_quantizer_matches_cpu(dtype, bits, sym):
    from hcache_deepspeed_b200.ops.quantizer import dequantize_int, quantize_int
    torch.manual_seed(0)
    x = (torch.randn(64, 1024) * 2 + (0 if sym else 1)).to(dtype)
    cg, pg = quantize_int(x.cuda(), 64, bits, sym)
    cc, pc = quantize_int(x, 64, bits, sym)
    torch.testing.assert_close(pg.cpu(), pc, rtol=1e-6, atol=1e-7)
    diff = (cg.cpu().to(torch.int32) - cc.to(torch.int32)) if bits == 8 else None
    if diff is not None:
        assert int(diff.abs().max()) <= 1 and float((diff != 0).float().mean()) < 1e-3  # ties may round differently
    back = dequantize_int(cg, pg, bits, dtype, x.shape).cpu().float()
    assert float((back - x.float()).abs().max()) <= float(pc[:, 0].max()) * 0.51 + 1e-2 * float(x.float().abs().max())


@pytest.mark.parametrize("q_bits", [4, 6, 8, 12])
def test_fp_quantizer_matches_cpu(q_bits):
    from hcache_deepspeed_b200.ops.fp_quantizer import fp_dequantize, fp_quantize
    torch.manual_seed(1)
    x = torch.randn(32, 512, dtype=torch.bfloat16)
    pg, sg = fp_quantize(x.cuda(), q_bits, 512)
    pc, sc = fp_quantize(x, q_bits, 512)
    torch.testing.assert_close(sg.cpu(), sc)
    assert float((pg.cpu() != pc).float().mean()) < 2e-3
    bg = fp_dequantize(pg, sg, q_bits, 512, torch.bfloat16, x.shape).cpu().float()
    bc = fp_dequantize(pc, sc, q_bits, 512, torch.bfloat16, x.shape).float()
    assert float((bg - bc).abs().max()) <= float(x.float().abs().max()) * 0.26
    idx = torch.tensor([5, 0, 31], dtype=torch.int32)
    sel = fp_dequantize(pg, sg, q_bits, 512, torch.bfloat16, (3, 512), group_idx=idx.cuda()).cpu()
    torch.testing.assert_close(sel.float(), bg[idx.long()])


@pytest.mark.parametrize("pdt,gdt", [(torch.float32, torch.float32), (torch.float32, torch.bfloat16), (torch.bfloat16, torch.bfloat16)])
@pytest.mark.parametrize("adamw", [True, False])
def test_fused_adam_matches_torch(pdt, gdt, adamw):
    from hcache_deepspeed_b200.ops.adam import FusedAdam
    torch.manual_seed(2)
    shapes = [(1000, 33), (7, ), (70001, ), (256, 256)] * 8  # > 24 tensors: several launches
    ps = [torch.nn.Parameter(torch.randn(*s, device="cuda").to(pdt)) for s in shapes]
    ref = [torch.nn.Parameter(p.detach().float().clone()) for p in ps]
    if pdt != gdt:
        for p in ps:
            p.grad_dtype = None  # low-precision gradients for fp32 master parameters
    opt = FusedAdam(ps, lr=1e-2, weight_decay=0.05, adam_w_mode=adamw)
    ropt = (torch.optim.AdamW if adamw else torch.optim.Adam)(ref, lr=1e-2, weight_decay=0.05)
    for _ in range(3):
        for p, r in zip(ps, ref):
            g = torch.randn_like(r)
            p.grad = g.to(gdt)
            r.grad = g.to(gdt).float()
        opt.step()
        ropt.step()
    tol = 1e-5 if pdt == torch.float32 else 3e-2
    for p, r in zip(ps, ref):
        torch.testing.assert_close(p.detach().float(), r.detach(), rtol=tol, atol=tol)


def test_fused_lion_matches_reference():
    from hcache_deepspeed_b200.ops.lion import FusedLion
    from hcache_deepspeed_b200.ops.lion.lion import lion_reference
    torch.manual_seed(3)
    p = torch.nn.Parameter(torch.randn(100000, device="cuda"))
    w, m = p.detach().cpu().clone(), torch.zeros(100000)
    opt = FusedLion([p], lr=1e-3, weight_decay=0.1)
    for _ in range(3):
        g = torch.randn(100000)
        p.grad = g.cuda()
        opt.step()
        lion_reference(w, g, m, 1e-3, 0.9, 0.999, 0.1)
    torch.testing.assert_close(p.detach().cpu(), w, rtol=1e-5, atol=1e-6)


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
def test_quantize_fp8_rows(dtype):
    from hcache_deepspeed_b200.ops import kernels as K
    torch.manual_seed(0)
    x = (torch.randn(37, 1000, device="cuda") * torch.logspace(-2, 2, 37, device="cuda")[:, None]).to(dtype)
    x[5] = 0
    q, s = K.quantize_fp8_rows(x)
    amax = x.float().abs().amax(dim=1)
    want_s = torch.where(amax > 0, amax / 448.0, torch.ones_like(amax))
    torch.testing.assert_close(s, want_s, rtol=1e-6, atol=0)
    want_q = (x.float() / want_s[:, None]).to(torch.float8_e4m3fn)
    got = q.view(torch.float8_e4m3fn).float()
    # identical rounding except where the fp32 division lands within an ulp of a tie
    assert (got != want_q.float()).float().mean() < 1e-3
    deq = got * s[:, None]
    assert ((deq - x.float()).abs() <= 0.0625 * amax[:, None] + 1e-12).all()


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
@pytest.mark.parametrize("M,N,Kd", [(1, 256, 128), (100, 1000, 2096), (300, 4096, 4096), (2048, 1376, 512)])
@pytest.mark.parametrize("mode", ["store", "gated"])
@pytest.mark.parametrize("bn", [0, 64, 128, 256])
def test_gemm_fp8(dtype, M, N, Kd, mode, bn):
    """E4M3 x E4M3 -> fp32 (TMEM) GEMM against an fp32 matmul of the de-quantised operands."""
    from hcache_deepspeed_b200.ops import kernels as K
    torch.manual_seed(M + N)
    a = torch.randn(M, Kd, device="cuda").to(dtype)
    w = (torch.randn(N, Kd, device="cuda") * Kd ** -0.5).to(dtype)
    bias = torch.randn(N, device="cuda").to(dtype)
    aq, a_s = K.quantize_fp8_rows(a)
    wq, w_s = K.quantize_fp8_rows(w)
    gated = mode == "gated"
    out = torch.empty(M, N // 2 if gated else N, dtype=dtype, device="cuda")
    K.gemm_fp8(aq, a_s, wq, w_s, out, bias=bias, act=K.ACT_SILU if gated else K.ACT_IDENTITY, gated=gated, bn=bn)
    ad = aq.view(torch.float8_e4m3fn).float() * a_s[:, None]
    wd = wq.view(torch.float8_e4m3fn).float() * w_s[:, None]
    want = ad @ wd.t() + bias.float()
    if gated:
        want = torch.nn.functional.silu(want[:, 0::2]) * want[:, 1::2]
    tol = dict(rtol=2e-2, atol=2e-2) if dtype == torch.bfloat16 else dict(rtol=4e-3, atol=4e-3)
    torch.testing.assert_close(out.float(), want, **tol)
    # and the quantised product tracks the unquantised one (fp8 error, not a wiring error)
    full = a.float() @ w.float().t() + bias.float()
    if not gated:
        assert (out.float() - full).abs().mean() < 0.05 * full.abs().mean() + 0.02


@pytest.mark.parametrize("M", [1, 8, 16, 33, 64])
@pytest.mark.parametrize("mode", ["store", "gated", "accum"])
@pytest.mark.parametrize("use_ws", [False, True])
def test_gemm_fp8_skinny(M, mode, use_ws):
    """Decode-size E4M3 GEMM on the feature-major kernel: ragged N / K, every token tile, split-K through scratch
    (left zeroed) or straight into the fp32 accumulator."""
    from hcache_deepspeed_b200.ops import kernels as K
    torch.manual_seed(M)
    N, Kd = (690 if mode == "gated" else 1000), 2128
    a = torch.randn(M, Kd, device="cuda").bfloat16()
    w = (torch.randn(N, Kd, device="cuda") * Kd ** -0.5).bfloat16()
    bias = torch.randn(N, device="cuda").bfloat16()
    aq, a_s = K.quantize_fp8_rows(a)
    wq, w_s = K.quantize_fp8_rows(w)
    want = (aq.view(torch.float8_e4m3fn).float() * a_s[:, None]) @ (wq.view(torch.float8_e4m3fn).float() * w_s[:, None]).t() \
        + bias.float()
    ws = torch.zeros(K.gemm_fp8_skinny_workspace_bytes(M, N), dtype=torch.uint8, device="cuda") if use_ws else None
    for rep in range(2):
        if mode == "accum":
            acc = torch.full((M, N), 0.5, device="cuda", dtype=torch.float32)
            K.gemm_fp8_accum(aq, a_s, wq, w_s, acc, bias=bias)
            torch.testing.assert_close(acc, want + 0.5, rtol=2e-3, atol=2e-3)
        else:
            gated = mode == "gated"
            out = torch.empty(M, N // 2 if gated else N, dtype=torch.bfloat16, device="cuda")
            K.gemm_fp8(aq, a_s, wq, w_s, out, bias=bias, act=K.ACT_SILU if gated else K.ACT_IDENTITY, gated=gated, workspace=ws)
            ref_out = torch.nn.functional.silu(want[:, 0::2]) * want[:, 1::2] if gated else want
            torch.testing.assert_close(out.float(), ref_out, rtol=2e-2, atol=2e-2)
    if ws is not None:
        assert int(ws.view(torch.int32).abs().max()) == 0


def _fp6_case(M, N, Kd, seed=0):
    from hcache_deepspeed_b200.ops import kernels as K
    from hcache_deepspeed_b200.ops.fp8 import Fp8Weight
    torch.manual_seed(seed)
    a = torch.randn(M, Kd, device="cuda").bfloat16()
    w = (torch.randn(N, Kd, device="cuda") * Kd ** -0.5).bfloat16()
    bias = torch.randn(N, device="cuda").bfloat16()
    aq, a_s = K.quantize_fp8_rows(a)
    w6 = Fp8Weight.from_tensor(w, "e3m2")
    want = (aq.view(torch.float8_e4m3fn).float() * a_s[:, None]) @ w6.dequantize().t() + bias.float()
    return K, aq, a_s, w6, bias, want


@pytest.mark.parametrize("M", [1, 8, 40, 64])
@pytest.mark.parametrize("mode", ["store", "gated", "accum"])
def test_gemm_fp6_weights_skinny(M, mode):
    """E3M2 (FP6) weights, packed 6 bits each in global memory and unpacked by the TMA, on the decode-size kernel: the
    result equals the fp32 product of the de-quantised operands (oracle: torch)."""
    N, Kd = (768 if mode == "gated" else 1000), 2176
    K, aq, a_s, w6, bias, want = _fp6_case(M, N, Kd, seed=M)
    if mode == "accum":
        acc = torch.full((M, N), 0.5, device="cuda", dtype=torch.float32)
        K.gemm_fp8_accum(aq, a_s, w6.q, w6.scale, acc, bias=bias, w_fmt=1)
        torch.testing.assert_close(acc, want + 0.5, rtol=2e-3, atol=2e-3)
    else:
        gated = mode == "gated"
        out = torch.empty(M, N // 2 if gated else N, dtype=torch.bfloat16, device="cuda")
        ws = torch.zeros(K.gemm_fp8_skinny_workspace_bytes(M, N), dtype=torch.uint8, device="cuda")
        K.gemm_fp8(aq, a_s, w6.q, w6.scale, out, bias=bias, act=K.ACT_SILU if gated else K.ACT_IDENTITY, gated=gated,
                   workspace=ws, w_fmt=1)
        ref_out = torch.nn.functional.silu(want[:, 0::2]) * want[:, 1::2] if gated else want
        torch.testing.assert_close(out.float(), ref_out, rtol=2e-2, atol=2e-2)


@pytest.mark.parametrize("M,N,Kd,bn", [(100, 1000, 2176, 0), (100, 512, 1024, 256), (300, 4096, 4096, 0), (2048, 1408, 512, 0)])
def test_gemm_fp6_weights_token_major(M, N, Kd, bn):
    """The same weight format on the token-major kernels: single-CTA (M <= 128 or explicit tile) and the CTA-pair kernel."""
    K, aq, a_s, w6, bias, want = _fp6_case(M, N, Kd, seed=3)
    out = torch.empty(M, N, dtype=torch.bfloat16, device="cuda")
    K.gemm_fp8(aq, a_s, w6.q, w6.scale, out, bias=bias, bn=bn, w_fmt=1)
    torch.testing.assert_close(out.float(), want, rtol=2e-2, atol=3e-2)


def test_fp6_linear_module_gpu():
    """quantization_mode='wf6af16' resolves to the fp6 tcgen05 linear; output tracks the unquantised product."""
    from types import SimpleNamespace
    from hcache_deepspeed_b200.inference.v2.modules import heuristics
    from hcache_deepspeed_b200.inference.v2.modules.configs import DSLinearConfig
    cfg = DSLinearConfig(in_channels=1024, out_channels=768, input_dtype=torch.bfloat16, output_dtype=torch.bfloat16)
    lin = heuristics.instantiate_linear(cfg, SimpleNamespace(quantization=SimpleNamespace(quantization_mode="wf6af16")))
    assert lin.name() == "fp6_tcgen05_linear"
    torch.manual_seed(0)
    w = (torch.randn(768, 1024, device="cuda") * 0.03).bfloat16()
    x = torch.randn(20, 1024, device="cuda").bfloat16()
    y = lin(x, lin.transform_param(w))
    want = x.float() @ w.float().t()
    cos = torch.nn.functional.cosine_similarity(y.float().flatten(), want.flatten(), dim=0)
    assert cos > 0.99, float(cos)
