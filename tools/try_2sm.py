"""Numerics + timing of the CTA-pair GEMM (bf16 and fp8) against the single-CTA kernels and cuBLAS."""
import os
import sys
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import torch
from hcache_deepspeed_b200.ops import kernels as K

dev = "cuda"
torch.manual_seed(0)


def timeit(fn, iters=10):
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    for _ in range(3):
        fn()
    ts = []
    for _ in range(iters):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    return sorted(ts)[len(ts) // 2]


def check(M, N, Kd, bn, gated=False):
    a = torch.randn(M, Kd, device=dev).bfloat16()
    w = (torch.randn(N, Kd, device=dev) * Kd ** -0.5).bfloat16()
    b = torch.randn(N, device=dev).bfloat16()
    out = K.gemm_2sm(a, w, bias=b, act=3 if gated else 0, gated=gated, bn=bn)
    torch.cuda.synchronize()
    want = a.float() @ w.float().t() + b.float()
    if gated:
        want = torch.nn.functional.silu(want[:, 0::2]) * want[:, 1::2]
    err = (out.float() - want).abs().max().item()
    print(f"bf16 M{M} N{N} K{Kd} bn{bn} gated={gated}: max err {err:.4f}", flush=True)
    assert err < 0.08, err
    aq, a_s = K.quantize_fp8_rows(a)
    wq, w_s = K.quantize_fp8_rows(w)
    o8 = torch.empty_like(out)
    K.gemm_fp8(aq, a_s, wq, w_s, o8, bias=b, act=3 if gated else 0, gated=gated, bn=1000 + bn)
    torch.cuda.synchronize()
    want8 = (aq.view(torch.float8_e4m3fn).float() * a_s[:, None]) @ (wq.view(torch.float8_e4m3fn).float() * w_s[:, None]).t() + b.float()
    if gated:
        want8 = torch.nn.functional.silu(want8[:, 0::2]) * want8[:, 1::2]
    err8 = (o8.float() - want8).abs().max().item()
    print(f"fp8  M{M} N{N} K{Kd} bn{bn} gated={gated}: max err {err8:.4f}", flush=True)
    assert err8 < 0.08, err8


for shape in [(256, 256, 128), (300, 1000, 2096), (1024, 4096, 4096), (4096, 1376, 512), (130, 512, 64)]:
    for bn in (128, 256):
        check(*shape, bn)
check(700, 2 * 688, 512, 256, gated=True)
check(700, 2 * 688, 512, 128, gated=True)
print("numerics ok", flush=True)

for (M, N, Kd) in [(4096, 4096, 4096), (4096, 12288, 4096), (4096, 4096, 11008), (8192, 8192, 8192), (4096, 22016, 4096)]:
    a = torch.randn(M, Kd, device=dev).bfloat16()
    w = (torch.randn(N, Kd, device=dev) * Kd ** -0.5).bfloat16()
    out = torch.empty(M, N, device=dev, dtype=torch.bfloat16)
    fl = 2.0 * M * N * Kd
    t2 = {bn: timeit(lambda: K.gemm_2sm(a, w, out=out, bn=bn)) for bn in (128, 256)}
    t1 = timeit(lambda: K.gemm(a, w, out=out, bn=256))
    tc = timeit(lambda: torch.matmul(a, w.t(), out=out))
    aq, a_s = K.quantize_fp8_rows(a)
    wq, w_s = K.quantize_fp8_rows(w)
    t8 = timeit(lambda: K.gemm_fp8(aq, a_s, wq, w_s, out, bn=256))
    t82 = {bn: timeit(lambda: K.gemm_fp8(aq, a_s, wq, w_s, out, bn=1000 + bn)) for bn in (128, 256)}
    tf = lambda t: round(fl / t / 1e9, 1)
    print(f"M{M} N{N} K{Kd} TFLOP/s: bf16 2sm bn128 {tf(t2[128])} bn256 {tf(t2[256])} | 1sm {tf(t1)} | cublas {tf(tc)} || "
          f"fp8 2sm bn128 {tf(t82[128])} bn256 {tf(t82[256])} | 1sm {tf(t8)}", flush=True)
