"""Where the gated epilogue's time goes: fp8 / bf16 CTA-pair GEMM, store vs gated, identity vs SiLU."""
import os
import sys
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import torch
from hcache_deepspeed_b200.ops import kernels as K

dev = "cuda"


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


M, N, Kd = 4160, 22016, 4096
a = torch.randn(M, Kd, device=dev).bfloat16()
w = (torch.randn(N, Kd, device=dev) * Kd ** -0.5).bfloat16()
aq, a_s = K.quantize_fp8_rows(a)
wq, w_s = K.quantize_fp8_rows(w)
full = torch.empty(M, N, device=dev, dtype=torch.bfloat16)
half = torch.empty(M, N // 2, device=dev, dtype=torch.bfloat16)
fl = 2.0 * M * N * Kd
for name, fn in [
    ("bf16 2sm store", lambda: K.gemm_2sm(a, w, out=full, bn=256)),
    ("bf16 2sm store+silu", lambda: K.gemm_2sm(a, w, out=full, act=3, bn=256)),
    ("bf16 2sm gated identity", lambda: K.gemm_2sm(a, w, out=half, act=0, gated=True, bn=256)),
    ("bf16 2sm gated silu", lambda: K.gemm_2sm(a, w, out=half, act=3, gated=True, bn=256)),
    ("fp8 2sm store", lambda: K.gemm_fp8(aq, a_s, wq, w_s, full, bn=1256)),
    ("fp8 2sm store+silu", lambda: K.gemm_fp8(aq, a_s, wq, w_s, full, act=3, bn=1256)),
    ("fp8 2sm gated identity", lambda: K.gemm_fp8(aq, a_s, wq, w_s, half, act=0, gated=True, bn=1256)),
    ("fp8 2sm gated silu", lambda: K.gemm_fp8(aq, a_s, wq, w_s, half, act=3, gated=True, bn=1256)),
]:
    t = timeit(fn)
    print(f"{name:28s} {t * 1000:8.1f} us  {fl / t / 1e9:8.1f} TFLOP/s", flush=True)
