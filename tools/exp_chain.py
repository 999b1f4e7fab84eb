"""Experiment: is the ~10 us fixed cost per decode GEMM a launch / shared-memory-carve-out effect?
Back-to-back GEMMs over distinct weight buffers (no L2 reuse), alone and interleaved with a tiny kernel."""
import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from hcache_deepspeed_b200.ops import kernels as K
dev = "cuda"
M, N, Kd, n = 16, 22016, 4096, 16
a = (torch.randn(M, Kd, device=dev)).to(torch.bfloat16)
ws = [(torch.randn(N, Kd, device=dev) * 0.02).to(torch.bfloat16) for _ in range(n)]
out = torch.empty(M, N // 2, device=dev, dtype=torch.bfloat16)
x, g, y = torch.randn(M, Kd, device=dev).to(torch.bfloat16), torch.ones(Kd, device=dev, dtype=torch.bfloat16), torch.empty(M, Kd, device=dev, dtype=torch.bfloat16)
small = torch.zeros(1024, device=dev)
def run(mode):
    for w in ws:
        K.gemm(a, w, act=3, gated=True, out=out, bn=128)
        if mode == "norm":
            K.norm(x, g, y, eps=1e-5)
        elif mode == "torch":
            small.add_(1.0)
for mode in ("none", "norm", "torch", "none"):
    for _ in range(3): run(mode)
    torch.cuda.synchronize()
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record(); run(mode); e.record(); torch.cuda.synchronize()
    ms = s.elapsed_time(e)
    print(f"mode={mode:6s} total {ms:.3f} ms -> {ms/n*1000:.1f} us per GEMM(+small), {N*Kd*2*n/ms/1e6:.0f} GB/s")
# CUDA graph of the same chain
gph = torch.cuda.CUDAGraph()
st = torch.cuda.Stream()
with torch.cuda.stream(st):
    run("none"); torch.cuda.synchronize()
    with torch.cuda.graph(gph, stream=st):
        run("none")
torch.cuda.synchronize()
for _ in range(2): gph.replay()
torch.cuda.synchronize()
s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
s.record(); gph.replay(); e.record(); torch.cuda.synchronize()
ms = s.elapsed_time(e)
print(f"graph replay: {ms/n*1000:.1f} us per GEMM, {N*Kd*2*n/ms/1e6:.0f} GB/s")
