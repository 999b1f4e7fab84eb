import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import importlib.util
spec = importlib.util.spec_from_file_location("kb", os.path.join(os.path.dirname(os.path.abspath(__file__)), "kernel_bench.py")); kb = importlib.util.module_from_spec(spec); sys.argv=["x"]; spec.loader.exec_module(kb)
K = kb.K
M = 16
for (N, Kd) in [(12288, 4096), (32000, 4096)]:
    a, w = kb.rand(M, Kd), kb.rand(N, Kd, scale=Kd ** -0.5)
    out = torch.empty(M, N, device="cuda", dtype=torch.bfloat16)
    for bn in (64, 128):
        med, mn = kb.timeit(lambda: K.gemm(a, w, out=out, bn=bn, splits=1))
        print(f"debug={os.environ.get('HDS_GEMM_DEBUG','0')} N{N} bn{bn}: {med*1000:.1f} us  {N*Kd*2/med/1e6:.0f} GB/s", flush=True)
