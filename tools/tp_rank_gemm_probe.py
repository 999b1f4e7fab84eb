"""Decode GEMMs at the per-rank shapes of Llama-2-7B under TP=8 (and TP=1 for reference): which kernel / tile / split
serves them best. CUDA events, L2 flushed between iterations. One JSON line per variant."""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from hcache_deepspeed_b200.ops import kernels as K  # noqa: E402

dev = "cuda"
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)


def timeit(fn, iters=12, warm=3):
    for _ in range(warm):
        fn()
    ts = []
    for _ in range(iters):
        flush.zero_()
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record()
        fn()
        e.record()
        torch.cuda.synchronize()
        ts.append(s.elapsed_time(e) * 1e3)
    ts.sort()
    return ts[len(ts) // 2]


def rnd(*s):
    return (torch.randn(*s, device=dev) * 0.05).to(torch.bfloat16)


ws = torch.zeros(64 << 20, dtype=torch.uint8, device=dev)
for M in (8, 64):
    for name, N, Kd, gated, accum in (("qkv", 1536, 4096, False, False), ("o_proj", 4096, 512, False, True),
                                      ("mlp1_gated", 2752, 4096, True, False), ("mlp2", 4096, 1376, False, True)):
        x, w = rnd(M, Kd), rnd(N, Kd)
        out = torch.empty(M, N // 2 if gated else N, dtype=torch.bfloat16, device=dev)
        acc = torch.zeros(M, N, dtype=torch.float32, device=dev)
        floor_us = (N * Kd * 2) / 6.5e6
        variants = []
        if accum:
            for bn in (0, -1, 64, 128):
                for sp in (0, 1, 2, 4):
                    variants.append((f"accum bn={bn} splits={sp}", lambda bn=bn, sp=sp: K.gemm_accum(x, w, acc, bn=bn, splits=sp)))
        else:
            for bn in (0, -1, 64, 128):
                for sp in (0, 1, 2, 4, 8):
                    variants.append((f"{'gated' if gated else 'store'} bn={bn} splits={sp}",
                                     lambda bn=bn, sp=sp: K.gemm(x, w, act=3 if gated else 0, gated=gated, out=out, bn=bn, splits=sp, workspace=ws)))
        best = None
        for label, fn in variants:
            try:
                us = timeit(fn)
            except Exception as e:  # noqa: BLE001
                print(json.dumps({"M": M, "gemm": name, "variant": label, "error": str(e)[:80]}))
                continue
            rec = {"M": M, "gemm": name, "N": N, "K": Kd, "variant": label, "us": round(us, 2), "hbm_floor_us": round(floor_us, 2)}
            print(json.dumps(rec), flush=True)
            if best is None or us < best[0]:
                best = (us, label)
        print(json.dumps({"M": M, "gemm": name, "BEST": best[1], "us": round(best[0], 2), "hbm_floor_us": round(floor_us, 2)}), flush=True)
# the norm that follows: 64 rows x 4096, fp32 delta
x = rnd(64, 4096)
d = torch.zeros(64, 4096, dtype=torch.float32, device=dev)
g = torch.ones(4096, dtype=torch.bfloat16, device=dev)
o = torch.empty_like(x)
print(json.dumps({"norm 64x4096 fp32 delta us": round(timeit(lambda: K.norm(x, g, o, delta=d, zero_delta=True, res_out=x)), 2)}))
