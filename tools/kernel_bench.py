"""Per-kernel timing on one B200: CUDA events, >=3 warm-ups, L2 flushed between timed iterations.
Prints one JSON line per case and writes gpurun_out/kernel_bench.jsonl. Roofline fractions are
against MEASURED_PEAKS.json (measured copy bandwidth / cuBLAS bf16), fallback 6650 GB/s, 1590 TFLOP/s.
"""
import json
import math
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from hcache_deepspeed_b200.ops import kernels as K  # noqa: E402
from hcache_deepspeed_b200.ops import reference as ref  # noqa: E402

PEAKS = {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0}
_pk = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "MEASURED_PEAKS.json")
if os.path.exists(_pk):
    PEAKS.update({k: v for k, v in json.load(open(_pk)).items() if k in ("hbm_gbs", "bf16_tflops", "bf16_tflops_sustained")})

dev = "cuda"
flush_buf = torch.empty(256 * 1024 * 1024, dtype=torch.uint8, device=dev)
results = []


def timeit(fn, iters=10, warmup=3, flush=True):
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    times = []
    for _ in range(iters):
        if flush:
            flush_buf.zero_()
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record()
        fn()
        e.record()
        torch.cuda.synchronize()
        times.append(s.elapsed_time(e))
    times.sort()
    return times[len(times) // 2], times[0]


def report(name, ms_med, ms_min, flops=None, bytes_=None, **extra):
    rec = {"case": name, "ms_median": round(ms_med, 4), "ms_min": round(ms_min, 4)}
    if flops:
        rec["tflops"] = round(flops / ms_med / 1e9, 1)
        rec["frac_of_measured_bf16"] = round(flops / ms_med / 1e9 / PEAKS["bf16_tflops"], 3)
    if bytes_:
        rec["gbs"] = round(bytes_ / ms_med / 1e6, 1)
        rec["frac_of_measured_hbm"] = round(bytes_ / ms_med / 1e6 / PEAKS["hbm_gbs"], 3)
    rec.update(extra)
    results.append(rec)
    print(json.dumps(rec), flush=True)


def rand(*shape, scale=1.0):
    return (torch.randn(*shape, device=dev) * scale).to(torch.bfloat16)


def bench_gemm():
    for (M, N, Kd) in [(4096, 4096, 4096), (4096, 12288, 4096), (4096, 4096, 11008), (8192, 8192, 8192), (2048, 8192, 4096)]:
        a, w = rand(M, Kd), rand(N, Kd, scale=Kd ** -0.5)
        out = torch.empty(M, N, device=dev, dtype=torch.bfloat16)
        for bn in (128, 256):
            med, mn = timeit(lambda: K.gemm(a, w, out=out, bn=bn))
            report(f"gemm_store M{M} N{N} K{Kd} bn{bn}", med, mn, flops=2.0 * M * N * Kd)
        med, mn = timeit(lambda: K.gemm_2sm(a, w, out=out, bn=256))
        report(f"gemm_2sm_store M{M} N{N} K{Kd} bn256", med, mn, flops=2.0 * M * N * Kd)
        med, mn = timeit(lambda: torch.matmul(a, w.t(), out=out))
        report(f"cublas M{M} N{N} K{Kd}", med, mn, flops=2.0 * M * N * Kd)
    # gated up-projection (Llama-2-7B: I=11008)
    M, N, Kd = 4096, 22016, 4096
    a, w = rand(M, Kd), rand(N, Kd, scale=Kd ** -0.5)
    out = torch.empty(M, N // 2, device=dev, dtype=torch.bfloat16)
    med, mn = timeit(lambda: K.gemm(a, w, act=3, gated=True, out=out, bn=256))
    report(f"gemm_gated M{M} N{N} K{Kd} bn256", med, mn, flops=2.0 * M * N * Kd)
    med, mn = timeit(lambda: K.gemm_2sm(a, w, act=3, gated=True, out=out, bn=256))
    report(f"gemm_2sm_gated M{M} N{N} K{Kd} bn256", med, mn, flops=2.0 * M * N * Kd)


def bench_decode_gemm():
    shapes = [(12288, 4096, "store"), (22016, 4096, "gated"), (4096, 4096, "accum"), (4096, 11008, "accum"), (32000, 4096, "store")]
    ms = (8, 32, 64)
    if os.environ.get("BIG_SHAPES"):   # Llama-2-70B / 13B layer shapes at decode size: wave quantisation of the feature tiles
        shapes = [(57344, 8192, "gated"), (10240, 8192, "store"), (8192, 28672, "accum"), (27648, 5120, "gated"), (15360, 5120, "store")]
        ms = (8, )
    for M in ms:
        for (N, Kd, kind) in shapes:
            a, w = rand(M, Kd), rand(N, Kd, scale=Kd ** -0.5)
            wbytes = N * Kd * 2
            if kind == "accum":
                acc = torch.zeros(M, N, device=dev, dtype=torch.float32)
                for bn, splits in ((0, 0), (0, 1), (0, 2), (0, 4), (0, 8), (0, 13), (128, 4), (128, 9), (64, 4)):
                    med, mn = timeit(lambda: K.gemm_accum(a, w, acc, bn=bn, splits=splits))
                    report(f"decode_gemm_accum M{M} N{N} K{Kd} bn{bn} splits{splits}", med, mn, bytes_=wbytes)
            else:
                out = torch.empty(M, N // 2 if kind == "gated" else N, device=dev, dtype=torch.bfloat16)
                ws = torch.zeros(K.gemm_splitk_workspace_bytes(M, N), dtype=torch.uint8, device=dev)
                for bn, splits in ((0, 0), (0, 1), (0, 2), (0, 4), (0, 8), (128, 1), (128, 3), (64, 2)):
                    med, mn = timeit(lambda: K.gemm(a, w, act=3 if kind == "gated" else 0, gated=kind == "gated", out=out,
                                                    bn=bn, splits=splits, workspace=ws))
                    report(f"decode_gemm_{kind} M{M} N{N} K{Kd} bn{bn} splits{splits}", med, mn, bytes_=wbytes)
            out2 = torch.empty(M, N, device=dev, dtype=torch.bfloat16)
            med, mn = timeit(lambda: torch.matmul(a, w.t(), out=out2))
            report(f"cublas_decode M{M} N{N} K{Kd}", med, mn, bytes_=wbytes)


def bench_tiled():
    """row-major vs pre-tiled weights on the decode / short-prefill GEMM shapes of Llama-2-7B"""
    for M in (8, 64, 4096):
        for (N, Kd, kind) in [(12288, 4096, "store"), (22016, 4096, "gated"), (4096, 4096, "accum"), (4096, 11008, "accum")]:
            a, w = rand(M, Kd), rand(N, Kd, scale=Kd ** -0.5)
            wt = K.TiledWeight.from_dense(w)
            wbytes = N * Kd * 2
            for name, ww in (("rowmajor", w), ("tiled", wt)):
                if kind == "accum":
                    acc = torch.zeros(M, N, device=dev, dtype=torch.float32)
                    med, mn = timeit(lambda: K.gemm_accum(a, ww, acc))
                else:
                    out = torch.empty(M, N // 2 if kind == "gated" else N, device=dev, dtype=torch.bfloat16)
                    ws = torch.zeros(K.gemm_splitk_workspace_bytes(min(M, 128), N), dtype=torch.uint8, device=dev)
                    med, mn = timeit(lambda: K.gemm(a, ww, act=3 if kind == "gated" else 0, gated=kind == "gated", out=out,
                                                    workspace=ws if M <= 128 else None))
                if M <= 128:
                    report(f"{name} {kind} M{M} N{N} K{Kd}", med, mn, bytes_=wbytes)
                else:
                    report(f"{name} {kind} M{M} N{N} K{Kd}", med, mn, flops=2.0 * M * N * Kd)


def bench_restore():
    for (T, H, n_kv, d) in [(4096, 4096, 32, 128), (4096, 5120, 40, 128), (1024, 4096, 32, 128)]:
        bs = 64
        nb = T // bs + 8
        lat = rand(T, H)
        w_kv = rand(2 * n_kv * d, H, scale=H ** -0.5)
        cache = torch.zeros(nb, 2, n_kv, bs, d, device=dev, dtype=torch.bfloat16)
        pos = torch.arange(T, device=dev, dtype=torch.int32)
        slot = torch.arange(T, device=dev, dtype=torch.int32)
        inv = ref.default_inv_freq(d, device=dev)
        med, mn = timeit(lambda: K.restore_kv_proj(lat, w_kv, cache, pos, slot, inv, n_kv, d, bs))
        report(f"restore_kv_proj T{T} H{H} kv{n_kv}x{d}", med, mn, flops=2.0 * T * H * 2 * n_kv * d,
               pcie_equiv_ms_at_55GBs=round(T * H * 2 / 55e6, 3))
        # the reference's formulation: full QKV GEMM (cuBLAS) then a RoPE/KV-copy kernel
        w_qkv = rand(3 * n_kv * d, H, scale=H ** -0.5)
        qkv = torch.empty(T, 3 * n_kv * d, device=dev, dtype=torch.bfloat16)
        q_out = torch.empty(T, n_kv * d, device=dev, dtype=torch.bfloat16)

        def ref_way():
            torch.matmul(lat, w_qkv.t(), out=qkv)
            K.rope_kv_append(qkv, q_out, cache, pos, slot, inv, n_kv, n_kv, d, d, bs)
        med, mn = timeit(ref_way)
        report(f"ref_style_restore(cublas full QKV + rope kernel) T{T} H{H}", med, mn, flops=2.0 * T * H * 3 * n_kv * d)


def bench_attention_decode():
    d, bs = 128, 64
    shapes = [(8, 4096, 32, 32), (16, 4096, 32, 32), (64, 4096, 32, 32), (8, 16384, 32, 32), (64, 4096, 8, 1), (1, 4096, 32, 32), (4, 4096, 32, 32)]
    sweep = (1, 2, 4, 8, 16)
    if os.environ.get("ATTN_SWEEP"):   # wave-quantisation study: more CTAs than resident slots, every split count
        shapes, sweep = [(8, 4096, 32, 32), (8, 4096, 40, 40), (12, 4096, 32, 32)], tuple(range(1, 11))
    for (n_seqs, ctx, nq, nkv) in shapes:
        blocks_per_seq = (ctx + 1 + bs - 1) // bs
        nb = n_seqs * blocks_per_seq
        cache = (torch.randn(nb, 2, nkv, bs, d, device=dev) * 0.3).to(torch.bfloat16)
        tables = torch.arange(nb, device=dev, dtype=torch.int32).view(n_seqs, blocks_per_seq)
        items = torch.tensor([(s, s, 1, ctx) for s in range(n_seqs)], dtype=torch.int32, device=dev)
        q = rand(n_seqs, nq, d)
        out = torch.empty_like(q)
        G = nq // nkv
        kv_bytes = n_seqs * (ctx + 1) * 2 * nkv * d * 2
        for splits in sweep:
            if n_seqs * nkv * splits > 148 * 32:
                continue
            ws = torch.zeros(max(16, K.splitkv_workspace_bytes(n_seqs, nq, nkv, d, splits, G)), dtype=torch.uint8, device=dev)
            med, mn = timeit(lambda: K.paged_attention_splitkv(q, out, cache, tables, items, n_seqs, G, nq, nkv, d, bs, splits,
                                                                1 / math.sqrt(d), ws))
            report(f"attn_decode seqs{n_seqs} ctx{ctx} heads{nq}/{nkv} splits{splits} warps{os.environ.get('HDS_ATTN_WARPS', 'auto')}", med, mn, bytes_=kv_bytes)


def bench_attention_prefill():
    d, bs, nq, nkv = 128, 64, 32, 32
    for (n_seqs, ctx, n_new) in [(1, 0, 4096), (1, 4096, 128), (8, 4096, 64)]:
        total = ctx + n_new
        blocks_per_seq = (total + bs - 1) // bs
        nb = n_seqs * blocks_per_seq
        cache = (torch.randn(nb, 2, nkv, bs, d, device=dev) * 0.3).to(torch.bfloat16)
        tables = torch.arange(nb, device=dev, dtype=torch.int32).view(n_seqs, blocks_per_seq)
        its = []
        row = 0
        for s in range(n_seqs):
            for off in range(0, n_new, 128):
                n = min(128, n_new - off)
                its.append((s, row, n, ctx + off))
                row += n
        items = torch.tensor(its, dtype=torch.int32, device=dev)
        q = rand(row, nq, d)
        out = torch.empty_like(q)
        flops = 0.0
        for (_, _, n, p0) in its:
            flops += 4.0 * nq * d * n * (p0 + (n + 1) / 2)
        for splits in (1, 4, 8, 16):
            if len(its) * nq * splits > 148 * 12:
                continue
            ws = torch.zeros(max(16, K.prefill_workspace_bytes(len(its), nq, d, splits)), dtype=torch.uint8, device=dev)
            med, mn = timeit(lambda: K.paged_attention_prefill(q, out, cache, tables, items, len(its), nq, nkv, d, bs,
                                                               1 / math.sqrt(d), n_splits=splits, workspace=ws))
            report(f"attn_prefill seqs{n_seqs} ctx{ctx} new{n_new} splits{splits}", med, mn, flops=flops)


def bench_norm():
    for rows in (64, 4096):
        x, g = rand(rows, 4096), rand(4096)
        out, res = torch.empty_like(x), torch.empty_like(x)
        delta = torch.zeros(rows, 4096, device=dev)
        med, mn = timeit(lambda: K.norm(x, g, out, delta=delta, zero_delta=True, res_out=res, eps=1e-5))
        report(f"pre_rmsnorm_f32delta rows{rows} cols4096", med, mn, bytes_=rows * 4096 * (2 + 4 + 4 + 2 + 2))


def bench_gemm_fp8():
    for (M, N, Kd) in [(4096, 4096, 4096), (4096, 12288, 4096), (4096, 4096, 11008), (8192, 8192, 8192), (32768, 12288, 4096)]:
        a, w = rand(M, Kd), rand(N, Kd, scale=Kd ** -0.5)
        aq, a_s = K.quantize_fp8_rows(a)
        wq, w_s = K.quantize_fp8_rows(w)
        out = torch.empty(M, N, device=dev, dtype=torch.bfloat16)
        for bn in (128, 256, 1256):
            med, mn = timeit(lambda: K.gemm_fp8(aq, a_s, wq, w_s, out, bn=bn))
            report(f"gemm_fp8{'_2sm' if bn > 1000 else ''} M{M} N{N} K{Kd} bn{bn % 1000}", med, mn, flops=2.0 * M * N * Kd)
        med, mn = timeit(lambda: K.quantize_fp8_rows(a, aq, a_s))
        report(f"quantize_fp8_rows M{M} K{Kd}", med, mn, bytes_=M * Kd * 3)
        med, mn = timeit(lambda: K.gemm(a, w, out=out, bn=256))
        report(f"gemm_bf16 M{M} N{N} K{Kd} bn256", med, mn, flops=2.0 * M * N * Kd)


if __name__ == "__main__":
    which = sys.argv[1:] or ["gemm", "decode_gemm", "restore", "attn_decode", "attn_prefill", "norm"]
    table = {"gemm": bench_gemm, "decode_gemm": bench_decode_gemm, "restore": bench_restore,
             "attn_decode": bench_attention_decode, "tiled": bench_tiled, "attn_prefill": bench_attention_prefill, "norm": bench_norm, "gemm_fp8": bench_gemm_fp8}
    for w in which:
        table[w]()
    os.makedirs("gpurun_out", exist_ok=True)
    with open("gpurun_out/kernel_bench.jsonl", "a") as f:
        for r in results:
            f.write(json.dumps(r) + "\n")
