"""Collective benchmark behind ``bin/hds_bench`` (role of the reference's ``benchmarks/communication/run_all.py`` behind
``bin/ds_bench``): sweeps message sizes for all_reduce / all_gather / reduce_scatter / all_to_all / broadcast / pt2pt through
the ``hcache_deepspeed_b200.comm`` facade and prints size, latency, algorithmic and bus bandwidth per collective.

Timing: CUDA events on the launching stream after warm-up on a GPU, ``perf_counter`` on the host; the reported time is the
MAX over ranks. Bus-bandwidth factors are the usual ones (all_reduce 2(n-1)/n, all_gather / reduce_scatter / all_to_all
(n-1)/n, broadcast and pt2pt 1)."""
from __future__ import annotations

import argparse
import os
import sys
import time
from typing import Callable, Dict, List

import torch

OPS = ("all_reduce", "all_gather", "reduce_scatter", "all_to_all", "broadcast", "pt2pt")


def bus_factor(op: str, n: int) -> float:
    if op == "all_reduce":
        return 2.0 * (n - 1) / n
    if op in ("all_gather", "reduce_scatter", "all_to_all"):
        return (n - 1) / n
    return 1.0


def _sizes(min_bytes: int, max_bytes: int) -> List[int]:
    out, b = [], max(4, min_bytes)
    while b <= max_bytes:
        out.append(b)
        b *= 2
    return out


def _make_op(dist, op: str, numel: int, dtype, device, rank: int, world: int) -> Callable[[], None]:
    per = max(1, numel // world)
    if op == "all_reduce":
        x = torch.ones(numel, dtype=dtype, device=device)
        return lambda: dist.all_reduce(x)
    if op == "all_gather":
        src, dst = torch.ones(per, dtype=dtype, device=device), torch.empty(per * world, dtype=dtype, device=device)
        return lambda: dist.all_gather_into_tensor(dst, src)
    if op == "reduce_scatter":
        src, dst = torch.ones(per * world, dtype=dtype, device=device), torch.empty(per, dtype=dtype, device=device)
        return lambda: dist.reduce_scatter_tensor(dst, src)
    if op == "all_to_all":
        src, dst = torch.ones(per * world, dtype=dtype, device=device), torch.empty(per * world, dtype=dtype, device=device)
        return lambda: dist.all_to_all_single(dst, src)
    if op == "broadcast":
        x = torch.ones(numel, dtype=dtype, device=device)
        return lambda: dist.broadcast(x, 0)
    x = torch.ones(numel, dtype=dtype, device=device)

    def pt2pt():
        if world < 2:
            return
        if rank == 0:
            dist.send(x, 1)
        elif rank == 1:
            dist.recv(x, 0)
    return pt2pt


def run(args, out=sys.stdout) -> List[Dict]:
    from .. import comm as dist
    dist.init_distributed(dist_backend=args.backend)
    rank, world = dist.get_rank(), dist.get_world_size()
    on_gpu = args.backend == "nccl" and torch.cuda.is_available()
    device = torch.device("cuda", int(os.environ.get("LOCAL_RANK", 0))) if on_gpu else torch.device("cpu")
    if on_gpu:
        torch.cuda.set_device(device)
    dtype = getattr(torch, args.dtype)
    esize = torch.empty((), dtype=dtype).element_size()
    rows = []
    for op in ([args.op] if args.op != "all" else OPS):
        if rank == 0:
            print(f"---- {op} ({world} ranks, {args.dtype}, backend {args.backend})", file=out)
            print(f"{'bytes':>14}{'latency us':>14}{'algbw GB/s':>14}{'busbw GB/s':>14}", file=out)
        for nbytes in _sizes(args.minsize, args.maxsize):
            fn = _make_op(dist, op, max(world, nbytes // esize), dtype, device, rank, world)
            for _ in range(args.warmups):
                fn()
            dist.barrier()
            if on_gpu:
                torch.cuda.synchronize()
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
                for _ in range(args.trials):
                    fn()
                e1.record()
                torch.cuda.synchronize()
                sec = e0.elapsed_time(e1) / 1e3 / args.trials
            else:
                t0 = time.perf_counter()
                for _ in range(args.trials):
                    fn()
                sec = (time.perf_counter() - t0) / args.trials
            t = torch.tensor([sec], dtype=torch.float64, device=device)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            sec = float(t.item())
            alg = nbytes / sec / 1e9
            row = {"op": op, "bytes": nbytes, "latency_us": sec * 1e6, "algbw_GBps": alg, "busbw_GBps": alg * bus_factor(op, world)}
            rows.append(row)
            if rank == 0:
                print(f"{nbytes:>14}{row['latency_us']:>14.1f}{alg:>14.2f}{row['busbw_GBps']:>14.2f}", file=out)
    return rows


def benchmark_parser() -> argparse.ArgumentParser:
    p = argparse.ArgumentParser(description="collective micro-benchmark (launch one process per rank, e.g. with bin/hds or torchrun)")
    p.add_argument("--op", default="all", choices=("all", ) + OPS)
    p.add_argument("--backend", default="nccl" if torch.cuda.is_available() else "gloo", choices=["nccl", "gloo", "ccl"],
                   help="ccl = gloo + this framework's shared-memory engine for host tensors")
    p.add_argument("--dtype", default="float32")
    p.add_argument("--minsize", type=int, default=1 << 10, help="smallest message in bytes")
    p.add_argument("--maxsize", type=int, default=1 << 26, help="largest message in bytes")
    p.add_argument("--trials", type=int, default=20)
    p.add_argument("--warmups", type=int, default=5)
    p.add_argument("--local_rank", type=int, default=int(os.environ.get("LOCAL_RANK", 0)))
    return p


def main(argv=None) -> int:
    args = benchmark_parser().parse_args(argv)
    needed = ("RANK", "WORLD_SIZE", "MASTER_ADDR", "MASTER_PORT")
    if not all(k in os.environ for k in needed):   # not under a launcher: run as a single rank
        os.environ.setdefault("RANK", "0")
        os.environ.setdefault("WORLD_SIZE", "1")
        os.environ.setdefault("LOCAL_RANK", "0")
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("MASTER_PORT", "29577")
    run(args)
    return 0


if __name__ == "__main__":
    sys.exit(main())
