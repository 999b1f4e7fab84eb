"""``hds_io``: file I/O micro-benchmark over the aio engine, and a small sweep to size the NVMe tier
(ref: deepspeed/nvme/ds_aio_basic.py, perf_run_sweep.py, bin/ds_io, bin/ds_nvme_tune)."""
from __future__ import annotations

import argparse
import itertools
import json
import os
import time

import torch

from ..io.aio import AsyncIOHandle
from ..utils.memory import pinned_empty


def run_once(path: str, nbytes: int, read: bool, block_size: int, queue_depth: int, threads: int, use_direct: bool,
             loops: int = 3) -> dict:
    h = AsyncIOHandle(block_size=block_size, queue_depth=queue_depth, intra_op_parallelism=threads, use_direct=use_direct)
    buf = pinned_empty((nbytes, ), torch.uint8)
    if read and not os.path.exists(path):
        buf.random_(0, 255)
        h.sync_pwrite(buf, path)
    best = 0.0
    for _ in range(loops):
        t0 = time.perf_counter()
        (h.sync_pread if read else h.sync_pwrite)(buf, path)
        dt = time.perf_counter() - t0
        best = max(best, nbytes / dt / 1e9)
    return {"op": "read" if read else "write", "GB/s": round(best, 3), "block_size": block_size,
            "queue_depth": queue_depth, "threads": threads, "o_direct": use_direct, "bytes": nbytes}


def sweep(folder: str, nbytes: int, read: bool):
    path = os.path.join(folder, "hds_io_sweep.bin")
    results = []
    for bs, th, direct in itertools.product((1 << 20, 4 << 20, 16 << 20), (1, 4, 8, 16), (False, True)):
        try:
            results.append(run_once(path, nbytes, read, bs, 32, th, direct, loops=2))
        except OSError as e:
            results.append({"error": str(e), "block_size": bs, "threads": th, "o_direct": direct})
    return sorted(results, key=lambda r: -r.get("GB/s", 0))


def main(argv=None):
    p = argparse.ArgumentParser(description="file I/O benchmark over the hcache_deepspeed_b200 aio engine")
    p.add_argument("--folder", required=True)
    p.add_argument("--io_size", type=int, default=1 << 30)
    p.add_argument("--read", action="store_true")
    p.add_argument("--block_size", type=int, default=4 << 20)
    p.add_argument("--queue_depth", type=int, default=32)
    p.add_argument("--threads", type=int, default=8)
    p.add_argument("--use_direct", action="store_true")
    p.add_argument("--sweep", action="store_true")
    a = p.parse_args(argv)
    os.makedirs(a.folder, exist_ok=True)
    if a.sweep:
        for r in sweep(a.folder, a.io_size, a.read):
            print(json.dumps(r))
    else:
        print(json.dumps(run_once(os.path.join(a.folder, "hds_io.bin"), a.io_size, a.read, a.block_size, a.queue_depth,
                                  a.threads, a.use_direct)))


if __name__ == "__main__":
    main()
