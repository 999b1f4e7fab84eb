"""Benchmark through the handle API (``deepspeed/nvme/ds_aio_handle.py``): one aio handle per process, pinned buffer,
timed ``pread`` / ``pwrite`` loops. Returns GB/s."""
import os

from .ds_io import run_once


def aio_handle_multiprocessing(args, read_op: bool) -> float:
    path = os.path.join(args.folder, "ds_aio_bench.bin")
    res = run_once(path, args.io_size, read_op, args.block_size, args.queue_depth, args.threads, args.use_direct, loops=args.loops)
    return float(res["GB/s"])
