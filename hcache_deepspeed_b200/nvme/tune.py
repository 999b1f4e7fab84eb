"""``hds_nvme_tune``: sweep the aio engine's knobs over a folder, log every run, and emit the ``"aio"`` config block
that maximises read + write bandwidth — what sizes the NVMe tier of the hidden-state store and of ZeRO-Infinity.

Reference: ``deepspeed/nvme/perf_run_sweep.py:24-320`` (sweep driver, one subprocess + log file per point),
``parse_nvme_stats.py:14-148`` (log scraping) and ``perf_generate_param.py:42-97`` (best-point selection). Here the
sweep runs in-process on the C++ aio engine and writes JSON-lines logs, so there is no text scraping step.
"""
from __future__ import annotations

import argparse
import itertools
import json
import os
from typing import Dict, Iterable, List, Optional, Tuple

from .ds_io import run_once

BENCH_LOG_DIR, READ_LOG_DIR, WRITE_LOG_DIR = "_aio_bench_logs", "read", "write"
READ_SPEED, WRITE_SPEED = "read_speed", "write_speed"
DEFAULT_SWEEP_CONFIG = {
    "block_size": ["1M", "8M"],
    "queue_depth": [32, 128],
    "io_parallel": [1, 8],
    "use_direct": [False, True],
}
_UNITS = {"K": 1 << 10, "M": 1 << 20, "G": 1 << 30}


def parse_size(v) -> int:
    """'8M' / '400M' / '1G' / 4096 -> bytes."""
    if isinstance(v, int):
        return v
    s = str(v).strip().upper().rstrip("B")
    return int(float(s[:-1]) * _UNITS[s[-1]]) if s[-1] in _UNITS else int(s)


def get_sweep_config_dict(path: Optional[str]) -> Dict[str, list]:
    if path is None:
        return dict(DEFAULT_SWEEP_CONFIG)
    with open(path) as f:
        return json.load(f)


def sweep_points(space: Dict[str, list]) -> List[Dict[str, object]]:
    keys = sorted(space)
    return [dict(zip(keys, vals)) for vals in itertools.product(*(space[k] for k in keys))]


def log_name(op: str, point: Dict[str, object]) -> str:
    return op + "".join(f"_{k}{point[k]}" for k in sorted(point)) + ".jsonl"


def run_sweep(nvme_dirs: Iterable[str], space: Dict[str, list], io_size: int, log_dir: str, read: bool = True,
              write: bool = True, loops: int = 1, verbose: bool = False) -> List[dict]:
    """Run every point for every op; append one JSON line per (point, folder) to ``<log_dir>/<op>/<point>.jsonl``."""
    results = []
    ops = [(WRITE_LOG_DIR, False)] * write + [(READ_LOG_DIR, True)] * read   # write first: reads reuse the file
    for sub, is_read in ops:
        os.makedirs(os.path.join(log_dir, sub), exist_ok=True)
        for point in sweep_points(space):
            for folder in nvme_dirs:
                os.makedirs(folder, exist_ok=True)
                path = os.path.join(folder, "hds_nvme_tune.bin")
                try:
                    r = run_once(path, io_size, is_read, parse_size(point.get("block_size", "1M")),
                                 int(point.get("queue_depth", 32)), int(point.get("io_parallel", 1)),
                                 bool(point.get("use_direct", False)), loops=max(1, loops))
                except OSError as e:  # e.g. O_DIRECT refused by the file system
                    r = {"op": sub, "error": str(e), "GB/s": 0.0}
                r.update(point=point, folder=folder)
                with open(os.path.join(log_dir, sub, log_name(sub, point)), "a") as f:
                    f.write(json.dumps(r) + "\n")
                if verbose:
                    print(json.dumps(r))
                results.append(r)
    for folder in nvme_dirs:
        try:
            os.remove(os.path.join(folder, "hds_nvme_tune.bin"))
        except OSError:
            pass
    return results


def get_sorted_results(log_dir: str, metric: str = READ_SPEED) -> Tuple[List[tuple], Dict[tuple, float]]:
    """Read one op's logs: ``{point key: GB/s summed over folders}`` and the keys sorted by bandwidth."""
    perf: Dict[tuple, float] = {}
    for name in sorted(os.listdir(log_dir)) if os.path.isdir(log_dir) else []:
        if not name.endswith(".jsonl"):
            continue
        with open(os.path.join(log_dir, name)) as f:
            for line in f:
                r = json.loads(line)
                key = tuple(sorted((k, str(v)) for k, v in r["point"].items()))
                perf[key] = perf.get(key, 0.0) + float(r.get("GB/s", 0.0))
    return sorted(perf, key=lambda k: perf[k]), perf


def convert_to_param(key: tuple) -> dict:
    p = dict(key)
    return {"block_size": parse_size(p.get("block_size", "1M")), "queue_depth": int(p.get("queue_depth", 32)),
            "intra_op_parallelism": int(p.get("io_parallel", 1)), "single_submit": False, "overlap_events": True,
            "use_gds": False, "use_direct": p.get("use_direct", "False") == "True"}


def generate_aio_param(read_log_dir: str, write_log_dir: str) -> dict:
    """The point with the best read + write bandwidth, as the ``"aio"`` section of a config."""
    _, rd = get_sorted_results(read_log_dir, READ_SPEED)
    _, wr = get_sorted_results(write_log_dir, WRITE_SPEED)
    keys = set(rd) & set(wr) if rd and wr else set(rd) | set(wr)
    assert keys, "no sweep results to choose from"
    best = max(keys, key=lambda k: rd.get(k, 0.0) + wr.get(k, 0.0))
    return {"aio": convert_to_param(best), "read_GBps": rd.get(best), "write_GBps": wr.get(best)}


def sweep_main(argv=None) -> dict:
    p = argparse.ArgumentParser(description="tune the aio engine for a set of NVMe folders")
    p.add_argument("--nvme_dir", nargs="+", required=True)
    p.add_argument("--sweep_config", default=None)
    p.add_argument("--no_read", action="store_true")
    p.add_argument("--no_write", action="store_true")
    p.add_argument("--io_size", default="400M")
    p.add_argument("--log_dir", default=BENCH_LOG_DIR)
    p.add_argument("--loops", type=int, default=1)
    p.add_argument("--verbose", action="store_true")
    a = p.parse_args(argv)
    run_sweep(a.nvme_dir, get_sweep_config_dict(a.sweep_config), parse_size(a.io_size), a.log_dir, not a.no_read,
              not a.no_write, a.loops, a.verbose)
    out = generate_aio_param(os.path.join(a.log_dir, READ_LOG_DIR), os.path.join(a.log_dir, WRITE_LOG_DIR))
    print(json.dumps(out, indent=3))
    return out


def generate_main(log_dir: str = BENCH_LOG_DIR) -> dict:
    out = generate_aio_param(os.path.join(log_dir, READ_LOG_DIR), os.path.join(log_dir, WRITE_LOG_DIR))
    print(json.dumps(out, indent=3))
    return out


if __name__ == "__main__":
    sweep_main()
