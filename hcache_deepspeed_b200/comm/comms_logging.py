"""Per-collective statistics (ref: ``deepspeed/utils/comms_logging.py:32-178``): count, total/avg latency,
algorithmic and bus bandwidth per op and message size."""
from __future__ import annotations

import math
from typing import Dict, List


def get_caller_func(frame: int = 3) -> str:
    import sys
    return sys._getframe(frame).f_code.co_name


def convert_size(size_bytes: int) -> str:
    if size_bytes == 0:
        return "0B"
    names = ("B", "KB", "MB", "GB", "TB", "PB")
    i = int(math.floor(math.log(size_bytes, 1024)))
    return "%s %s" % (round(size_bytes / math.pow(1024, i), 2), names[i])


def calc_bw_log(comm_op: str, size: int, duration_ms: float, n: int):
    """(algbw, busbw) in Gbps given bytes, milliseconds and world size — NCCL-tests conventions."""
    duration_s = max(duration_ms, 1e-9) / 1e3
    if comm_op in ("all_to_all_single", "all_to_all"):
        tput = size / duration_s
        busbw = tput * ((n - 1) / n)
    elif comm_op in ("all_gather", "all_gather_into_tensor", "reduce_scatter", "reduce_scatter_tensor"):
        size *= n
        tput = size / duration_s
        busbw = tput * ((n - 1) / n)
    elif comm_op in ("all_reduce", "all_reduce_coalesced", "inference_all_reduce"):
        tput = size * 2 / duration_s
        busbw = (size / duration_s) * (2 * (n - 1) / n)
    else:  # send/recv/broadcast/reduce/gather/scatter/barrier
        tput = size / duration_s
        busbw = tput
    return tput * 8 / 1e9, busbw * 8 / 1e9


class CommsLogger:

    def __init__(self):
        self.comms_dict: Dict[str, Dict[int, List]] = {}
        self.verbose = False
        self.debug = False
        self.prof_ops: List[str] = []
        self.prof_all = True
        self.enabled = False

    def configure(self, comms_config=None, enabled=None, prof_all=None, prof_ops=None, verbose=None, debug=None):
        if comms_config is not None:
            # either the wrapper (``DeepSpeedCommsConfig``: comms_logger_enabled + comms_logger) or the ``comms_logger`` block
            # itself (``CommsLoggerConfig``, what the training config carries)
            cl = getattr(comms_config, "comms_logger", None)
            if cl is None and hasattr(comms_config, "prof_all"):
                cl = comms_config
                self.enabled = bool(cl.enabled)
            else:
                self.enabled = bool(getattr(comms_config, "comms_logger_enabled", self.enabled)) and \
                    bool(getattr(cl, "enabled", True))
            if cl is not None:
                self.verbose, self.debug = cl.verbose, cl.debug
                self.prof_ops, self.prof_all = cl.prof_ops, cl.prof_all
        for name, val in (("enabled", enabled), ("prof_all", prof_all), ("prof_ops", prof_ops), ("verbose", verbose),
                          ("debug", debug)):
            if val is not None:
                setattr(self, name, val)

    def start_profiling_comms(self):
        self.prof_all = True

    def stop_profiling_comms(self):
        self.prof_all = False

    def start_profiling_op(self, op_name_list):
        self.prof_ops = list(set(self.prof_ops) | set(op_name_list))

    def stop_profiling_op(self, op_name_list):
        self.prof_ops = [op for op in self.prof_ops if op not in op_name_list]

    def append(self, raw_name: str, record_name: str, latency_ms: float, msg_size: int, world_size: int = 1):
        algbw, busbw = calc_bw_log(raw_name, msg_size, latency_ms, max(1, world_size))
        per_op = self.comms_dict.setdefault(record_name, {})
        if msg_size in per_op:
            per_op[msg_size][0] += 1
            per_op[msg_size][1].append(latency_ms)
            per_op[msg_size][2].append(algbw)
            per_op[msg_size][3].append(busbw)
        else:
            per_op[msg_size] = [1, [latency_ms], [algbw], [busbw]]
        if self.verbose:
            from ..utils.logging import log_dist
            log_dist(f"comm op: {record_name} | time (ms): {latency_ms:.2f} | msg size: {convert_size(msg_size)} | "
                     f"algbw (Gbps): {algbw:.2f} | busbw (Gbps): {busbw:.2f}", [0])

    def log_all(self, print_log: bool = True, show_straggler: bool = False):
        lines = [f"{'Comm. Op': <20}{'Message Size': <20}{'Count': <20}{'Total Latency(ms)': <20}"
                 f"{'Avg Latency(ms)': <20}{'tput_avg (Gbps)': <20}{'busbw_avg (Gbps)': <20}"]
        for record_name, sizes in self.comms_dict.items():
            lines.append(record_name)
            for msg_size, (count, lats, algs, buss) in sorted(sizes.items()):
                total = sum(lats)
                lines.append(f"{' ': <20}{convert_size(msg_size): <20}{count: <20}{total: <20.2f}"
                             f"{total / count: <20.2f}{sum(algs) / count: <20.2f}{sum(buss) / count: <20.2f}")
        if show_straggler:
            import torch
            import torch.distributed as dist
            lines.append("_______________________________ Straggler effect (min over ranks vs local) _______________")
            for record_name, sizes in self.comms_dict.items():
                for msg_size, (count, lats, _, _) in sorted(sizes.items()):
                    t = torch.tensor(lats)
                    mn = t.clone()
                    if dist.is_initialized():
                        dist.all_reduce(mn, op=dist.ReduceOp.MIN)
                    lines.append(f"{record_name: <20}{convert_size(msg_size): <20}{count: <20}"
                                 f"{float((t - mn).sum()): <20.2f}")
        out = "\n".join(lines)
        if print_log:
            print(out)
        return out
