"""NUMA helpers for the launcher's ``--bind_cores_to_rank`` (``deepspeed/utils/numa.py:24-200``): discover nodes and
build the ``numactl`` prefix that pins a local rank to its share of the cores."""
from __future__ import annotations

import os
import re
import subprocess
from typing import List, Optional, Tuple


def get_numa_cores() -> List[List[int]]:
    try:
        out = subprocess.check_output(["numactl", "--hardware"], stderr=subprocess.DEVNULL).decode()
        nodes = []
        for line in out.splitlines():
            m = re.match(r"node (\d+) cpus: (.*)", line)
            if m:
                nodes.append([int(c) for c in m.group(2).split()])
        if nodes:
            return nodes
    except (OSError, subprocess.CalledProcessError):
        pass
    return [list(range(os.cpu_count() or 1))]


def parse_range_list(s: str) -> List[int]:
    """"0-3,8,10-11" -> [0,1,2,3,8,10,11] (strictly increasing)."""
    out = []
    for part in s.split(","):
        if "-" in part:
            a, b = part.split("-")
            a, b = int(a), int(b)
            if b < a:
                raise ValueError(f"invalid range {part}")
            out += list(range(a, b + 1))
        else:
            out.append(int(part))
    if any(out[i] >= out[i + 1] for i in range(len(out) - 1)):
        raise ValueError("core list must be strictly increasing")
    return out


def get_numactl_cmd(bind_core_list: Optional[str], num_local_procs: int, local_rank: int) -> Tuple[int, List[str]]:
    cores = parse_range_list(bind_core_list) if bind_core_list else list(range(os.cpu_count() or 1))
    per = max(1, len(cores) // num_local_procs)
    mine = cores[per * local_rank:per * (local_rank + 1)]
    cmd = ["numactl"]
    nodes = get_numa_cores()
    owning = [i for i, n in enumerate(nodes) if set(mine) <= set(n)]
    if owning:
        cmd += ["-m", str(owning[0])]
    cmd += ["-C", ",".join(map(str, mine))]
    return per, cmd


def bind_to_gpu_numa(gpu_index: int) -> List[int]:
    """Pin this process to the CPUs that are local to ``gpu_index`` (NVML's CPU affinity mask), so pinned host buffers
    it allocates afterwards are first-touched on the GPU's own NUMA node: with 8 ranks streaming hidden states at once
    the host links only reach their rate when every copy reads node-local memory. Returns the CPU list ([] if NVML or
    the affinity call is unavailable; the process is left untouched then)."""
    try:
        import pynvml
        pynvml.nvmlInit()
        h = pynvml.nvmlDeviceGetHandleByIndex(gpu_index)
        n_words = (os.cpu_count() + 63) // 64
        mask = pynvml.nvmlDeviceGetCpuAffinity(h, n_words)
        cpus = [w * 64 + b for w, word in enumerate(mask) for b in range(64) if (word >> b) & 1]
        cpus = [c for c in cpus if c in os.sched_getaffinity(0)]
        if cpus:
            os.sched_setaffinity(0, cpus)
        return cpus
    except Exception:
        return []


def parse_range(rng: str) -> range:
    """``"5"`` -> range(5, 6); ``"2-4"`` -> range(2, 5)."""
    try:
        v = int(rng)
        return range(v, v + 1)
    except ValueError:
        parts = rng.split("-")
        if len(parts) != 2:
            raise ValueError(f"Bad range: '{rng}', range must be either a number or two number separated by dash")
        lo, hi = int(parts[0]), int(parts[1])
        if lo > hi:
            raise ValueError(f"Bad range: '{rng}', range end must larger than or equal to start")
        return range(lo, hi + 1)


def check_for_numactl_pkg() -> bool:
    """Is the ``numactl`` tool usable here (the launcher's ``--bind_cores_to_rank`` needs it)? The binary on PATH is what
    counts; when it is missing the distribution's package manager is asked so the message can name the install command."""
    import shutil
    import subprocess
    if shutil.which("numactl") is not None:
        return True
    for mgr, (flag, tool) in {"dpkg": ("-l", "apt"), "pacman": ("-Q", "pacman"), "rpm": ("-q", "yum")}.items():
        if shutil.which(mgr) is None:
            continue
        ok = subprocess.run([mgr, flag, "numactl"], stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL).returncode == 0
        if not ok:
            print(f"please install the numactl package with {tool}")
        return ok
    return False
