"""``hds_report``: environment and native-op status (ref: deepspeed/env_report.py)."""
from __future__ import annotations

import os
import shutil
import subprocess
import sys

GREEN, RED, YELLOW, END = "\033[92m", "\033[91m", "\033[93m", "\033[0m"
OKAY, NO = f"{GREEN}[OKAY]{END}", f"{RED}[NO]{END}"


def op_report(verbose: bool = True):
    from .ops import build
    rows = []
    for name, path, tool in (("sm_100a kernel library (libhds_b200.so)", build.CUDA_LIB, "nvcc"),
                             ("host runtime (libhds_host.so)", build.HOST_LIB, "g++")):
        rows.append((name, path.exists(), shutil.which(tool) is not None or os.path.exists("/usr/local/cuda/bin/" + tool)))
    print("-" * 74)
    print("native op report")
    print("-" * 74)
    print(f"{'op name':<45}{'built':<12}{'buildable here':<12}")
    for name, built, can in rows:
        print(f"{name:<45}{OKAY if built else NO:<21}{OKAY if can else NO}")
    # one line per op builder: pre-built in the in-tree library / can be JIT-built here, and why not if not
    from .ops.op_builder import ALL_OPS
    print("-" * 74)
    print(f"{'op builder':<30}{'pre-built':<12}{'compatible':<12}translation units")
    for name, b in sorted(ALL_OPS.items()):
        problems = b.compatibility_problems()
        units = ", ".join(p.name for p in b.all_sources()) or "(torch formulation)"
        print(f"{name:<30}{OKAY if b.installed() else NO:<21}{OKAY if not problems else NO:<21}{units}")
        if verbose:
            for w in problems:
                print(f"{'':<30}{YELLOW}{w}{END}")
        rows.append((name, b.installed(), not problems))
    if os.environ.get("HDS_REPORT_GDS", "0") == "1":      # opens the cuFile driver in a child process: slow where it blocks
        try:
            from .io import gds_available
            print(f"{'GPUDirect Storage (libcufile)':<45}{OKAY if gds_available() else NO}")
        except Exception:
            pass
    return rows


def debug_report():
    import torch
    from . import __version__
    nvcc = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    nvcc_ver = "n/a"
    if os.path.exists(nvcc):
        out = subprocess.run([nvcc, "--version"], capture_output=True, text=True).stdout
        nvcc_ver = out.strip().splitlines()[-2] if out else "n/a"
    rows = [("torch version", torch.__version__), ("torch cuda version", torch.version.cuda),
            ("cuda available", torch.cuda.is_available()),
            ("device", torch.cuda.get_device_name(0) if torch.cuda.is_available() else "cpu"),
            ("nvcc", nvcc_ver), ("hcache_deepspeed_b200", __version__), ("python", sys.version.split()[0]),
            ("nccl", ".".join(map(str, torch.cuda.nccl.version())) if torch.cuda.is_available() else "n/a")]
    print("-" * 74)
    print("general environment info")
    print("-" * 74)
    for k, v in rows:
        print(f"{k:.<30} {v}")
    return rows


def ninja_installed() -> bool:
    try:
        import ninja  # noqa: F401
        return True
    except ImportError:
        return shutil.which("ninja") is not None


def nvcc_version():
    """Release string of the nvcc that would build the ops (e.g. ``"12.9"``), or ``None`` without one."""
    nvcc = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    if not os.path.exists(nvcc):
        return None
    out = subprocess.run([nvcc, "-V"], capture_output=True, text=True).stdout
    for tok in out.split():
        if tok.endswith(",") and tok[:-1].replace(".", "").isdigit() and "release" in out.split()[out.split().index(tok) - 1]:
            return tok[:-1]
    return None


def installed_cann_path():
    """Ascend toolkits are not a target of this framework (sm_100a only); kept so callers of the reference's report work."""
    return None


def installed_cann_version():
    return None


def get_shm_size():
    """(bytes of /dev/shm, warnings): the host collectives engine and PyTorch data-loader workers live there."""
    try:
        st = os.statvfs("/dev/shm")
    except OSError:
        return None, ["/dev/shm is not available"]
    size = st.f_frsize * st.f_blocks
    warn = []
    if size < 512 * 1024 ** 2:
        warn.append(f" {YELLOW}[WARNING] /dev/shm size might be too small, if running in docker increase to at least "
                    f"--shm-size='1gb'{END}")
    return size, warn


def human_readable_size(size) -> str:
    units = ["B", "KB", "MB", "GB", "TB", "PB"]
    i, v = 0, float(size)
    while v >= 1024 and i < len(units) - 1:
        v /= 1024
        i += 1
    return f"{v:.2f} {units[i]}"


def parse_arguments(argv=None):
    import argparse
    p = argparse.ArgumentParser(description="environment and native-op report")
    p.add_argument("--hide_operator_status", action="store_true", help="skip the op compatibility table")
    p.add_argument("--hide_errors_and_warnings", action="store_true", help="no per-op explanations")
    return p.parse_args(argv)


def main(hide_operator_status: bool = False, hide_errors_and_warnings: bool = False):
    if not hide_operator_status:
        op_report(verbose=not hide_errors_and_warnings)
    rows = debug_report()
    size, warn = get_shm_size()
    if size is not None:
        print(f"{'shared memory (/dev/shm) size':.<30} {human_readable_size(size)}")
    print(f"{'ninja':.<30} {OKAY if ninja_installed() else NO}")
    if not hide_errors_and_warnings:
        for w in warn:
            print(w)
    return rows


def cli_main(argv=None):
    args = parse_arguments(argv)
    main(hide_operator_status=args.hide_operator_status, hide_errors_and_warnings=args.hide_errors_and_warnings)


if __name__ == "__main__":
    cli_main()
