"""In-tree native build: nvcc for the sm_100a kernel library, g++ for the host runtime library.

Both libraries are plain C-ABI shared objects loaded with ctypes (no torch headers, so a full rebuild
takes well under a minute and the .so files travel with the source tree).

The reference builds its ops through ``op_builder/builder.py:523-607`` (``OpBuilder.load`` /
``jit_load`` -> ``torch.utils.cpp_extension``) into a per-user JIT cache; here the artefacts are
deliberately kept inside the package (``ops/lib``).
"""
from __future__ import annotations

import os
import shutil
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor
from pathlib import Path

OPS_DIR = Path(__file__).resolve().parent
CSRC = OPS_DIR / "csrc"
HOST_SRC = OPS_DIR / "csrc_host"
LIB_DIR = OPS_DIR / "lib"
OBJ_DIR = OPS_DIR / "lib" / "obj"
# public headers: <repo>/include in a source checkout, <package>/include in an installed tree (setup.py copies them there)
REPO_INCLUDE = OPS_DIR.parent / "include" if (OPS_DIR.parent / "include" / "hds_b200.h").exists() else OPS_DIR.parent.parent / "include"

CUDA_LIB = LIB_DIR / "libhds_b200.so"
HOST_LIB = LIB_DIR / "libhds_host.so"

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-lineinfo", "-O3", "-std=c++17", "--expt-relaxed-constexpr",
    "-Xcompiler", "-fPIC", "-Xcompiler", "-fvisibility=default",
    "-diag-suppress", "177",
]


def _nvcc() -> str:
    cand = os.environ.get("NVCC") or shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    if not os.path.exists(cand):
        raise RuntimeError("nvcc not found: cannot build the sm_100a kernel library")
    return cand


def _newer(src_paths, target: Path) -> bool:
    if not target.exists():
        return True
    t = target.stat().st_mtime
    return any(Path(p).stat().st_mtime > t for p in src_paths)


def _run(cmd) -> None:
    res = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    if res.returncode != 0:
        raise RuntimeError("build command failed:\n  " + " ".join(map(str, cmd)) + "\n" + res.stdout)


def build_cuda(force: bool = False, verbose: bool = False) -> Path:
    """Compile every ``csrc/*.cu`` for sm_100a and link ``libhds_b200.so``."""
    LIB_DIR.mkdir(exist_ok=True)
    OBJ_DIR.mkdir(exist_ok=True)
    sources = sorted(CSRC.glob("*.cu"))
    headers = sorted(CSRC.glob("*.cuh")) + sorted(REPO_INCLUDE.glob("*.h"))
    nvcc = _nvcc()
    jobs = []
    for src in sources:
        obj = OBJ_DIR / (src.stem + ".o")
        if force or _newer([src, *headers], obj):
            jobs.append([nvcc, *NVCC_FLAGS, "-I", str(REPO_INCLUDE), "-c", str(src), "-o", str(obj)])
    if jobs:
        if verbose:
            print(f"[hds build] compiling {len(jobs)} CUDA translation unit(s) for sm_100a", file=sys.stderr)
        with ThreadPoolExecutor(max_workers=min(len(jobs), os.cpu_count() or 4)) as pool:
            list(pool.map(_run, jobs))
    objs = [OBJ_DIR / (s.stem + ".o") for s in sources]
    if force or jobs or _newer(objs, CUDA_LIB):
        # static cudart (nvcc default): no runtime dependency beyond the driver; streams and pointers
        # created by torch's runtime are driver handles and are valid here
        _run([nvcc, "-shared", "-o", str(CUDA_LIB), *map(str, objs), "-gencode", "arch=compute_100a,code=sm_100a"])
    return CUDA_LIB


def build_host(force: bool = False, verbose: bool = False) -> Path:
    """Compile the host runtime (block allocator, batch builder, async file IO) with g++."""
    LIB_DIR.mkdir(exist_ok=True)
    sources = sorted(HOST_SRC.glob("*.cpp"))
    headers = sorted(HOST_SRC.glob("*.h")) + sorted(REPO_INCLUDE.glob("*.h"))
    if not sources:
        return HOST_LIB
    if force or _newer([*sources, *headers], HOST_LIB):
        if verbose:
            print("[hds build] compiling host runtime", file=sys.stderr)
        _run(["g++", "-O2", "-ftree-vectorize", "-std=c++17", "-fPIC", "-shared", "-pthread", "-I", str(REPO_INCLUDE), "-o", str(HOST_LIB),
              *map(str, sources), "-lrt"])
    return HOST_LIB


def build_all(force: bool = False, verbose: bool = True) -> None:
    build_host(force=force, verbose=verbose)
    build_cuda(force=force, verbose=verbose)


if __name__ == "__main__":
    build_all(force="--force" in sys.argv, verbose=True)
    print(CUDA_LIB, HOST_LIB)
