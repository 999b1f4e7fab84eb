"""Per-op builders: compatibility probing, on-demand (JIT) compilation and loading of one native op at a time.

Role of the reference's ``op_builder/builder.py`` (``OpBuilder.is_compatible / load / jit_load``, :523-607, and the
``CUDAOpBuilder`` arch / version checks). What is different here:

* the ops are plain C-ABI shared objects loaded with ctypes (no torch headers, no pybind), so a JIT build of one op is a
  few nvcc / g++ invocations and no ninja project;
* the default artefacts are the two in-tree libraries every op is linked into (``ops/lib/libhds_b200.so`` and
  ``libhds_host.so``, built by ``ops/build.py``). ``load()`` returns those when they exist and are not older than the
  op's sources; otherwise (or with ``HDS_FORCE_JIT=1``) it compiles THIS op's translation units plus the ones they need
  into ``ops/lib/jit/<op>-<hash>.so``, keyed by a hash of sources, headers, flags and compiler version, under a file lock;
* only one architecture exists (sm_100a), so arch-list handling collapses to "does this nvcc know sm_100a".

Environment: ``HDS_BUILD_OPS=0`` forbids JIT builds (load() then raises if the op is not pre-built),
``HDS_FORCE_JIT=1`` ignores the pre-built libraries, ``HDS_SKIP_CUDA_CHECK=1`` skips the nvcc-vs-torch CUDA version check,
``HDS_JIT_DIR`` moves the JIT cache, ``NVCC`` / ``CXX`` pick the compilers.
"""
from __future__ import annotations

import ctypes
import fcntl
import hashlib
import os
import re
import shutil
import subprocess
import tempfile
from functools import lru_cache
from pathlib import Path
from typing import Dict, List, Optional, Sequence, Tuple

from .. import build as _build

MIN_CUDA_FOR_SM100A = (12, 8)


class OpBuildError(RuntimeError):
    pass


# ---- tool-chain probes ------------------------------------------------------------------------------------------------
def _cxx() -> Optional[str]:
    return os.environ.get("CXX") or shutil.which("g++") or shutil.which("c++")


def _nvcc() -> Optional[str]:
    try:
        return _build._nvcc()
    except RuntimeError:
        return None


@lru_cache(maxsize=None)
def nvcc_version() -> Optional[Tuple[int, int]]:
    nvcc = _nvcc()
    if nvcc is None:
        return None
    out = subprocess.run([nvcc, "--version"], stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True).stdout
    m = re.search(r"release (\d+)\.(\d+)", out)
    return (int(m.group(1)), int(m.group(2))) if m else None


@lru_cache(maxsize=None)
def cxx_version() -> Optional[str]:
    cxx = _cxx()
    if cxx is None:
        return None
    return subprocess.run([cxx, "-dumpfullversion", "-dumpversion"], stdout=subprocess.PIPE, stderr=subprocess.STDOUT,
                          text=True).stdout.strip()


@lru_cache(maxsize=None)
def nvcc_knows_sm100a() -> bool:
    nvcc = _nvcc()
    if nvcc is None:
        return False
    out = subprocess.run([nvcc, "--list-gpu-arch"], stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True).stdout
    if "compute_100" in out:
        return True
    v = nvcc_version()
    return v is not None and v >= MIN_CUDA_FOR_SM100A


def _try_compile(code: str, compiler: Sequence[str], suffix: str, extra: Sequence[str] = ()) -> bool:
    with tempfile.TemporaryDirectory(prefix="hds_probe_") as tmp:
        src = Path(tmp) / f"probe{suffix}"
        src.write_text(code)
        res = subprocess.run([*compiler, *extra, "-c", str(src), "-o", str(Path(tmp) / "probe.o")], stdout=subprocess.PIPE,
                             stderr=subprocess.STDOUT)
        return res.returncode == 0


@lru_cache(maxsize=None)
def has_header(header: str, cuda: bool = False) -> bool:
    """Can ``#include <header>`` be compiled (by nvcc when ``cuda``, else by the host compiler)?"""
    comp = _nvcc() if cuda else _cxx()
    if comp is None:
        return False
    return _try_compile(f"#include <{header}>\nint hds_probe() {{ return 0; }}\n", [comp], ".cu" if cuda else ".cpp",
                        ["-std=c++17"])


@lru_cache(maxsize=None)
def has_function(name: str, headers: Tuple[str, ...] = (), libs: Tuple[str, ...] = ()) -> bool:
    """Does a call to ``name`` compile AND link against ``libs`` (the reference's ``has_function`` for libaio etc.)?"""
    cxx = _cxx()
    if cxx is None:
        return False
    inc = "".join(f"#include <{h}>\n" for h in headers)
    code = f"{inc}int main() {{ return reinterpret_cast<long>(&{name}) != 0 ? 0 : 1; }}\n"
    with tempfile.TemporaryDirectory(prefix="hds_probe_") as tmp:
        src = Path(tmp) / "probe.cpp"
        src.write_text(code)
        res = subprocess.run([cxx, "-std=c++17", str(src), "-o", str(Path(tmp) / "probe"), *[f"-l{lib}" for lib in libs]],
                             stdout=subprocess.PIPE, stderr=subprocess.STDOUT)
        return res.returncode == 0


def has_shared_library(*names: str) -> bool:
    for n in names:
        try:
            ctypes.CDLL(n)
            return True
        except OSError:
            continue
    return False


def torch_cuda_mismatch() -> Optional[str]:
    """A reason string when nvcc's CUDA major differs from the one torch was built with (the extensions share the
    driver with torch, not the runtime, so a minor difference is harmless)."""
    if os.environ.get("HDS_SKIP_CUDA_CHECK", "0") == "1":
        return None
    try:
        import torch
        tv = torch.version.cuda
    except Exception:
        return None
    nv = nvcc_version()
    if tv is None or nv is None:
        return None
    if int(tv.split(".")[0]) != nv[0]:
        return f"nvcc is CUDA {nv[0]}.{nv[1]} but torch was built with CUDA {tv} (set HDS_SKIP_CUDA_CHECK=1 to ignore)"
    return None


def find_library_file(*names: str) -> Optional[str]:
    """Path of a shared library WITHOUT loading it (some driver libraries block in their constructors)."""
    dirs = [d for d in os.environ.get("LD_LIBRARY_PATH", "").split(":") if d]
    dirs += ["/usr/local/cuda/lib64", "/usr/local/cuda/targets/x86_64-linux/lib", "/usr/lib/x86_64-linux-gnu", "/usr/lib64",
             "/usr/lib", "/lib/x86_64-linux-gnu"]
    for d in dirs:
        for n in names:
            for p in sorted(Path(d).glob(n + "*")) if Path(d).is_dir() else ():
                return str(p)
    return None


# ---- the builder ---------------------------------------------------------------------------------------------------
class OpBuilder:
    """One native op: its translation units, the units those need, the symbols it must export, and how to get them."""
    NAME = "op"
    LIB = "cuda"                      # "cuda": nvcc, sm_100a, linked into libhds_b200.so; "host": g++, libhds_host.so
    SOURCES: Tuple[str, ...] = ()     # file names under ops/csrc (cuda) or ops/csrc_host (host)
    DEPENDS: Tuple[str, ...] = ()     # further translation units whose symbols SOURCES reference
    SYMBOLS: Tuple[str, ...] = ()     # extern "C" entry points the op must provide
    HOST_LIBS: Tuple[str, ...] = ("rt", )

    def __init__(self, name: Optional[str] = None):
        self.name = name or self.NAME
        self.error_log: List[str] = []
        self._jit_handle = None

    # -- description ------------------------------------------------------------------------------------------------------
    def absolute_name(self) -> str:
        return f"hcache_deepspeed_b200.ops.{self.name}"

    def src_dir(self) -> Path:
        return _build.CSRC if self.LIB == "cuda" else _build.HOST_SRC

    def sources(self) -> List[str]:
        return [str(self.src_dir() / s) for s in self.SOURCES]

    def all_sources(self) -> List[Path]:
        seen, out = set(), []
        for s in (*self.SOURCES, *self.DEPENDS):
            if s not in seen:
                seen.add(s)
                out.append(self.src_dir() / s)
        return out

    def headers(self) -> List[Path]:
        pat = ("*.cuh", "*.h") if self.LIB == "cuda" else ("*.h", )
        hs = [p for g in pat for p in sorted(self.src_dir().glob(g))]
        return hs + sorted(_build.REPO_INCLUDE.glob("*.h"))

    def include_paths(self) -> List[str]:
        return [str(_build.REPO_INCLUDE)]

    def nvcc_args(self) -> List[str]:
        return list(_build.NVCC_FLAGS)

    def cxx_args(self) -> List[str]:
        return ["-O2", "-ftree-vectorize", "-std=c++17", "-fPIC", "-pthread"]

    def extra_ldflags(self) -> List[str]:
        return [f"-l{lib}" for lib in self.HOST_LIBS] if self.LIB == "host" else []

    # -- compatibility --------------------------------------------------------------------------------------------------
    def compatibility_problems(self) -> List[str]:
        """Reasons this op cannot be BUILT here (empty = can build). A pre-built library makes the op usable anyway."""
        why: List[str] = []
        for p in self.all_sources():
            if not p.exists():
                why.append(f"source file {p.name} is missing")
        if self.LIB == "cuda":
            if _nvcc() is None:
                why.append("nvcc not found (set NVCC or install the CUDA toolkit)")
            elif not nvcc_knows_sm100a():
                v = nvcc_version()
                why.append(f"nvcc {v[0]}.{v[1] if v else '?'} cannot target sm_100a (CUDA >= {MIN_CUDA_FOR_SM100A[0]}.{MIN_CUDA_FOR_SM100A[1]} needed)")
            else:
                mm = torch_cuda_mismatch()
                if mm:
                    why.append(mm)
        elif _cxx() is None:
            why.append("no host C++ compiler (g++) found")
        return why + self.extra_problems()

    def extra_problems(self) -> List[str]:
        return []

    def is_compatible(self, verbose: bool = False) -> bool:
        self.error_log = self.compatibility_problems()
        if verbose:
            for w in self.error_log:
                print(f"[hds op_builder] {self.name}: {w}")
        return not self.error_log or self.installed()

    # -- pre-built ------------------------------------------------------------------------------------------------------
    def prebuilt_path(self) -> Path:
        return _build.CUDA_LIB if self.LIB == "cuda" else _build.HOST_LIB

    def installed(self) -> bool:
        """The in-tree library exists and exports every symbol of this op."""
        path = self.prebuilt_path()
        if not path.exists():
            return False
        return not self.SYMBOLS or not self.missing_symbols(path)

    def missing_symbols(self, lib_path: Path) -> List[str]:
        have = exported_symbols(str(lib_path), lib_path.stat().st_mtime_ns)
        return [s for s in self.SYMBOLS if s not in have]

    def prebuilt_is_stale(self) -> bool:
        path = self.prebuilt_path()
        if not path.exists():
            return True
        t = path.stat().st_mtime
        return any(p.exists() and p.stat().st_mtime > t for p in (*self.all_sources(), *self.headers()))

    # -- JIT ------------------------------------------------------------------------------------------------------------
    def build_hash(self) -> str:
        h = hashlib.sha256()
        for p in (*self.all_sources(), *self.headers()):
            h.update(p.name.encode())
            h.update(p.read_bytes() if p.exists() else b"<missing>")
        flags = self.nvcc_args() if self.LIB == "cuda" else self.cxx_args()
        h.update(" ".join([*flags, *self.include_paths(), *self.extra_ldflags()]).encode())
        h.update(str(nvcc_version() if self.LIB == "cuda" else cxx_version()).encode())
        return h.hexdigest()[:16]

    @staticmethod
    def jit_dir() -> Path:
        return Path(os.environ.get("HDS_JIT_DIR") or (_build.LIB_DIR / "jit"))

    def jit_path(self) -> Path:
        return self.jit_dir() / f"{self.name}-{self.build_hash()}.so"

    def jit_build(self, verbose: bool = True) -> Path:
        """Compile this op's translation units (and the ones they need) into its own shared object; no-op when an object
        with the same hash is already there. Safe against concurrent ranks (file lock, atomic rename)."""
        problems = self.compatibility_problems()
        if problems:
            raise OpBuildError(f"cannot build op {self.name}: " + "; ".join(problems))
        target = self.jit_path()
        if target.exists():
            return target
        if os.environ.get("HDS_BUILD_OPS", "1") == "0":
            raise OpBuildError(f"op {self.name} is not built and HDS_BUILD_OPS=0 forbids building it now")
        target.parent.mkdir(parents=True, exist_ok=True)
        with open(target.parent / f".{self.name}.lock", "w") as lock:
            fcntl.flock(lock, fcntl.LOCK_EX)
            try:
                if target.exists():          # another rank built it while this one waited
                    return target
                if verbose:
                    print(f"[hds op_builder] building op {self.name} ({len(self.all_sources())} translation unit(s)) -> {target.name}")
                with tempfile.TemporaryDirectory(prefix=f"hds_jit_{self.name}_", dir=str(target.parent)) as tmp:
                    out = Path(tmp) / target.name
                    self._compile_and_link(Path(tmp), out)
                    os.replace(out, target)
                for old in target.parent.glob(f"{self.name}-*.so"):     # one artefact per op: drop earlier hashes
                    if old != target:
                        old.unlink(missing_ok=True)
            finally:
                fcntl.flock(lock, fcntl.LOCK_UN)
        return target

    def _compile_and_link(self, tmp: Path, out: Path) -> None:
        inc = [a for p in self.include_paths() for a in ("-I", p)]
        srcs = self.all_sources()
        if self.LIB == "host":
            _build._run([_cxx(), *self.cxx_args(), "-shared", *inc, "-o", str(out), *map(str, srcs), *self.extra_ldflags()])
            return
        nvcc = _nvcc()
        objs = []
        jobs = []
        for s in srcs:
            o = tmp / (s.stem + ".o")
            objs.append(o)
            jobs.append([nvcc, *self.nvcc_args(), *inc, "-c", str(s), "-o", str(o)])
        from concurrent.futures import ThreadPoolExecutor
        with ThreadPoolExecutor(max_workers=min(len(jobs), os.cpu_count() or 4)) as pool:
            list(pool.map(_build._run, jobs))
        _build._run([nvcc, "-shared", "-o", str(out), *map(str, objs), "-gencode", "arch=compute_100a,code=sm_100a"])

    def jit_load(self, verbose: bool = True):
        if self._jit_handle is None:
            path = self.jit_build(verbose=verbose)
            lib = ctypes.CDLL(str(path))
            missing = [s for s in self.SYMBOLS if not hasattr(lib, s)]
            if missing:
                raise OpBuildError(f"op {self.name}: {path.name} lacks {missing}; SOURCES / DEPENDS of the builder are incomplete")
            self._jit_handle = lib
        return self._jit_handle

    # -- load -------------------------------------------------------------------------------------------------------------
    def load(self, verbose: bool = True):
        """Handle of a library that provides this op: the in-tree one (with typed signatures) when it is built and up to
        date, else this op's own JIT-built object."""
        from .. import native
        force_jit = os.environ.get("HDS_FORCE_JIT", "0") == "1"
        if not force_jit and self.installed() and not self.prebuilt_is_stale():
            return native.cuda_lib() if self.LIB == "cuda" else native.host_lib()
        if not force_jit and self.installed() and self.compatibility_problems():
            return native.cuda_lib() if self.LIB == "cuda" else native.host_lib()   # stale but cannot rebuild here: use it
        return self.jit_load(verbose=verbose)

    builder = load


@lru_cache(maxsize=8)
def exported_symbols(lib_path: str, mtime_ns: int) -> frozenset:
    """Dynamic symbols a shared object defines (``nm -D --defined-only``); ``mtime_ns`` only keys the cache."""
    nm = shutil.which("nm")
    if nm is None:
        lib = ctypes.CDLL(lib_path)

        class _Probe(frozenset):
            def __contains__(self, s):
                return hasattr(lib, s)
        return _Probe()
    out = subprocess.run([nm, "-D", "--defined-only", lib_path], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True).stdout
    return frozenset(line.split()[-1] for line in out.splitlines() if line.strip())


def object_dependencies(builder: OpBuilder) -> Dict[str, List[str]]:
    """For a CUDA op whose objects sit in the incremental build cache: which ``hds::`` symbols each translation unit
    leaves undefined (what DEPENDS has to cover). Diagnostic used by the tests to keep the tables honest."""
    nm = shutil.which("nm")
    out: Dict[str, List[str]] = {}
    if nm is None:
        return out
    for s in builder.all_sources():
        o = _build.OBJ_DIR / (s.stem + ".o")
        if not o.exists():
            continue
        txt = subprocess.run([nm, "-C", "--undefined-only", str(o)], stdout=subprocess.PIPE, text=True).stdout
        out[s.name] = sorted({line.split(None, 1)[1].split("(")[0] for line in txt.splitlines() if " hds::" in line or " hds_" in line})
    return out


def unresolved_symbols(builder: OpBuilder) -> List[str]:
    """``hds`` symbols the op's translation units reference but none of SOURCES + DEPENDS defines (needs the incremental
    build's objects; empty list = the op links on its own)."""
    nm = shutil.which("nm")
    if nm is None or builder.LIB != "cuda":
        return []
    defined, wanted = set(), set()
    for s in builder.all_sources():
        o = _build.OBJ_DIR / (s.stem + ".o")
        if not o.exists():
            return []
        for line in subprocess.run([nm, str(o)], stdout=subprocess.PIPE, text=True).stdout.splitlines():
            parts = line.split()
            if len(parts) == 2 and parts[0] == "U":
                if "hds" in parts[1]:
                    wanted.add(parts[1])
            elif len(parts) == 3 and parts[1] in "TtWwBbDdVv":
                defined.add(parts[2])
    return sorted(wanted - defined)
