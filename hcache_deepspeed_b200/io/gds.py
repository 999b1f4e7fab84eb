"""GPUDirect Storage path (cuFile) for the NVMe tier: file -> device memory without a host bounce buffer.

Counterpart of the reference's ``gds_handle`` (``csrc/gds/py_lib/deepspeed_gds_op.cpp:22-144``,
``deepspeed_py_gds_handle.cpp:59-62``): ``cuFileDriverOpen``, ``cuFileHandleRegister``, ``cuFileBufRegister``,
``cuFileRead`` / ``cuFileWrite`` at explicit file and device offsets. libcufile is bound with ctypes at run time;
when the library, the nvidia-fs driver or a GDS-capable file system is missing, :func:`gds_available` is False and
the store falls back to the async-I/O + pinned-staging path.
"""
from __future__ import annotations

import ctypes
import os
from typing import Optional

import torch

_lib: Optional[ctypes.CDLL] = None
_driver_open = False


class _CUfileDescr(ctypes.Structure):
    _fields_ = [("type", ctypes.c_int), ("fd", ctypes.c_int), ("_pad", ctypes.c_char * 8), ("fs_ops", ctypes.c_void_p)]


class _CUfileError(ctypes.Structure):
    _fields_ = [("err", ctypes.c_int), ("cu_err", ctypes.c_int)]


def _load() -> Optional[ctypes.CDLL]:
    global _lib
    if _lib is not None:
        return _lib
    for name in ("libcufile.so.0", "libcufile.so", "/usr/local/cuda/lib64/libcufile.so"):
        try:
            lib = ctypes.CDLL(name)
        except OSError:
            continue
        lib.cuFileDriverOpen.restype = _CUfileError
        lib.cuFileDriverClose.restype = _CUfileError
        lib.cuFileHandleRegister.restype = _CUfileError
        lib.cuFileHandleRegister.argtypes = [ctypes.POINTER(ctypes.c_void_p), ctypes.POINTER(_CUfileDescr)]
        lib.cuFileHandleDeregister.argtypes = [ctypes.c_void_p]
        lib.cuFileBufRegister.restype = _CUfileError
        lib.cuFileBufRegister.argtypes = [ctypes.c_void_p, ctypes.c_size_t, ctypes.c_int]
        lib.cuFileBufDeregister.restype = _CUfileError
        lib.cuFileBufDeregister.argtypes = [ctypes.c_void_p]
        for fn in (lib.cuFileRead, lib.cuFileWrite):
            fn.restype = ctypes.c_ssize_t
            fn.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_size_t, ctypes.c_longlong, ctypes.c_longlong]
        _lib = lib
        return lib
    return None


_probe_result: Optional[str] = None   # None = not probed yet; "" = usable; otherwise the reason it is not

_PROBE = r"""
import ctypes, os, sys, tempfile
sys.path.insert(0, {root!r})
import torch
from hcache_deepspeed_b200.io import gds
lib = gds._load()
if lib is None:
    print("REASON libcufile not found"); sys.exit(0)
st = lib.cuFileDriverOpen()
if st.err != 0:
    print("REASON cuFileDriverOpen error", st.err); sys.exit(0)
gds._driver_open = True
gds._probe_result = ""
h = gds.GDSHandle(block_size=1 << 20, threads=2)
x = torch.arange(1 << 18, dtype=torch.int32, device="cuda")
path = os.path.join({tmp!r}, "hds_gds_probe_%d.bin" % os.getpid())
try:
    h.sync_pwrite(x, path)
    y = torch.zeros_like(x)
    h.sync_pread(y, path)
    torch.cuda.synchronize()
    print("OK" if torch.equal(x, y) else "REASON cuFile round trip returned different bytes")
finally:
    if os.path.exists(path):
        os.remove(path)
"""


def probe_gds(timeout_s: float = 45.0, directory: Optional[str] = None) -> str:
    """Try libcufile end to end (driver open, one write, one read) in a CHILD process with a time limit and return ""
    when it works or the reason when it does not. cuFile calls cannot be interrupted from Python: on a machine whose
    nvidia-fs / file-system set-up is broken ``cuFileDriverOpen`` or the first transfer may block forever (seen on the
    B200 boxes of this project's pod), and a serving process must find that out without hanging itself."""
    import subprocess
    import sys
    import tempfile
    root = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    code = _PROBE.format(root=root, tmp=directory or tempfile.gettempdir())
    proc = subprocess.Popen([sys.executable, "-c", code], stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True,
                            start_new_session=True)
    try:
        stdout, stderr = proc.communicate(timeout=timeout_s)
    except subprocess.TimeoutExpired:
        proc.kill()   # no wait(): a child stuck in an uninterruptible driver call would block the reaper as well
        return f"libcufile did not answer within {timeout_s:.0f} s (driver open or first transfer blocked)"
    out = (stdout or "").strip().splitlines()
    if out and out[-1] == "OK":
        return ""
    if out and out[-1].startswith("REASON"):
        return out[-1][7:]
    return f"probe process failed (rc={proc.returncode}): {(stderr or '').strip()[-200:]}"


def gds_available() -> bool:
    """True when libcufile works on this machine: checked ONCE per process by :func:`probe_gds` (a child process with a
    time limit), only then is the driver opened here. ``HDS_GDS=0`` switches the path off without probing."""
    global _driver_open, _probe_result
    if not torch.cuda.is_available() or os.environ.get("HDS_GDS", "1") == "0":
        return False
    if _probe_result is None:
        _probe_result = "libcufile not found" if _load() is None else probe_gds()
    if _probe_result:
        return False
    if not _driver_open:
        st = _load().cuFileDriverOpen()
        _driver_open = st.err == 0
    return _driver_open


def gds_unavailable_reason() -> str:
    """Why :func:`gds_available` is False ("" when it is True or has not been asked yet)."""
    return _probe_result or ""


def gds_mode() -> str:
    """'nvidia-fs' when the kernel module that does the peer-to-peer DMA is loaded, 'compat' when cuFile serves the calls
    through its own pinned bounce buffers (POSIX reads + cudaMemcpy inside the library), 'unavailable' otherwise."""
    if not gds_available():
        return "unavailable"
    return "nvidia-fs" if os.path.exists("/proc/driver/nvidia-fs/stats") else "compat"


class GDSHandle:
    """cuFile reads / writes between a file and a device tensor. ``sync_*`` block the caller; ``async_*`` cut the
    transfer into ``block_size`` pieces executed by ``threads`` worker threads (cuFile is synchronous per calling thread,
    so parallel calls at different offsets are how the reference keeps a drive busy too:
    ``csrc/gds/py_lib/deepspeed_gds_op.cpp:72-140``) and return a request id for :meth:`wait`."""

    def __init__(self, block_size: int = 8 << 20, threads: int = 4) -> None:
        if not gds_available():
            raise RuntimeError("GPUDirect Storage is not available on this system (libcufile / nvidia-fs)")
        self._lib = _load()
        self._registered = {}
        self._block_size, self._threads = max(4096, block_size), max(1, threads)
        self._pool = None
        self._requests = {}
        self._next_id = 1

    # ---- asynchronous API (ref: deepspeed_py_gds_handle.cpp:59-62 async_pread / async_pwrite / wait) -------------------
    def _submit(self, write: bool, t: torch.Tensor, path: str, file_offset: int, nbytes: Optional[int]) -> int:
        from concurrent.futures import ThreadPoolExecutor
        if self._pool is None:
            self._pool = ThreadPoolExecutor(max_workers=self._threads, thread_name_prefix="hds-gds")
        n = nbytes if nbytes is not None else t.numel() * t.element_size()
        fd, handle = self._open(path, write=write)
        fn = self._lib.cuFileWrite if write else self._lib.cuFileRead
        base = ctypes.c_void_p(t.data_ptr())
        dev = t.device

        def piece(off, length):
            torch.cuda.set_device(dev)
            got = fn(handle, base, length, file_offset + off, off)
            if got < 0:
                raise OSError(f"cuFile{'Write' if write else 'Read'} failed with {got}")
            return int(got)

        futs = [self._pool.submit(piece, off, min(self._block_size, n - off)) for off in range(0, max(n, 1), self._block_size)]
        rid = self._next_id
        self._next_id += 1
        self._requests[rid] = (futs, fd, handle, t)
        return rid

    def async_pread(self, device_tensor: torch.Tensor, path: str, file_offset: int = 0, nbytes: Optional[int] = None) -> int:
        return self._submit(False, device_tensor, path, file_offset, nbytes)

    def async_pwrite(self, device_tensor: torch.Tensor, path: str, file_offset: int = 0, nbytes: Optional[int] = None) -> int:
        return self._submit(True, device_tensor, path, file_offset, nbytes)

    def wait(self, request_id: int = 0) -> int:
        """Bytes moved by ``request_id`` (0: by all outstanding requests); raises the first error of a failed piece."""
        ids = [request_id] if request_id else list(self._requests)
        total = 0
        for rid in ids:
            futs, fd, handle, _ = self._requests.pop(rid)
            try:
                total += sum(f.result() for f in futs)
            finally:
                self._lib.cuFileHandleDeregister(handle)
                os.close(fd)
        return total

    def pending(self) -> int:
        return sum(1 for futs, *_ in self._requests.values() if not all(f.done() for f in futs))

    def pin_device_tensor(self, t: torch.Tensor) -> None:
        """Register a device buffer once so repeated transfers skip per-call registration."""
        st = self._lib.cuFileBufRegister(ctypes.c_void_p(t.data_ptr()), t.numel() * t.element_size(), 0)
        if st.err != 0:
            raise RuntimeError(f"cuFileBufRegister failed: {st.err}")
        self._registered[t.data_ptr()] = t

    def unpin_device_tensor(self, t: torch.Tensor) -> None:
        if self._registered.pop(t.data_ptr(), None) is not None:
            self._lib.cuFileBufDeregister(ctypes.c_void_p(t.data_ptr()))

    def new_pinned_device_tensor(self, num_elem: int, example: torch.Tensor) -> torch.Tensor:
        t = torch.empty(num_elem, dtype=example.dtype, device=example.device)
        self.pin_device_tensor(t)
        return t

    def _open(self, path: str, write: bool):
        flags = (os.O_WRONLY | os.O_CREAT if write else os.O_RDONLY) | getattr(os, "O_DIRECT", 0)
        fd = os.open(path, flags, 0o644)
        descr = _CUfileDescr(type=1, fd=fd)  # CU_FILE_HANDLE_TYPE_OPAQUE_FD
        handle = ctypes.c_void_p()
        st = self._lib.cuFileHandleRegister(ctypes.byref(handle), ctypes.byref(descr))
        if st.err != 0:
            os.close(fd)
            raise RuntimeError(f"cuFileHandleRegister({path}) failed: {st.err}")
        return fd, handle

    def sync_pread(self, device_tensor: torch.Tensor, path: str, file_offset: int = 0, nbytes: Optional[int] = None,
                   device_offset: int = 0) -> int:
        n = nbytes if nbytes is not None else device_tensor.numel() * device_tensor.element_size()
        fd, handle = self._open(path, write=False)
        try:
            got = self._lib.cuFileRead(handle, ctypes.c_void_p(device_tensor.data_ptr()), n, file_offset, device_offset)
        finally:
            self._lib.cuFileHandleDeregister(handle)
            os.close(fd)
        if got < 0:
            raise OSError(f"cuFileRead failed with {got}")
        return int(got)

    def sync_pwrite(self, device_tensor: torch.Tensor, path: str, file_offset: int = 0, nbytes: Optional[int] = None,
                    device_offset: int = 0) -> int:
        n = nbytes if nbytes is not None else device_tensor.numel() * device_tensor.element_size()
        fd, handle = self._open(path, write=True)
        try:
            got = self._lib.cuFileWrite(handle, ctypes.c_void_p(device_tensor.data_ptr()), n, file_offset, device_offset)
        finally:
            self._lib.cuFileHandleDeregister(handle)
            os.close(fd)
        if got < 0:
            raise OSError(f"cuFileWrite failed with {got}")
        return int(got)
