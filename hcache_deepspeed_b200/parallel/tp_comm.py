"""Tensor-parallel collectives over NVLink 5 / NVSwitch, done by this library's own kernels.

The reference finishes every attention-out / MLP-down GEMM with a separate NCCL all-reduce kernel and assembles the
vocabulary-sharded logits with an NCCL all-gather plus a permute copy
(``deepspeed/inference/v2/model_implementations/llama_v2/model.py:159-160, 168-169, 195-197``). Here:

* every rank's row-parallel GEMM leaves its partial sums in its OWN buffer in *symmetric memory* (same offset on every
  rank, peers mapped into each other's address space, plus an NVLS multicast alias when the switch offers one) and its
  last CTA releases them with flag stores over NVLink (``tp_signal_tail``, ``ops/csrc/hds_ptx.cuh``);
* the next norm kernel (``hds_norm_tp``) acquires the flags and does the collective itself: for decode-size batches a
  one-shot all-reduce (``multimem.ld_reduce`` — the switch adds the ranks' copies — or 16-byte peer loads), for larger
  batches reduce-scatter + all-gather (each rank reduces the rows it owns and ``multimem.st``-broadcasts the normalised
  rows into every rank's activation buffer); the consuming GEMM's TMA producer waits on the flags before its first
  activation load (all-gather -> GEMM);
* the LM-head GEMM signals the same way and ``hds_tp_gather_cols`` pulls the peers' logit slices straight into the
  final ``[S, vocab]`` layout.

Symmetric allocation and handle exchange are plumbing and use ``torch.distributed._symmetric_memory``; flags, data
movement and reductions are this library's kernels. :class:`LoopbackWorld` runs the same protocol for ``n`` emulated
ranks on ONE GPU (one host thread per rank, a single stream, a host-side rendezvous before every waiting launch so
that no kernel ever spins on a flag whose producer has not been enqueued yet) — the single-GPU test vehicle.
"""
from __future__ import annotations

import os
import threading
from typing import List, Optional, Sequence, Tuple

import torch

from ..models.transformer import TPComm
from ..utils.logging import logger

PAD_WORDS = 64  # 256 bytes: arrive[8] | ... | ticket, sent (ops/csrc/hds_ptx.cuh)


class FlagTPComm(TPComm):
    """Shared logic of the real and the loopback communicator: buffer registry + protocol calls."""
    fused = True
    mode = "flags"

    def __init__(self, group, size: int, rank: int, device: torch.device) -> None:
        super().__init__(group, size, rank)
        self.device = device
        self._bufs: List[torch.Tensor] = []
        self._peer_bases: List[List[int]] = []     # per buffer: every rank's base address (this rank's mapping)
        self._mc_bases: List[int] = []             # per buffer: multicast alias base (0 = none)
        self.ctx = -1

    # ---- to be provided -----------------------------------------------------------------------------------
    def _alloc(self, shape, dtype) -> Tuple[torch.Tensor, List[int], int]:
        raise NotImplementedError

    def pre_wait(self) -> None:
        """Called on the host right before a launch that waits on peers' flags (loopback: rendezvous)."""

    # ---- buffers ------------------------------------------------------------------------------------------
    def alloc_symmetric(self, shape, dtype: torch.dtype) -> torch.Tensor:
        """A zeroed tensor in symmetric memory (peers can address it: :meth:`peer_ptrs_raw`, :meth:`multicast_ptr`)."""
        t, bases, mc = self._alloc(tuple(shape), dtype)
        self._bufs.append(t)
        self._peer_bases.append([int(b) for b in bases])
        self._mc_bases.append(int(mc or 0))
        return t

    def alloc_delta(self, n_tokens: int, hidden: int, count: int = 2) -> List[torch.Tensor]:
        return [self.alloc_symmetric((n_tokens, hidden), torch.float32) for _ in range(count)]

    def _locate(self, t: torch.Tensor) -> Tuple[int, int]:
        base = t.data_ptr()
        for i, buf in enumerate(self._bufs):
            b0 = buf.data_ptr()
            if b0 <= base < b0 + buf.numel() * buf.element_size():
                return i, base - b0
        raise ValueError("tensor is not a symmetric allocation of this communicator")

    def is_symmetric(self, t: torch.Tensor) -> bool:
        try:
            self._locate(t)
            return True
        except ValueError:
            return False

    def peer_ptrs_raw(self, t: torch.Tensor, byte_offset: int = 0) -> List[int]:
        """Address of ``t`` (+offset) in every rank's copy of its symmetric allocation, this rank's mapping."""
        i, off = self._locate(t)
        return [b + off + byte_offset for b in self._peer_bases[i]]

    #: which halves of the NVLS path are in use (``HDS_TP_NVLS_LD`` / ``HDS_TP_NVLS_ST`` = 0 switch one off for experiments)
    nvls_ld = True
    nvls_st = True

    def multicast_ptr(self, t: torch.Tensor, kind: str = "ld") -> Optional[int]:
        """NVLS alias of ``t`` (``multimem.*`` address) or None. ``kind``: "ld" (``multimem.ld_reduce`` source) or "st"
        (``multimem.st`` destination)."""
        i, off = self._locate(t)
        on = self.use_nvls and (self.nvls_ld if kind == "ld" else self.nvls_st)
        return self._mc_bases[i] + off if self._mc_bases[i] and on else None

    # ---- protocol -----------------------------------------------------------------------------------------
    def _make_ctx(self) -> None:
        from ..ops import kernels
        self.pad = self.alloc_symmetric((PAD_WORDS, ), torch.int32)
        self.ctx = kernels.tp_ctx_create(self.rank, self.size, self.peer_ptrs_raw(self.pad))

    def close(self) -> None:
        """Release the library-side context slot (the buffers go with the tensors)."""
        if self.ctx >= 0:
            from ..ops import kernels
            kernels.tp_ctx_destroy(self.ctx)
            self.ctx = -1

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    #: experiments only (tools/tp_decode_probe.py): drop every wait and signal — results are WRONG, timing shows what the
    #: flag exchanges cost
    debug_nosync = False

    def arm(self, wait: bool = False, signal: bool = False) -> None:
        """Attach a flag wait / signal to the next GEMM-family launch of this thread."""
        from ..ops import kernels
        if self.debug_nosync:
            return
        if kernels.tp_disarm():
            raise RuntimeError("a tensor-parallel wait/signal was armed but the launch that should carry it never happened")
        if wait:
            self.pre_wait()
        kernels.tp_arm(self.ctx, wait, signal)

    def signal(self) -> None:
        from ..ops import kernels
        if self.debug_nosync:
            return
        kernels.tp_signal(self.ctx)

    def wait(self) -> None:
        from ..ops import kernels
        if self.debug_nosync:
            return
        self.pre_wait()
        kernels.tp_wait(self.ctx)

    def barrier(self) -> None:
        """Device-side barrier of the ranks' current streams."""
        self.signal()
        self.wait()

    def all_reduce_delta(self, delta32: torch.Tensor) -> None:
        """Nothing to launch: the producing GEMM signalled, the consuming norm kernel reduces (``hds_norm_tp``)."""

    def gather_logits(self, local: torch.Tensor, out_full: torch.Tensor, vocab_ranges) -> torch.Tensor:
        """``out_full[:, a_p:b_p] =`` rank p's ``local`` for every rank, pulled over NVLink by one kernel that first
        acquires the flags the LM-head GEMMs signalled (``local`` must be a symmetric allocation)."""
        from ..ops import kernels
        self.pre_wait()
        kernels.tp_gather_cols(self.ctx, out_full, self.peer_ptrs_raw(local), local.stride(0), [a for a, _ in vocab_ranges],
                               [b - a for a, b in vocab_ranges], local.shape[0], wait=not self.debug_nosync)
        return out_full


class FusedTPComm(FlagTPComm):
    """One process per GPU; buffers from ``torch.distributed._symmetric_memory``. ``HDS_TP_NVLS=0`` keeps the NVLS
    multicast aliases unused (peer loads / stores instead of ``multimem.*``)."""

    def __init__(self, group, size: int, rank: int, device: torch.device, mode: str = "flags") -> None:
        super().__init__(group, size, rank, device)
        import torch.distributed._symmetric_memory as symm_mem
        self._symm = symm_mem
        self._group_name = group.group_name
        self._handles = []
        try:
            symm_mem.enable_symm_mem_for_group(self._group_name)
        except Exception:
            pass
        self.use_nvls = os.environ.get("HDS_TP_NVLS", "1") != "0"
        self.nvls_ld = os.environ.get("HDS_TP_NVLS_LD", "1") != "0"
        self.nvls_st = os.environ.get("HDS_TP_NVLS_ST", "1") != "0"
        self._make_ctx()
        self.use_nvls = self.use_nvls and bool(self._mc_bases[0])
        logger.info(f"TP rank {self.rank}/{self.size}: flag protocol ready ({'NVLS multimem' if self.use_nvls else 'peer loads/stores'})")

    def _alloc(self, shape, dtype):
        t = self._symm.empty(shape, dtype=dtype, device=self.device)
        hdl = self._symm.rendezvous(t, group=self._group_name)
        t.zero_()
        torch.cuda.synchronize(self.device)
        hdl.barrier(channel=0)   # init-time only: every rank's zero fill is done before anyone touches a peer's copy
        self._handles.append(hdl)
        mc = int(getattr(hdl, "multicast_ptr", 0) or 0)
        return t, [int(p) for p in hdl.buffer_ptrs], mc

    def peer_view(self, t: torch.Tensor, peer: int) -> torch.Tensor:
        """Tensor aliasing ``t``'s storage in rank ``peer``'s copy of the symmetric allocation (same shape)."""
        i, off = self._locate(t)
        return self._handles[i].get_buffer(peer, tuple(t.shape), t.dtype, off // t.element_size())


class LoopbackWorld:
    """``n`` emulated tensor-parallel ranks on one GPU. Each rank is a host thread with its own communicator; all launch
    on the same stream; ``pre_wait`` is a host rendezvous, so a waiting kernel is always enqueued AFTER the kernels of
    every rank that produce what it waits for, and its spin ends at the first probe. (Ranks as concurrently running
    kernels on one GPU are NOT emulated: nothing guarantees co-scheduling of kernels that wait on one another.)"""

    def __init__(self, n: int, device: torch.device) -> None:
        self.n, self.device = n, device
        self._lock = threading.Lock()
        self._rendezvous = threading.Barrier(n)
        self._allocs: List[List[torch.Tensor]] = []   # [allocation index][rank]
        self.comms = [LoopbackTPComm(self, r) for r in range(n)]
        for c in self.comms:
            c._make_ctx()

    def buffers(self, index: int, shape, dtype) -> List[torch.Tensor]:
        with self._lock:
            while len(self._allocs) <= index:
                self._allocs.append([])
            if not self._allocs[index]:
                self._allocs[index] = [torch.zeros(shape, dtype=dtype, device=self.device) for _ in range(self.n)]
            bufs = self._allocs[index]
            if tuple(bufs[0].shape) != tuple(shape) or bufs[0].dtype != dtype:
                raise RuntimeError("emulated ranks allocated symmetric buffers in different orders")
            return bufs

    def run(self, fn, *args):
        """``fn(rank, comm, *args)`` on one thread per rank; returns the per-rank results (re-raises the first error)."""
        out, err = [None] * self.n, [None] * self.n

        def body(r):
            try:
                torch.cuda.set_device(self.device)
                out[r] = fn(r, self.comms[r], *args)
            except BaseException as e:  # noqa: BLE001
                err[r] = e
                self._rendezvous.abort()

        threads = [threading.Thread(target=body, args=(r, )) for r in range(self.n)]
        for t in threads:
            t.start()
        for t in threads:
            t.join()
        self._rendezvous.reset()
        for e in err:
            if e is not None and not isinstance(e, threading.BrokenBarrierError):
                raise e
        for e in err:
            if e is not None:
                raise e
        return out


class LoopbackTPComm(FlagTPComm):
    use_nvls = False

    def __init__(self, world: LoopbackWorld, rank: int) -> None:
        super().__init__(None, world.n, rank, world.device)
        self.world = world
        self._n_alloc = 0

    def _alloc(self, shape, dtype):
        bufs = self.world.buffers(self._n_alloc, shape, dtype)
        self._n_alloc += 1
        return bufs[self.rank], [b.data_ptr() for b in bufs], 0

    def pre_wait(self) -> None:
        self.world._rendezvous.wait()

    def all_gather_logits(self, local, out_full, vocab_ranges):
        raise RuntimeError("the loopback communicator has no NCCL fallback")


def make_tp_comm(group, size: int, rank: int, device: torch.device, mode: str = "flags", strict: Optional[bool] = None) -> TPComm:
    """The flag-protocol communicator. When symmetric memory cannot be set up (driver / topology) this raises if
    ``strict`` (default: ``HDS_TP_STRICT=1``), otherwise it logs the reason and returns the NCCL baseline."""
    if strict is None:
        strict = os.environ.get("HDS_TP_STRICT", "0") == "1"
    try:
        return FusedTPComm(group, size, rank, device)
    except Exception as e:
        if strict:
            raise
        logger.warning(f"fused tensor-parallel collectives unavailable ({type(e).__name__}: {e}); using NCCL all-reduce")
        return TPComm(group=group, size=size, rank=rank)
