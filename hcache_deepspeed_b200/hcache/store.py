"""Tiered store for saved session state: pinned host DRAM first, NVMe behind it.

The reference keeps nothing (latents are returned to the caller as pageable CPU tensors,
``llama_v2/model.py:218``); the closest thing it has to a tier manager is ZeRO-Infinity's parameter swapper
(``runtime/swap_tensor/partitioned_param_swapper.py:37-422``: pinned buffer pool + aio handle), whose behaviour
this follows: bounded pinned pool, asynchronous spill / fetch through the aio handle, aligned file segments.

DRAM tier: states live in pinned tensors (directly usable as cudaMemcpyAsync sources). When the pool exceeds its
budget the least-recently-used states are written to ``nvme_dir`` (one file per session, ``hcache/format.py``)
and dropped from DRAM. ``load`` brings a state back (aio reads straight into pinned buffers); ``prefetch`` starts
that read early so it overlaps the previous request's decode.
"""
from __future__ import annotations

import os
import threading
from collections import OrderedDict
from dataclasses import dataclass, field
from typing import Dict, List, Optional

import torch

from ..io.aio import AsyncIOHandle
from ..utils.memory import pinned_empty
from .format import HEADER_RESERVED, StateHeader, plan_layout


@dataclass
class SessionState:
    """Everything needed to bring a session's context back without recomputing it (beyond the planned prefix)."""
    tokens: torch.Tensor                       # [n_tokens] int32/int64, host
    kinds: List[str]                           # per layer: "hidden" | "kv" | "none"
    hidden: Optional[torch.Tensor] = None      # [n_hidden_layers, n_tokens, hidden] pinned
    kv: Optional[torch.Tensor] = None          # [n_blocks, n_kv_layers, 2, kv_heads, block_size, head_dim] pinned
    meta: Dict = field(default_factory=dict)
    kv_layer_major: bool = False               # kv is [n_kv_layers, n_blocks, ...]: one contiguous transfer per layer

    @property
    def n_tokens(self) -> int:
        return int(self.tokens.numel())

    def layers_of(self, kind: str) -> List[int]:
        return [i for i, k in enumerate(self.kinds) if k == kind]

    def nbytes(self) -> int:
        n = self.tokens.numel() * self.tokens.element_size()
        for t in (self.hidden, self.kv):
            if t is not None:
                n += t.numel() * t.element_size()
        return n


class HiddenStateStore:

    def __init__(self, dram_bytes: int = 8 << 30, nvme_dir: Optional[str] = None, aio: Optional[AsyncIOHandle] = None,
                 use_direct: bool = False) -> None:
        self.dram_budget = dram_bytes
        self.nvme_dir = nvme_dir
        if nvme_dir:
            os.makedirs(nvme_dir, exist_ok=True)
        self._aio = aio or (AsyncIOHandle(block_size=4 << 20, queue_depth=32, intra_op_parallelism=8, use_direct=use_direct)
                            if nvme_dir else None)
        self._dram: "OrderedDict[int, SessionState]" = OrderedDict()  # LRU order: oldest first
        self._dram_bytes = 0
        self._on_nvme: Dict[int, StateHeader] = {}
        self._prefetching: Dict[int, tuple] = {}
        self._lock = threading.Lock()
        self.stats = {"saves": 0, "dram_hits": 0, "nvme_loads": 0, "spills": 0, "bytes_spilled": 0, "bytes_loaded": 0}

    # ---- paths -------------------------------------------------------------------------------------------
    def _path(self, uid: int) -> str:
        return os.path.join(self.nvme_dir, f"session_{uid}.hcache")

    def __contains__(self, uid: int) -> bool:
        return uid in self._dram or uid in self._on_nvme

    def tier_of(self, uid: int) -> Optional[str]:
        if uid in self._dram:
            return "dram"
        if uid in self._on_nvme:
            return "nvme"
        return None

    # ---- save ----------------------------------------------------------------------------------------------
    def save(self, uid: int, state: SessionState) -> None:
        """Keep ``state`` in the DRAM tier (its tensors should be pinned); spill older sessions if over budget."""
        with self._lock:
            old = self._dram.pop(uid, None)
            if old is not None:
                self._dram_bytes -= old.nbytes()
            self._on_nvme.pop(uid, None)
            self._dram[uid] = state
            self._dram_bytes += state.nbytes()
            self.stats["saves"] += 1
            self._enforce_budget(keep=uid)

    def _enforce_budget(self, keep: Optional[int] = None) -> None:
        while self._dram_bytes > self.dram_budget and len(self._dram) > 1:
            victim = next(u for u in self._dram if u != keep)
            st = self._dram.pop(victim)
            self._dram_bytes -= st.nbytes()
            if self.nvme_dir:
                self._write(victim, st)
            # without an NVMe tier the state is simply evicted (the session will be recomputed)

    def _write(self, uid: int, st: SessionState) -> None:
        dtype = (st.hidden if st.hidden is not None else st.kv).dtype
        hid = st.hidden.shape[2] if st.hidden is not None else 0
        kv_shape = [st.kv.shape[1 if st.kv_layer_major else 0]] + list(st.kv.shape[2:]) if st.kv is not None else None
        tok = st.tokens.to(torch.int32).contiguous()
        hdr = plan_layout(dtype, st.n_tokens, hid, st.kinds, kv_shape,
                          kv_heads=st.kv.shape[3] if st.kv is not None else 0,
                          head_dim=st.kv.shape[5] if st.kv is not None else 0,
                          block_size=st.kv.shape[4] if st.kv is not None else 0, tokens_nbytes=tok.numel() * 4)
        hdr.extra = dict(st.meta)
        path = self._path(uid)
        head = torch.zeros(HEADER_RESERVED, dtype=torch.uint8)
        raw = hdr.to_bytes()
        head[:len(raw)] = torch.frombuffer(bytearray(raw), dtype=torch.uint8)
        reqs = [self._aio.async_pwrite(head, path, 0)]
        hi = ki = 0
        for seg in hdr.layers:
            if seg.kind == "hidden":
                reqs.append(self._aio.async_pwrite(st.hidden[hi].contiguous(), path, seg.offset))
                hi += 1
            elif seg.kind == "kv":
                reqs.append(self._aio.async_pwrite(st.kv[ki] if st.kv_layer_major else st.kv[:, ki].contiguous(), path,
                                                   seg.offset))
                ki += 1
        reqs.append(self._aio.async_pwrite(tok, path, hdr.tokens_offset))
        for r in reqs:
            self._aio.wait(r)
        self._on_nvme[uid] = hdr
        self.stats["spills"] += 1
        self.stats["bytes_spilled"] += hdr.payload_bytes

    # ---- load ----------------------------------------------------------------------------------------------
    def prefetch(self, uid: int) -> None:
        """Start reading an NVMe-resident state into pinned buffers (returns immediately)."""
        with self._lock:
            if uid in self._dram or uid in self._prefetching or uid not in self._on_nvme:
                return
            self._prefetching[uid] = self._start_read(uid)

    def _start_read(self, uid: int):
        hdr = self._on_nvme[uid]
        path = self._path(uid)
        dt = hdr.torch_dtype
        n_h = sum(1 for s in hdr.layers if s.kind == "hidden")
        n_k = sum(1 for s in hdr.layers if s.kind == "kv")
        hidden = pinned_empty((n_h, hdr.n_tokens, hdr.hidden), dt) if n_h else None
        kv_seg = next((s for s in hdr.layers if s.kind == "kv"), None)
        kv_layers = [pinned_empty(tuple(kv_seg.shape), dt) for _ in range(n_k)] if n_k else []
        tokens = torch.empty(hdr.tokens_nbytes // 4, dtype=torch.int32)
        reqs = []
        hi = ki = 0
        for seg in hdr.layers:
            if seg.kind == "hidden":
                reqs.append(self._aio.async_pread(hidden[hi], path, seg.offset))
                hi += 1
            elif seg.kind == "kv":
                reqs.append(self._aio.async_pread(kv_layers[ki], path, seg.offset))
                ki += 1
        reqs.append(self._aio.async_pread(tokens, path, hdr.tokens_offset))
        return hdr, hidden, kv_layers, tokens, reqs

    def load(self, uid: int) -> SessionState:
        """Return the state of ``uid`` (from DRAM, or read back from NVMe) and mark it most-recently used."""
        with self._lock:
            st = self._dram.get(uid)
            if st is not None:
                self._dram.move_to_end(uid)
                self.stats["dram_hits"] += 1
                return st
            if uid not in self._on_nvme:
                raise KeyError(f"no saved state for session {uid}")
            pending = self._prefetching.pop(uid, None) or self._start_read(uid)
        hdr, hidden, kv_layers, tokens, reqs = pending
        for r in reqs:
            self._aio.wait(r)
        # segments are stored one layer at a time: come back layer-major, the layout the pipelined reload streams
        kv = torch.stack(kv_layers, dim=0) if kv_layers else None
        if kv is not None and torch.cuda.is_available():
            kv = kv.pin_memory()
        st = SessionState(tokens=tokens, kinds=[s.kind for s in hdr.layers], hidden=hidden, kv=kv, meta=dict(hdr.extra),
                          kv_layer_major=True)
        with self._lock:
            self.stats["nvme_loads"] += 1
            self.stats["bytes_loaded"] += hdr.payload_bytes
            self._dram[uid] = st
            self._dram_bytes += st.nbytes()
            self._enforce_budget(keep=uid)
        return st

    def drop(self, uid: int) -> None:
        with self._lock:
            st = self._dram.pop(uid, None)
            if st is not None:
                self._dram_bytes -= st.nbytes()
            if self._on_nvme.pop(uid, None) is not None:
                try:
                    os.remove(self._path(uid))
                except OSError:
                    pass

    @property
    def dram_bytes_used(self) -> int:
        return self._dram_bytes
