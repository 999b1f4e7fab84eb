"""Ragged batch container: everything one forward needs to know about its tokens and sequences.

All metadata lives in ONE pinned int32 staging buffer and its device mirror; ``finalize`` fills it
with a single native call (``hds_build_batch``) and ships it with a single asynchronous H2D copy.
The reference builds Python lists per token and issues five copies of separate buffers, and hands
kernels raw device pointers to per-sequence block lists
(``deepspeed/inference/v2/ragged/ragged_wrapper.py:31-292``); here kernels get a dense
``block_tables[n_seqs, stride]`` plus per-token ``tok_pos`` / ``tok_slot`` and ready-made attention
work items.
"""
from __future__ import annotations

import ctypes
from typing import List, Optional, Tuple

import torch

from ....ops import native
from ....ops.host_bindings import SeqIn
from ....utils.memory import pinned_zeros
from . import fast_host_buffer
from .manager_configs import DSStateManagerConfig
from .sequence_descriptor import DSSequenceDescriptor

#: tokens per tensor-core attention item (rows of one tcgen05 Q tile)
LARGE_ITEM_TOKENS = 128
#: query rows (tokens x q-heads-per-kv-head) a split-KV attention item may carry
SMALL_ITEM_ROWS = 8


def to_padded(original_size: int) -> int:
    """Pad token counts to a multiple of 64 (128 above 512) — CUDA-graph friendly bucket sizes."""
    def _pad(size, m):
        return ((size + m - 1) // m) * m
    return _pad(original_size, 64) if original_size <= 512 else _pad(original_size, 128)


class RaggedBatchWrapper:

    def __init__(self, config: DSStateManagerConfig, block_size: int = 64, max_blocks_per_seq: Optional[int] = None,
                 group: int = 1, device: Optional[torch.device] = None) -> None:
        self._config = config
        self._T = config.max_ragged_batch_size
        self._S = config.max_ragged_sequence_count
        self._device = device if device is not None else (torch.device("cuda", torch.cuda.current_device())
                                                          if torch.cuda.is_available() else torch.device("cpu"))
        self._lib = native.host_lib()
        self._seqs = (SeqIn * self._S)()
        self._counts = (ctypes.c_int32 * 8)()
        self.configure_kv(block_size, max_blocks_per_seq or -(-config.max_context // block_size), group)
        self.clear()

    def configure_kv(self, block_size: int, max_blocks_per_seq: int, group: int) -> None:
        """(Re)size the staging buffers for a KV geometry: block size, table stride, q heads per kv head."""
        T, S = self._T, self._S
        self._block_size = block_size
        self._stride = max_blocks_per_seq
        self._group = max(1, group)
        self._I = S + -(-T // LARGE_ITEM_TOKENS) + 1
        off = 0
        self._off = {}
        for name, n in (("ids", T), ("pos", T), ("slot", T), ("seq", T), ("last", S), ("small", 4 * self._I),
                        ("large", 4 * self._I), ("tables", S * self._stride)):
            self._off[name] = (off, n)
            off += n
        mode = fast_host_buffer.staging_mode() if self._device.type == "cuda" else None
        if mode is not None:    # opt-in: natively allocated mapped (+ write-combined) staging memory
            self._fast = fast_host_buffer.FastHostBuffer(off, torch.int32, mapped=True, write_combined=(mode == "wc"))
            self._host = self._fast.tensor
        else:
            self._host = pinned_zeros((off, ), torch.int32)
        self._host_np = self._host.numpy()   # same memory: scalar / slice writes without tensor-op dispatch overhead
        self._dev = self._host if self._device.type == "cpu" else torch.zeros(off, dtype=torch.int32,
                                                                               device=self._device)

    # ---- building ---------------------------------------------------------------------------------
    def clear(self) -> None:
        self._n_tokens = 0
        self._n_seqs = 0
        self._finalized = False
        self._inflight: List[Tuple[int, int, int]] = []
        self._descs: List[DSSequenceDescriptor] = []
        self._n_small = self._n_large = self._rows_small = self._max_kv_small = self._max_kv_large = 0

    def insert_sequence(self, seq_descriptor: DSSequenceDescriptor, tokens: torch.Tensor, do_checks: bool = True) -> None:
        n = int(tokens.numel())
        if do_checks:
            if self._n_tokens + n > self._T:
                raise RuntimeError(f"Ragged batch is full: {self._n_tokens} + {n} tokens exceeds {self._T}")
            if self._n_seqs >= self._S:
                raise RuntimeError(f"Ragged batch is full: sequence limit {self._S} reached")
        start = self._off["ids"][0] + self._n_tokens
        if tokens.device.type == "cpu":
            # host tokens (the normal case): numpy view of the pinned staging buffer, ~1 us instead of ~10 us per sequence
            if n == 1:
                self._host_np[start] = tokens.item()
            else:
                self._host_np[start:start + n] = tokens.reshape(-1).numpy()
        else:
            self._host[start:start + n] = tokens.reshape(-1).to(torch.int32)
        s = self._seqs[self._n_seqs]
        s.n_new = n
        s.seen = seq_descriptor.seen_tokens
        s.n_blocks = seq_descriptor.cur_allocated_blocks
        s.block_ids = seq_descriptor.kv_blocks_ptr()
        self._inflight.append((self._n_tokens, n, seq_descriptor.seen_tokens))
        self._descs.append(seq_descriptor)
        self._n_tokens += n
        self._n_seqs += 1

    def _hptr(self, name: str) -> int:
        return self._host.data_ptr() + 4 * self._off[name][0]

    def finalize(self, padding: bool = False) -> None:
        """Derive per-token and per-item metadata natively, then one async H2D of the staging prefix."""
        rc = self._lib.hds_build_batch(self._seqs, self._n_seqs, self._block_size, self._group, SMALL_ITEM_ROWS,
                                       LARGE_ITEM_TOKENS, self._T, self._I, self._stride, self._hptr("pos"),
                                       self._hptr("slot"), self._hptr("seq"), self._hptr("last"), self._hptr("tables"),
                                       self._hptr("small"), self._hptr("large"), self._counts)
        if rc == -2:
            raise RuntimeError("a sequence in the batch owns fewer KV blocks than its tokens need")
        if rc != 0:
            raise RuntimeError("ragged batch exceeds the configured token / item / context capacity")
        assert self._counts[0] == self._n_tokens
        self._n_small, self._n_large = self._counts[1], self._counts[2]
        self._rows_small, self._max_kv_small, self._max_kv_large = self._counts[3], self._counts[4], self._counts[5]
        self._padded_tokens = to_padded(self._n_tokens) if padding else self._n_tokens
        if padding and self._padded_tokens > self._n_tokens:
            pad = slice(self._n_tokens, min(self._padded_tokens, self._T))
            for name, fill in (("ids", -1), ("pos", 0), ("slot", -1), ("seq", 0)):
                o = self._off[name][0]
                self._host[o + pad.start:o + pad.stop] = fill
        if self._dev is not self._host:
            used = self._off["tables"][0] + self._n_seqs * self._stride
            self._dev[:used].copy_(self._host[:used], non_blocking=True)
        self._finalized = True

    # ---- views (device unless on_device=False) --------------------------------------------------------
    def _view(self, name: str, n: int, on_device: bool = True) -> torch.Tensor:
        o = self._off[name][0]
        return (self._dev if on_device else self._host)[o:o + n]

    def input_ids(self, on_device: bool = True) -> torch.Tensor:
        return self._view("ids", self._n_tokens, on_device)

    def tok_pos(self, on_device: bool = True) -> torch.Tensor:
        return self._view("pos", self._n_tokens, on_device)

    def tok_slot(self, on_device: bool = True) -> torch.Tensor:
        return self._view("slot", self._n_tokens, on_device)

    def tokens_to_seq(self, on_device: bool = True) -> torch.Tensor:
        return self._view("seq", self._n_tokens, on_device)

    def last_rows(self, on_device: bool = True) -> torch.Tensor:
        return self._view("last", self._n_seqs, on_device)

    def block_tables(self, on_device: bool = True) -> torch.Tensor:
        return self._view("tables", self._n_seqs * self._stride, on_device).view(self._n_seqs, self._stride)

    def items_small(self, on_device: bool = True) -> torch.Tensor:
        return self._view("small", 4 * self._n_small, on_device).view(self._n_small, 4)

    def items_large(self, on_device: bool = True) -> torch.Tensor:
        return self._view("large", 4 * self._n_large, on_device).view(self._n_large, 4)

    # ---- scalars -------------------------------------------------------------------------------------
    @property
    def current_tokens(self) -> int:
        return self._n_tokens

    @property
    def current_sequences(self) -> int:
        return self._n_seqs

    @property
    def tensor_toks(self) -> int:
        return self._n_tokens

    @property
    def n_items_small(self) -> int:
        return self._n_small

    @property
    def n_items_large(self) -> int:
        return self._n_large

    @property
    def rows_per_small_item(self) -> int:
        return self._rows_small

    @property
    def max_kv_len_small(self) -> int:
        return self._max_kv_small

    @property
    def max_kv_len_large(self) -> int:
        return self._max_kv_large

    @property
    def block_size(self) -> int:
        return self._block_size

    @property
    def table_stride(self) -> int:
        return self._stride

    @property
    def inflight_seq_descriptors_shadow(self) -> List[Tuple[int, int, int]]:
        """Host copy of ``(start_token, n_tokens, seen_tokens)`` per sequence of the batch."""
        return self._inflight

    # name kept for code written against the reference (engine_v2.py:173)
    _inflight_seq_descriptors_shadow = inflight_seq_descriptors_shadow

    # ---- accessors of the reference's wrapper, expressed over this wrapper's staging layout (ragged_wrapper.py:230-276) ----
    def inflight_seq_descriptors(self, on_device: bool = True) -> torch.Tensor:
        """int32 ``[n_seqs, 4]``: (first token of the sequence in the batch, tokens in flight, tokens seen before, 0)."""
        rows = [(b, n, seen, 0) for b, n, seen in self._inflight] or [(0, 0, 0, 0)][:0]
        t = torch.tensor(rows, dtype=torch.int32).reshape(-1, 4)
        return t.to(self._device) if on_device and self._device.type != "cpu" else t

    def batch_metadata_buffer(self, on_device: bool = True) -> torch.Tensor:
        """int32 ``[2]``: (tokens, sequences) of the in-flight batch."""
        t = torch.tensor([self._n_tokens, self._n_seqs], dtype=torch.int32)
        return t.to(self._device) if on_device and self._device.type != "cpu" else t

    def kv_ptrs(self, on_device: bool = True) -> torch.Tensor:
        """Where each sequence's KV block ids live: here the rows of the block table (``block_tables()``), not raw pointers."""
        return self._view("tables", self._n_seqs * self._stride, on_device).view(self._n_seqs, self._stride)

    def masks(self, on_device: bool = True):
        """Placeholder of the reference for non-causal masks (BERT-like models): not supported there either."""
        return None

    @property
    def sequences(self) -> List[DSSequenceDescriptor]:
        return self._descs
