"""Sub-allocator over ONE pre-allocated flat tensor, with compaction: tensors handed out are views into the buffer; when
no free run is long enough the live tensors are slid towards the start (their contents move, parameters that were bound
to them are re-pointed) and the request is served from the tail. Used for ZeRO-3 parameter gathering buffers so that
steady-state training never calls the device allocator.

Reference: ``deepspeed/runtime/zero/contiguous_memory_allocator.py:16-287`` (five dicts keyed by address / tensor id).
Here the state is one address-sorted list of live segments; free space is whatever lies between them.
"""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import Dict, List, Optional, Tuple

import torch


@dataclass(eq=False)
class _Segment:
    start: int
    size: int
    tensor: torch.Tensor
    params: List[Tuple[torch.nn.Parameter, int, tuple]] = field(default_factory=list)   # (param, numel, shape)


class ContiguousMemoryAllocator:

    def __init__(self, size: int, dtype: torch.dtype, device) -> None:
        self.buffer = torch.zeros(size, dtype=dtype, device=device)
        self.total_size = size
        self.total_free = size
        self.max_allocated = 0
        self._segments: List[_Segment] = []          # sorted by start
        self._by_id: Dict[int, _Segment] = {}

    # ---- queries ---------------------------------------------------------------------------------------------------
    def _free_runs(self) -> List[Tuple[int, int]]:
        runs, cursor = [], 0
        for seg in self._segments:
            if seg.start > cursor:
                runs.append((cursor, seg.start - cursor))
            cursor = seg.start + seg.size
        if cursor < self.total_size:
            runs.append((cursor, self.total_size - cursor))
        return runs

    @property
    def largest_contiguous(self) -> int:
        return max((n for _, n in self._free_runs()), default=0)

    @property
    def contiguous_sizes(self) -> Dict[int, int]:
        """address -> length of every free run (the reference's bookkeeping dict, derived on demand)."""
        return dict(self._free_runs())

    def print_allocation(self, resolution: int = 200) -> str:
        cells = ["."] * resolution
        for seg in self._segments:
            a = seg.start * resolution // self.total_size
            b = max(a + 1, (seg.start + seg.size) * resolution // self.total_size)
            for i in range(a, min(b, resolution)):
                cells[i] = "x"
        line = "".join(cells)
        print(line)
        return line

    # ---- allocation ------------------------------------------------------------------------------------------------
    def allocate_tensor(self, size: int) -> torch.Tensor:
        """A view of ``size`` elements. Fails if the buffer does not have that much free space in total; compacts first
        when it has, but not in one piece."""
        assert size <= self.total_free, "Not enough memory in buffer. Allocation failed"
        run = self._first_fit(size)
        if run is None:
            self._defragment_memory()
            run = self._first_fit(size)
            assert run is not None
        seg = _Segment(run, size, self.buffer.narrow(0, run, size))
        self._insert(seg)
        self.total_free -= size
        self.max_allocated = max(self.max_allocated, self.total_size - self.total_free)
        return seg.tensor

    def _first_fit(self, size: int) -> Optional[int]:
        for start, n in self._free_runs():
            if n >= size:
                return start
        return None

    def _insert(self, seg: _Segment) -> None:
        i = 0
        while i < len(self._segments) and self._segments[i].start < seg.start:
            i += 1
        self._segments.insert(i, seg)
        self._by_id[id(seg.tensor)] = seg

    def assign_to_param(self, tensor: torch.Tensor, param: torch.nn.Parameter, numel: int, shape) -> None:
        """Point ``param.data`` at the first ``numel`` elements of an allocated tensor; compaction keeps it pointed there."""
        seg = self._by_id.get(id(tensor))
        assert seg is not None, "No such tensor allocated by the allocator."
        assert numel <= seg.size, f"Assert tensor buffer does is not large enough"
        seg.params.append((param, numel, tuple(shape)))
        param.data = seg.tensor.narrow(0, 0, numel).view(shape)

    def release_tensor(self, tensor: torch.Tensor) -> None:
        self.release_tensor_with_id(id(tensor))

    def release_tensor_with_id(self, tensor_id: int) -> None:
        seg = self._by_id.pop(tensor_id, None)
        assert seg is not None, "Invalid tensor id"
        self._segments.remove(seg)
        self.total_free += seg.size

    # ---- compaction ------------------------------------------------------------------------------------------------
    def _defragment_memory(self) -> None:
        """Slide every live segment to the lowest free address, in address order (so a move never overwrites data that
        has not been moved yet; overlapping moves go through a temporary)."""
        cursor = 0
        for seg in self._segments:
            if seg.start != cursor:
                dst = self.buffer.narrow(0, cursor, seg.size)
                src = seg.tensor
                if cursor + seg.size > seg.start:          # ranges overlap
                    src = src.clone()
                dst.copy_(src)
                seg.start = cursor
                seg.tensor.data = dst          # the caller's handle stays valid: same tensor object, new storage window
                for param, numel, shape in seg.params:
                    param.data = dst.narrow(0, 0, numel).view(shape)
            cursor += seg.size

