"""Block-sparsity layouts for sparse attention: dense, fixed (Sparse Transformer), variable, BigBird, BSLongformer,
local sliding window. ``make_layout(seq_len)`` returns an int tensor ``[heads, blocks, blocks]`` of 0/1.
Reference: ``deepspeed/ops/sparse_attention/sparsity_config.py:9-730``."""
from __future__ import annotations

import random

import torch


class SparsityConfig:

    def __init__(self, num_heads: int, block: int = 16, different_layout_per_head: bool = False):
        self.num_heads, self.block, self.different_layout_per_head = num_heads, block, different_layout_per_head
        self.num_layout_heads = num_heads if different_layout_per_head else 1

    def setup_layout(self, seq_len: int) -> torch.Tensor:
        if seq_len % self.block != 0:
            raise ValueError(f"Sequence Length, {seq_len}, needs to be dividable by Block size {self.block}!")
        n = seq_len // self.block
        return torch.zeros((self.num_heads, n, n), dtype=torch.int64)

    def check_and_propagate_first_head_layout(self, layout: torch.Tensor) -> torch.Tensor:
        if not self.different_layout_per_head:
            layout[1:self.num_heads] = layout[0]
        return layout


class DenseSparsityConfig(SparsityConfig):
    def make_layout(self, seq_len: int) -> torch.Tensor:
        layout = self.setup_layout(seq_len)
        layout[:] = 1
        return layout


class FixedSparsityConfig(SparsityConfig):

    def __init__(self, num_heads, block=16, different_layout_per_head=False, num_local_blocks=4, num_global_blocks=1,
                 attention="bidirectional", horizontal_global_attention=False, num_different_global_patterns=1):
        super().__init__(num_heads, block, different_layout_per_head)
        if num_local_blocks % num_global_blocks != 0:
            raise ValueError("Number of blocks in a local window must be dividable by number of global blocks")
        if attention not in ("unidirectional", "bidirectional"):
            raise NotImplementedError('only "uni/bi-directional" attentions are supported for now!')
        if attention != "bidirectional" and horizontal_global_attention:
            raise ValueError('only "bi-directional" attentions can support horizontal global attention!')
        if num_different_global_patterns > 1 and not different_layout_per_head:
            raise ValueError("Number of different layouts cannot be more than one when you have set a single layout for all heads!")
        if num_different_global_patterns > num_local_blocks // num_global_blocks:
            raise ValueError("Number of layout versions cannot be larger than local window size / global blocks")
        self.num_local_blocks, self.num_global_blocks, self.attention = num_local_blocks, num_global_blocks, attention
        self.horizontal_global_attention, self.num_different_global_patterns = horizontal_global_attention, num_different_global_patterns

    def set_local_layout(self, h, layout):
        n = layout.shape[1]
        for i in range(0, n, self.num_local_blocks):
            end = min(i + self.num_local_blocks, n)
            for r in range(i, end):
                for c in range(i, (r + 1 if self.attention == "unidirectional" else end)):
                    layout[h, r, c] = 1
        return layout

    def set_global_layout(self, h, layout):
        n = layout.shape[1]
        first = self.num_local_blocks - (1 + h % self.num_different_global_patterns) * self.num_global_blocks
        end = n - n % self.num_local_blocks
        for i in range(first, end, self.num_local_blocks):
            first_row = 0 if self.attention == "bidirectional" else i
            layout[h, first_row:, i:i + self.num_global_blocks] = 1
            if self.horizontal_global_attention:
                layout[h, i:i + self.num_global_blocks, :] = 1
        if end < n:
            start = min(end + first, n - self.num_global_blocks)
            first_row = 0 if self.attention == "bidirectional" else start
            layout[h, first_row:, start:start + self.num_global_blocks] = 1
            if self.horizontal_global_attention:
                layout[h, start:start + self.num_global_blocks, :] = 1
        return layout

    def make_layout(self, seq_len):
        layout = self.setup_layout(seq_len)
        for h in range(self.num_layout_heads):
            layout = self.set_global_layout(h, self.set_local_layout(h, layout))
        return self.check_and_propagate_first_head_layout(layout)


class VariableSparsityConfig(SparsityConfig):

    def __init__(self, num_heads, block=16, different_layout_per_head=False, num_random_blocks=0, local_window_blocks=(4, ),
                 global_block_indices=(0, ), global_block_end_indices=None, attention="bidirectional",
                 horizontal_global_attention=False):
        super().__init__(num_heads, block, different_layout_per_head)
        self.num_random_blocks, self.local_window_blocks = num_random_blocks, list(local_window_blocks)
        self.global_block_indices = list(global_block_indices)
        if global_block_end_indices is not None:
            if len(global_block_indices) != len(global_block_end_indices):
                raise ValueError("Global block start indices length must be same as global block end indices length")
            for s, e in zip(global_block_indices, global_block_end_indices):
                if s >= e:
                    raise ValueError("Global block start index must be smaller than global block end index")
        self.global_block_end_indices = None if global_block_end_indices is None else list(global_block_end_indices)
        self.attention, self.horizontal_global_attention = attention, horizontal_global_attention

    def set_random_layout(self, h, layout):
        n = layout.shape[1]
        if n < self.num_random_blocks:
            raise ValueError("Number of random blocks must be smaller than overall number of blocks in a row")
        for r in range(n):
            cols = random.sample(range(n), self.num_random_blocks) if self.attention == "bidirectional" else \
                random.sample(range(r + 1), min(self.num_random_blocks, r + 1))
            layout[h, r, cols] = 1
        return layout

    def set_local_layout(self, h, layout):
        n = layout.shape[1]
        start = 0
        sizes = list(self.local_window_blocks)
        while start < n:
            size = sizes.pop(0) if len(sizes) > 1 else sizes[0]
            end = min(start + size, n)
            for r in range(start, end):
                for c in range(start, (r + 1 if self.attention == "unidirectional" else end)):
                    layout[h, r, c] = 1
            start = end
        return layout

    def set_global_layout(self, h, layout):
        n = layout.shape[1]
        ends = self.global_block_end_indices or [i + 1 for i in self.global_block_indices]
        for s, e in zip(self.global_block_indices, ends):
            if s < n:
                e = min(e, n)
                first_row = 0 if self.attention == "bidirectional" else s
                layout[h, first_row:, s:e] = 1
                if self.horizontal_global_attention:
                    layout[h, s:e, :] = 1
        return layout

    def make_layout(self, seq_len):
        layout = self.setup_layout(seq_len)
        for h in range(self.num_layout_heads):
            layout = self.set_global_layout(h, self.set_local_layout(h, self.set_random_layout(h, layout)))
        return self.check_and_propagate_first_head_layout(layout)


class BigBirdSparsityConfig(SparsityConfig):

    def __init__(self, num_heads, block=16, different_layout_per_head=False, num_random_blocks=1, num_sliding_window_blocks=3,
                 num_global_blocks=1, attention="bidirectional"):
        super().__init__(num_heads, block, different_layout_per_head)
        self.num_random_blocks, self.num_sliding_window_blocks, self.num_global_blocks, self.attention = \
            num_random_blocks, num_sliding_window_blocks, num_global_blocks, attention

    def make_layout(self, seq_len):
        layout = self.setup_layout(seq_len)
        n = layout.shape[1]
        if n < max(self.num_random_blocks, self.num_sliding_window_blocks, self.num_global_blocks):
            raise ValueError("number of random / sliding-window / global blocks must be smaller than blocks in a row")
        w = self.num_sliding_window_blocks // 2
        for h in range(self.num_layout_heads):
            for r in range(n):
                hi = n if self.attention == "bidirectional" else r + 1
                layout[h, r, random.sample(range(hi), min(self.num_random_blocks, hi))] = 1
                layout[h, r, max(0, r - w):min(r + w + 1, n)] = 1
            layout[h, :self.num_global_blocks, :] = 1
            layout[h, :, :self.num_global_blocks] = 1
            if self.attention == "unidirectional":
                layout[h] = torch.tril(layout[h])
        return self.check_and_propagate_first_head_layout(layout)


class BSLongformerSparsityConfig(SparsityConfig):

    def __init__(self, num_heads, block=16, different_layout_per_head=False, num_sliding_window_blocks=3,
                 global_block_indices=(0, ), global_block_end_indices=None, attention="bidirectional"):
        super().__init__(num_heads, block, different_layout_per_head)
        self.num_sliding_window_blocks, self.global_block_indices = num_sliding_window_blocks, list(global_block_indices)
        self.global_block_end_indices = None if global_block_end_indices is None else list(global_block_end_indices)
        self.attention = attention

    def make_layout(self, seq_len):
        layout = self.setup_layout(seq_len)
        n = layout.shape[1]
        w = self.num_sliding_window_blocks // 2
        ends = self.global_block_end_indices or [i + 1 for i in self.global_block_indices]
        for h in range(self.num_layout_heads):
            for r in range(n):
                layout[h, r, max(0, r - w):min(r + w + 1, n)] = 1
            for s, e in zip(self.global_block_indices, ends):
                if s < n:
                    layout[h, s:min(e, n), :] = 1
                    layout[h, :, s:min(e, n)] = 1
            if self.attention == "unidirectional":
                layout[h] = torch.tril(layout[h])
        return self.check_and_propagate_first_head_layout(layout)


class LocalSlidingWindowSparsityConfig(SparsityConfig):

    def __init__(self, num_heads, block=16, num_sliding_window_blocks=3, attention="unidirectional"):
        super().__init__(num_heads, block)
        self.num_sliding_window_blocks, self.attention = num_sliding_window_blocks, attention

    def make_layout(self, seq_len):
        layout = self.setup_layout(seq_len)
        n = layout.shape[1]
        w = self.num_sliding_window_blocks // 2
        for r in range(n):
            end = min(r + w + 1, n) if self.attention == "bidirectional" else r + 1
            layout[0, r, max(0, r - w):end] = 1
        return self.check_and_propagate_first_head_layout(layout)
