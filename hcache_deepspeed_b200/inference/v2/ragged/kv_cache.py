"""Paged KV cache.

One device tensor ``[num_layers, num_blocks, 2 (K,V), n_kv_heads_local, block_size, head_dim]``.
Compared with the reference layout ``(layers, blocks, block_size, 2, heads, dim)``
(``deepspeed/inference/v2/ragged/kv_cache.py:40-208``) the token axis sits inside the head axis, so
the K (or V) page of one head is one contiguous ``block_size x head_dim`` tile: exactly one
SWIZZLE_128B TMA box per 64 head dims, and directly a UMMA operand (K: K-major, V: MN-major).

The cache is zero-filled at allocation: the tensor-core attention kernel multiplies masked (zero)
probabilities with whatever lies in the unused tail of a page, which therefore must be finite.

Implemented here and absent in the reference (``kv_cache.py:166-184`` raise NotImplementedError):
``offload`` / ``restore`` of whole blocks to pinned host memory (plain KV offload, the alternative the
HCache planner weighs against hidden-state restore).
"""
from __future__ import annotations

import operator
from functools import reduce
from typing import Iterable, Optional, Tuple

import torch

from ....utils.memory import pinned_empty
from ....utils.logging import inference_logger
from .manager_configs import AllocationMode, KVCacheConfig, MemoryConfig


def split_kv(kv_cache_layer: torch.Tensor) -> Tuple[torch.Tensor, torch.Tensor]:
    """Views of K and V of one layer: each ``[blocks, heads, block_size, head_dim]``."""
    if kv_cache_layer.dim() != 5 or kv_cache_layer.shape[1] != 2:
        raise ValueError(f"KV-cache layer must be [blocks, 2, heads, block_size, head_dim], got {kv_cache_layer.shape}")
    return kv_cache_layer[:, 0], kv_cache_layer[:, 1]


class BlockedKVCache:

    def __init__(self, configs: Tuple[KVCacheConfig, ...], memory_config: MemoryConfig, mp_group=None,
                 offload: bool = False, device: Optional[torch.device] = None) -> None:
        if len(configs) != 1:
            raise ValueError("exactly one KV cache group is supported")
        self._configs = configs
        self._memory_config = memory_config
        self._enable_offload = offload
        cfg = configs[0]
        self._device = device if device is not None else (torch.device("cuda", torch.cuda.current_device())
                                                          if torch.cuda.is_available() else torch.device("cpu"))
        num_layers, n_heads, head_dim = cfg.cache_shape
        dtype = cfg.cache_dtype.value if hasattr(cfg.cache_dtype, "value") else cfg.cache_dtype
        if not isinstance(dtype, torch.dtype):
            from ....config import DtypeEnum
            dtype = DtypeEnum(dtype).value
        self._dtype = dtype
        elem = torch.tensor([], dtype=dtype).element_size()
        self._block_bytes_per_layer = 2 * n_heads * cfg.block_size * head_dim * elem
        per_block_bytes = num_layers * self._block_bytes_per_layer

        mode = memory_config.mode
        mode = mode if isinstance(mode, AllocationMode) else AllocationMode(mode)
        if mode == AllocationMode.RESERVE:
            if self._device.type == "cuda":
                free, _total = torch.cuda.mem_get_info(self._device)
                usable = free - memory_config.size
            else:
                usable = 256 * 1024 * 1024
            if usable <= 0:
                raise ValueError(f"Insufficient memory for the KV cache: reserve {memory_config.size} bytes leaves none")
        else:
            usable = memory_config.size
        num_blocks = int(usable // per_block_bytes)
        if mp_group is not None:
            # every tensor-parallel rank must agree on the block count
            import torch.distributed as dist
            if dist.is_available() and dist.is_initialized() and dist.get_world_size(group=mp_group) > 1:
                t = torch.tensor([num_blocks], dtype=torch.int32, device=self._device)
                dist.all_reduce(t, op=dist.ReduceOp.MIN, group=mp_group)
                num_blocks = int(t.item())
        if num_blocks < 1:
            raise ValueError("KV cache memory budget is smaller than one block")
        self._num_blocks = num_blocks
        inference_logger().info(
            f"Allocating KV-cache: {num_blocks} blocks x {per_block_bytes} B = {num_blocks * per_block_bytes / 2**30:.2f} GiB "
            f"(layers={num_layers}, kv_heads={n_heads}, block={cfg.block_size}, head_dim={head_dim}, {dtype})")
        self._cache = torch.zeros((num_layers, num_blocks, 2, n_heads, cfg.block_size, head_dim), dtype=dtype,
                                 device=self._device)

    # ---- accessors ---------------------------------------------------------------------------
    @property
    def num_blocks(self) -> int:
        return self._num_blocks

    @property
    def num_caches(self) -> int:
        return 1

    @property
    def block_size(self) -> int:
        return self._configs[0].block_size

    @property
    def block_bytes_all_layers(self) -> int:
        return self._configs[0].cache_shape[0] * self._block_bytes_per_layer

    def get_cache(self, cache_id: int = 0, cache_group: int = 0) -> torch.Tensor:
        """Layer ``cache_id``'s pages: ``[blocks, 2, heads, block_size, head_dim]``."""
        return self._cache[cache_id]

    @property
    def full_cache(self) -> torch.Tensor:
        return self._cache

    # ---- whole-block offload / restore ------------------------------------------------------------
    def offload(self, blocks: Iterable[int], host_out: Optional[torch.Tensor] = None, stream=None) -> torch.Tensor:
        """Copy the listed blocks (all layers) into pinned host memory ``[n, layers, 2, heads, bs, d]``."""
        ids = torch.as_tensor(list(blocks) if not isinstance(blocks, torch.Tensor) else blocks, dtype=torch.int32)
        n = ids.numel()
        shape = (n, ) + (self._cache.shape[0], ) + tuple(self._cache.shape[2:])
        if host_out is None:
            host_out = pinned_empty(shape, self._dtype)
        if self._device.type != "cuda":
            host_out.copy_(self._cache[:, ids.long()].permute(1, 0, 2, 3, 4, 5))
            return host_out
        from ....ops import kernels
        staging = torch.empty(shape, dtype=self._dtype, device=self._device)
        kernels.kv_blocks_copy(self._cache, staging, ids.to(self._device), to_staging=True)
        host_out.copy_(staging, non_blocking=True)
        return host_out

    def offload_layers(self, blocks: Iterable[int], layers, layer_major: bool = False) -> torch.Tensor:
        """Pages of selected layers only: pinned ``[n_blocks, len(layers), 2, heads, block_size, head_dim]``, or with
        ``layer_major`` ``[len(layers), n_blocks, ...]`` (every layer contiguous: the layout the layer-pipelined reload
        streams). The KV half of a mixed HCache plan; cold path — runs once at the end of a turn."""
        ids = torch.as_tensor(list(blocks) if not isinstance(blocks, torch.Tensor) else blocks, dtype=torch.long)
        lay = torch.as_tensor(list(layers), dtype=torch.long)
        sel = self._cache[lay.to(self._device)][:, ids.to(self._device)]          # [n_l, n_b, 2, h, bs, d]
        sel = sel.contiguous() if layer_major else sel.permute(1, 0, 2, 3, 4, 5).contiguous()
        host = pinned_empty(tuple(sel.shape), self._dtype)
        host.copy_(sel)
        return host

    # ---- block bookkeeping (the allocator belongs to the state manager, which binds it here; ref: kv_cache.py:144-201) ------
    def bind_allocator(self, allocator) -> None:
        self._allocator = allocator

    def reserve(self, num_blocks: int, cache_group: int = 0) -> torch.Tensor:
        return self._allocator.allocate(num_blocks)

    def free(self, blocks: Iterable[int], cache_group: int = 0) -> None:
        self._allocator.free(blocks if isinstance(blocks, torch.Tensor) else torch.as_tensor(list(blocks), dtype=torch.int32))

    @property
    def free_blocks(self):
        return [self._allocator.free_blocks]

    def restore(self, blocks: Iterable[int], host_in: torch.Tensor) -> None:
        """Inverse of :meth:`offload` into (possibly different) block ids."""
        ids = torch.as_tensor(list(blocks) if not isinstance(blocks, torch.Tensor) else blocks, dtype=torch.int32)
        if self._device.type != "cuda":
            self._cache[:, ids.long()] = host_in.permute(1, 0, 2, 3, 4, 5).to(self._dtype)
            return
        from ....ops import kernels
        staging = host_in.to(self._device, non_blocking=True)
        kernels.kv_blocks_copy(self._cache, staging, ids.to(self._device), to_staging=False)
