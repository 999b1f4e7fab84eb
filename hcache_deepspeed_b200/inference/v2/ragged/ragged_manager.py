"""Tracks every live sequence and owns the paged KV cache.

Counterpart of ``DSStateManager`` (``deepspeed/inference/v2/ragged/ragged_manager.py:19-213``), with
whole-sequence KV offload / reload implemented (the reference's ``offload`` config flag is inert).
"""
from __future__ import annotations

from typing import Any, Dict, Optional, Tuple

import torch

from ....utils.logging import inference_logger
from .blocked_allocator import BlockedAllocator
from .kv_cache import BlockedKVCache
from .manager_configs import DSStateManagerConfig, KVCacheConfig
from .sequence_descriptor import DSSequenceDescriptor


class DSStateManager:

    def __init__(self, config: DSStateManagerConfig, kv_configs: Tuple[KVCacheConfig, ...], base_mp_group: Optional[Any] = None,
                 device: Optional[torch.device] = None) -> None:
        self._config = config
        self._kv_configs = kv_configs
        self._seqs: Dict[int, DSSequenceDescriptor] = {}
        self._free_tracking_ids = list(range(config.max_tracked_sequences - 1, -1, -1))
        self._kv_cache = BlockedKVCache(kv_configs, config.memory_config, mp_group=base_mp_group, offload=config.offload,
                                        device=device)
        self._allocator = BlockedAllocator(self._kv_cache.num_blocks)
        self._kv_cache.bind_allocator(self._allocator)

    # ---- sequences --------------------------------------------------------------------------------
    def get_sequence(self, uid: int) -> Optional[DSSequenceDescriptor]:
        return self._seqs.get(uid, None)

    def get_or_create_sequence(self, uid: int) -> DSSequenceDescriptor:
        seq = self._seqs.get(uid, None)
        if seq is not None:
            return seq
        return self._create_sequence(uid)

    def _create_sequence(self, uid: int) -> DSSequenceDescriptor:
        if uid in self._seqs:
            raise ValueError(f"Sequence {uid} already exists.")
        if not self._free_tracking_ids:
            raise RuntimeError(f"Unable to create tracking slot for sequence {uid} since the metadata buffers are full.")
        tid = self._free_tracking_ids.pop()
        seq = DSSequenceDescriptor(tid, self._kv_configs[0].max_blocks_per_allocation_group)
        self._seqs[uid] = seq
        return seq

    def flush_sequence(self, uid: int) -> None:
        """Forget a sequence and release its KV blocks."""
        seq = self._seqs.pop(uid, None)
        if seq is None:
            inference_logger().warning(f"Attempting to flush sequence {uid} which does not exist.")
            return
        ids = seq.free_kv_cache()
        if ids.numel():
            self._allocator.free(ids)
        self._free_tracking_ids.append(seq.tracking_id)

    @property
    def tracked_sequences(self) -> Dict[int, DSSequenceDescriptor]:
        return self._seqs

    @property
    def n_tracked_sequences(self) -> int:
        return len(self._seqs)

    # ---- KV cache ---------------------------------------------------------------------------------
    @property
    def kv_block_size(self) -> int:
        return self._kv_configs[0].block_size

    @property
    def n_kv_cache_groups(self) -> int:
        return 1

    @property
    def free_blocks(self) -> torch.Tensor:
        return torch.tensor([self._allocator.free_blocks], dtype=torch.int32)

    def allocate_blocks(self, n_blocks: int, cache_group: int = 0) -> torch.Tensor:
        return self._allocator.allocate(n_blocks)

    def free_blocks_ids(self, ids: torch.Tensor) -> None:
        self._allocator.free(ids)

    def get_cache(self, cache_id: int, cache_group: int = 0) -> torch.Tensor:
        return self._kv_cache.get_cache(cache_id, cache_group)

    @property
    def kv_cache(self) -> BlockedKVCache:
        return self._kv_cache

    def restore_kv(self, uid: int, latent, qk_layer, qk_w, b=None, start: int = 0) -> int:
        """Rebuild a sequence's K/V pages from saved per-layer hidden states with a caller-supplied projection.

        ``latent[l]`` is ``[tokens, hidden]`` for layer ``l``; ``qk_layer(latent_l, qk_w[l], b=bias)`` returns that layer's
        ``[tokens, 2 * kv_heads * head_dim]`` (K then V; positional encoding already applied). Pages are allocated as needed
        and the sequence's seen-token count advanced. This is the model-agnostic slow path; the model's own ``restore_kv``
        (models/transformer.py) is the pipelined one the engine uses. The reference has this method as an unfinished stub
        (ragged_manager.py:170) - here it does what its signature says. Returns the number of tokens restored.
        """
        seq = self.get_or_create_sequence(uid)
        n_tok = int(latent[0].shape[0])
        bs = self.kv_block_size
        need = -(-(start + n_tok) // bs) - seq.cur_allocated_blocks
        if need > 0:
            seq.extend_kv_cache(self._allocator.allocate(need))
        ids = seq.all_block_ids.to(torch.long)
        pos = torch.arange(start, start + n_tok)
        blk, off = ids[pos // bs], pos % bs
        for layer_id, latent_layer in enumerate(latent):
            cache = self.get_cache(layer_id)                      # [blocks, 2, heads, block_size, head_dim]
            w = qk_w[layer_id] if isinstance(qk_w, (list, tuple)) else qk_w
            bias = b[layer_id] if isinstance(b, (list, tuple)) else b
            kv = qk_layer(latent_layer.to(cache.device), w, b=bias)
            heads, hd = cache.shape[2], cache.shape[4]
            kv = kv.reshape(n_tok, 2, heads, hd).to(cache.dtype)
            dblk, doff = blk.to(cache.device), off.to(cache.device)
            cache[dblk, 0, :, doff] = kv[:, 0]
            cache[dblk, 1, :, doff] = kv[:, 1]
        if start + n_tok > seq.seen_tokens:
            seq.pre_forward(start + n_tok - seq.seen_tokens)
            seq.post_forward()
        return n_tok

    # ---- whole-sequence KV offload (plain KV offload path) --------------------------------------------
    def offload_sequence(self, uid: int) -> torch.Tensor:
        """Move a sequence's KV pages to pinned host memory and release its device blocks."""
        if not self._config.offload:
            raise RuntimeError("KV offload is disabled: set state_manager.offload = True")
        seq = self._seqs[uid]
        if seq.offloaded_kv is not None:
            return seq.offloaded_kv
        ids = seq.all_block_ids.clone()
        host = self._kv_cache.offload(ids)
        if host.is_pinned() and torch.cuda.is_available():
            torch.cuda.current_stream().synchronize()
        self._allocator.free(seq.free_kv_cache())
        seq.offloaded_kv = host
        return host

    def reload_sequence(self, uid: int, host_kv: Optional[torch.Tensor] = None) -> None:
        """Bring an offloaded sequence back into freshly allocated device blocks."""
        seq = self._seqs[uid]
        host = host_kv if host_kv is not None else seq.offloaded_kv
        if host is None:
            raise RuntimeError(f"sequence {uid} has no offloaded KV")
        ids = self._allocator.allocate(host.shape[0])
        seq.extend_kv_cache(ids)
        self._kv_cache.restore(ids, host)
        seq.offloaded_kv = None
