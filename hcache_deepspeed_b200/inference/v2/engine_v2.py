"""The ragged inference engine.

Public surface of the reference's ``InferenceEngineV2`` (``deepspeed/inference/v2/engine_v2.py:30-301``):
``put`` (returns ``(logits, per-sequence latents)`` as in the HCache fork, lines 131-189), ``restore_kv``
(lines 107-128), ``query``, ``can_schedule``, ``get_remaining_block_capacity``, ``flush``, ``serialize``,
``free_blocks``, ``n_kv_cache_groups``, ``model``.

Defects of the fork that are NOT reproduced:
* ``put`` works for every model family, not only Llama-2 (the fork unpacks a 2-tuple that only the Llama
  model returns);
* ``restore_kv`` skips ``None`` latents without dereferencing missing sequences, and compares with ``is None``
  rather than an elementwise tensor ``==``;
* capture can be switched off (globally or per call) and does not force a blocking ``.cpu()`` per forward;
* under tensor parallelism only rank 0 copies latents to the host (they are replicated across ranks).
"""
from __future__ import annotations

import json
import os
import pickle
from typing import Iterable, List, Optional, Sequence, Tuple, Union

import torch

from ...utils.logging import inference_logger
from .config_v2 import RaggedInferenceEngineConfig
from .ragged import DSStateManager, PlaceholderSequenceDescriptor, RaggedBatchWrapper
from .scheduling_utils import SchedulingError, SchedulingResult

INFERENCE_MODEL_TIMER = "model-forward-inference"


class LatentHandle:
    """Hidden states of one sequence whose device-to-host copy may still be in flight (capture='async')."""

    def __init__(self, tensor: torch.Tensor, event) -> None:
        self._tensor, self._event = tensor, event

    def wait(self) -> torch.Tensor:
        if self._event is not None:
            self._event.synchronize()
            self._event = None
        return self._tensor

    @property
    def shape(self):
        return self._tensor.shape

    def ready(self) -> bool:
        return self._event is None or self._event.query()


class InferenceEngineV2:

    def __init__(self, policy, engine_config: RaggedInferenceEngineConfig) -> None:
        """``policy``: an :class:`InferenceV2Policy` (model description + checkpoint engine)."""
        self._config = engine_config
        self._policy = policy
        self._base_mp_group = self._initialize_tp_group()
        inference_logger().info("Building model...")
        self._model = self._policy.build_model(self._config, self._base_mp_group)
        inference_logger().info("Model built.")
        kv_cfg = self._model.kv_cache_config()
        group = max(1, self._model.n_q // max(1, self._model.n_kv))
        self._batch = RaggedBatchWrapper(self._config.state_manager, block_size=kv_cfg[0].block_size,
                                         max_blocks_per_seq=kv_cfg[0].max_blocks_per_allocation_group, group=group,
                                         device=self._model.device)
        self._state_manager = DSStateManager(self._config.state_manager, kv_cfg, base_mp_group=self._base_mp_group,
                                             device=self._model.device)
        self._model.set_state_manager(self._state_manager)
        self._capture_stream = torch.cuda.Stream(device=self._model.device) if self._model.device.type == "cuda" else None
        self.last_put_stats = {}
        self._decode_graphs = {}   # (n_seqs, attention splits) -> (CUDAGraph, static logits)
        self._decode_warm = {}     # eager runs seen per key before capturing
        self.graph_replays = 0

    # ------------------------------------------------------------------------------------------------
    def _initialize_tp_group(self):
        tp = self._config.tensor_parallel.tp_size
        local_rank = int(os.getenv("LOCAL_RANK", 0))
        if torch.cuda.is_available():
            torch.cuda.set_device(local_rank % max(1, torch.cuda.device_count()))
        if tp == 1:
            return None
        from ...comm import comm as dist
        dist.init_distributed()
        if local_rank >= tp:
            raise RuntimeError("Local rank is greater than TP size, ensure that the TP config is correct.")
        return dist.new_group(ranks=list(range(tp)))

    @property
    def free_blocks(self) -> torch.Tensor:
        """Free KV blocks per cache group (shape ``[n_kv_cache_groups]``)."""
        return self._state_manager.free_blocks

    @property
    def n_kv_cache_groups(self) -> int:
        return self._state_manager.n_kv_cache_groups

    def model(self):
        return self._model

    @property
    def state_manager(self) -> DSStateManager:
        return self._state_manager

    @property
    def config(self) -> RaggedInferenceEngineConfig:
        return self._config

    # ------------------------------------------------------------------------------------------------
    # HCache restore
    # ------------------------------------------------------------------------------------------------
    def restore_kv(self, batch_uids: Iterable[int], batch_tokens: Iterable[torch.Tensor],
                   batch_latents: Iterable[Optional[Union[torch.Tensor, LatentHandle]]],
                   layers: Optional[Sequence[int]] = None, do_checks: bool = True,
                   defer: Optional[bool] = None, sharded: bool = False, layer_ready=None) -> None:
        """Rebuild the KV cache of ``batch_uids`` from saved hidden states instead of recomputing the prompt.

        ``batch_tokens`` only supply the token counts (ref: engine_v2.py:108-128); ``batch_latents[i]`` is the host
        tensor ``[n_layers, n_tokens_i, hidden]`` returned by earlier ``put`` calls (concatenated along dim 1 by the
        caller if it spans several calls), or None to skip that uid. After the call the sequences have
        ``seen_tokens`` advanced; follow with a normal ``put`` of the new turn's tokens.

        ``defer`` (default: ``hcache.overlap_restore``): only start the transfers now and let the next ``put``
        interleave each layer's K/V projection with its own layers (hides the new turn's prefill under PCIe time).

        ``sharded`` (tensor parallel, one sequence per call): every rank passes only ITS row slice of the hidden
        states (``latent_row_range``), moves 1/tp of the bytes over its own PCIe link, and the projection kernels
        gather the slices from peer memory over NVLink — restore bandwidth scales with the number of GPUs.

        ``layer_ready(i)``: host callback that blocks until saved layer ``i`` has arrived in the (pinned) latent tensors;
        called right before that layer's host-to-device copy is queued, so storage reads issued up front overlap the
        copies and projections of the layers before them (NVMe tier).
        """
        uids, toks, lats = list(batch_uids), list(batch_tokens), list(batch_latents)
        todo = [(u, t, (l.wait() if isinstance(l, LatentHandle) else l)) for u, t, l in zip(uids, toks, lats)
                if l is not None]
        if not todo:
            return
        if do_checks:
            check = self.can_schedule([u for u, _, _ in todo], [int(t.numel()) for _, t, _ in todo])
            if check != SchedulingResult.Success:
                raise SchedulingError(check)
        for u, t, l in todo:
            if sharded:
                continue
            if l.dim() != 3 or int(l.shape[1]) != int(t.numel()):
                raise ValueError(f"latents of uid {u} have shape {tuple(l.shape)}, expected [layers, {int(t.numel())}, hidden]")
        self._batch.clear()
        for uid, tokens, _ in todo:
            seq = self._state_manager.get_or_create_sequence(uid)
            self._model.maybe_allocate_kv(seq, tokens.numel())
            seq.pre_forward(tokens.numel())
            self._batch.insert_sequence(seq, tokens, do_checks=do_checks)
        self._batch.finalize()
        if defer is None:
            defer = self._config.hcache.overlap_restore
        self._model.restore_kv(self._batch, [l for _, _, l in todo], layers=layers, defer=bool(defer), sharded=sharded,
                               layer_ready=layer_ready)
        for uid, _, _ in todo:
            self._state_manager.get_sequence(uid).post_forward()

    def restore_session(self, uid: int, tokens: torch.Tensor, hidden: Optional[torch.Tensor] = None,
                        hidden_layers: Sequence[int] = (), kv: Optional[torch.Tensor] = None,
                        kv_layers: Sequence[int] = (), recompute_layers: Sequence[int] = (),
                        defer: Optional[bool] = None, interleave: float = 2.5, kv_layer_major: bool = False,
                        sharded: bool = False, layer_ready=None) -> None:
        """Bring one session's context back following a mixed plan (``hcache/planner.py``):

        * ``recompute_layers`` — must be the prefix ``0..r-1``: recomputed from ``tokens`` (uses the GPU while the
          host link is the bottleneck);
        * ``kv_layers`` — ``kv [n_blocks, len(kv_layers), 2, kv_heads, block, head_dim]`` copied into the pages; with
          ``kv_layer_major`` the tensor is ``[len(kv_layers), n_blocks, ...]`` and streams layer by layer, overlapped
          with the next forward like the hidden states;
        * ``hidden_layers`` — ``hidden [len(hidden_layers), n_tokens, hidden]`` through the K/V projection kernel.

        ``interleave``: saved layers projected per recomputed layer (layer-forward time / per-layer transfer time).
        ``sharded`` (tensor parallel): ``hidden`` holds only this rank's row slice (``latent_row_range``) and the
        projection gathers the peers' slices over NVLink; ``kv`` always holds this rank's own KV heads.
        """
        tokens = tokens.reshape(-1)
        n = int(tokens.numel())
        rl = list(recompute_layers)
        if rl != list(range(len(rl))):
            raise ValueError("recomputed layers must form a prefix 0..r-1 (layer l needs layer l-1's output)")
        model = self._model
        model.finish_pending_restore()
        if len(rl) + len(list(hidden_layers)) + len(list(kv_layers)) != model.spec.n_layers:
            raise ValueError("the plan does not cover every layer exactly once")
        self._batch.clear()
        seq = self._state_manager.get_or_create_sequence(uid)
        if seq.seen_tokens != 0:
            raise RuntimeError(f"sequence {uid} already holds {seq.seen_tokens} tokens; flush it before restoring")
        model.maybe_allocate_kv(seq, n)
        seq.pre_forward(n)
        self._batch.insert_sequence(seq, tokens)
        self._batch.finalize()
        if defer is None:
            defer = self._config.hcache.overlap_restore
        n_hidden = len(list(hidden_layers))
        # start the host-to-device stream of the hidden states FIRST: it then runs under the recomputed prefix
        if n_hidden:
            model.restore_kv(self._batch, [hidden], layers=list(hidden_layers), defer=True, sharded=sharded,
                             layer_ready=layer_ready)
        if len(list(kv_layers)):
            model.restore_kv_pages(seq.all_block_ids, kv, list(kv_layers), defer=bool(defer) and kv_layer_major,
                                   layer_major=kv_layer_major)
        if rl:
            model.forward(self._batch, capture=False, n_layers_limit=len(rl), restore_interleave=interleave)
        if n_hidden and not defer:
            model.finish_pending_restore()
        seq.post_forward()

    # ------------------------------------------------------------------------------------------------
    # forward
    # ------------------------------------------------------------------------------------------------
    def put(self, batch_uids: Iterable[int], batch_tokens: Iterable[torch.Tensor], do_checks: bool = True,
            capture_latents: Optional[bool] = None, all_logits: bool = False) -> Tuple[torch.Tensor, List]:
        """One forward over a ragged batch. Returns ``(logits [len(batch_uids), vocab], latents)`` where
        ``latents[i]`` is sequence i's ``[n_layers, n_tokens_i, hidden]`` host tensor (pinned), a
        :class:`LatentHandle` in async mode, or None when capture is off."""
        uids, toks = list(batch_uids), list(batch_tokens)
        if do_checks:
            check = self.can_schedule(uids, [int(t.numel()) for t in toks])
            if check != SchedulingResult.Success:
                raise SchedulingError(check)
        self._batch.clear()
        for uid, tokens in zip(uids, toks):
            seq = self._state_manager.get_or_create_sequence(uid)
            self._model.maybe_allocate_kv(seq, tokens.numel())
            seq.pre_forward(tokens.numel())
            self._batch.insert_sequence(seq, tokens, do_checks=do_checks)
        self._batch.finalize()
        self._model.prepare_batch(self._batch)

        hc = self._config.hcache
        capture = capture_latents
        if capture is None:
            capture = hc.enabled and hc.capture != "off"
            if capture and not hc.capture_decode_steps and self._batch.current_tokens == self._batch.current_sequences:
                capture = False
        if (not capture and not all_logits and self._config.use_cuda_graph_decode and self._model.device.type == "cuda"
                and self._batch.current_tokens == self._batch.current_sequences and self._batch.n_items_large == 0):
            logits, latent_dev = self._graphed_decode(), None
        else:
            logits, latent_dev = self._model.forward(self._batch, capture=bool(capture), all_logits=all_logits)
        latents = self._export_latents(latent_dev) if capture else [None] * len(uids)
        assert logits.shape[0] == (self._batch.current_tokens if all_logits else self._batch.current_sequences)
        for uid in uids:
            seq = self._state_manager.get_sequence(uid)
            seq.post_forward()
            self._model.maybe_free_kv(seq)
        return logits, latents

    def _graphed_decode(self) -> torch.Tensor:
        """Pure decode step through a CUDA graph. The batch metadata (token ids, slots, block tables, work items)
        lives in fixed device buffers refreshed by ``finalize`` before every replay; only launch geometry is baked
        in, so one graph serves a given (sequence count, KV split count) until the heuristic wants more splits."""
        from ...ops.backend import pick_decode_splits, pick_packed_splits, use_packed_decode
        model, batch = self._model, self._batch
        model.finish_pending_restore()
        pick = pick_packed_splits if use_packed_decode(model.n_q, model.n_kv, model.d) else pick_decode_splits
        key = (batch.current_sequences, pick(batch.n_items_small, model.n_kv, batch.max_kv_len_small),
               batch.rows_per_small_item)
        entry = self._decode_graphs.get(key)
        if entry is None:
            seen = self._decode_warm.get(key, 0)
            self._decode_warm[key] = seen + 1
            if seen < 2:  # first steps run eagerly (lazy kernel attribute setup, allocator warm-up)
                return model.forward(batch, capture=False)[0]
            from ...ops import kernels as _k
            from ...utils.cuda_graph import capture
            before = _k.launch_count
            try:
                graph, static_logits = capture(lambda: model.forward(batch, capture=False)[0], device=model.device)
            except Exception as e:  # keep serving eagerly; a failed capture launched nothing
                _k.launch_count = before
                inference_logger().warning(f"CUDA-graph capture of the decode step failed ({type(e).__name__}: {e}); "
                                           f"batch shape {key} stays on the eager path")
                self._decode_graphs[key] = entry = (None, None, 0)
                return model.forward(batch, capture=False)[0]
            entry = (graph, static_logits, _k.launch_count - before)
            _k.launch_count = before  # capture launched nothing; replays account for the kernels they run
            self._decode_graphs[key] = entry
        if entry[0] is None:
            return model.forward(batch, capture=False)[0]
        entry[0].replay()
        from ...ops import kernels as _k
        _k.launch_count += entry[2]
        self.graph_replays += 1
        return entry[1]

    def _export_latents(self, latent_dev: torch.Tensor) -> List:
        """Device ``[L, T, H]`` -> pinned host memory, split into per-sequence views (engine_v2.py:172-179).
        Under tensor parallelism the hidden states are replicated: rank 0 exports them (other ranks return None),
        or with ``hcache.shard_capture`` every rank exports its 1/tp row slice of each sequence."""
        shadow = self._batch.inflight_seq_descriptors_shadow
        mode = self._config.hcache.capture
        tp, rank = self._model.tp_size, self._model.tp_rank
        shard = tp > 1 and self._config.hcache.shard_capture
        if tp > 1 and rank != 0 and not shard:
            return [None] * len(shadow)
        spans = [(b, n) for b, n, _ in shadow]
        if shard:
            from ...models.transformer import shard_rows
            spans = []
            for b, n, _ in shadow:
                lo, hi, _ = shard_rows(n, tp, rank)
                spans.append((b + lo, hi - lo))
        if latent_dev.device.type != "cuda":
            return [latent_dev[:, b:b + n].clone() for b, n in spans]
        total = sum(n for _, n in spans)
        host = torch.empty((latent_dev.shape[0], total, latent_dev.shape[2]), dtype=latent_dev.dtype, pin_memory=True)
        comp = torch.cuda.current_stream()
        cs = self._capture_stream
        cs.wait_stream(comp)
        views, off = [], 0
        with torch.cuda.stream(cs):
            contiguous = all(spans[i][0] + spans[i][1] == spans[i + 1][0] for i in range(len(spans) - 1))
            if contiguous and spans:
                host.copy_(latent_dev[:, spans[0][0]:spans[0][0] + total], non_blocking=True)
            for b, n in spans:
                if not contiguous and n:
                    host[:, off:off + n].copy_(latent_dev[:, b:b + n], non_blocking=True)
                views.append(host[:, off:off + n])
                off += n
            ev = torch.cuda.Event()
            ev.record(cs)
        # the model's latent buffer is reused by the next forward: make that forward wait for this copy
        comp.wait_event(ev)
        if mode == "async":
            return [LatentHandle(v, ev) for v in views]
        ev.synchronize()
        return views

    def latent_row_range(self, n_tokens: int) -> Tuple[int, int]:
        """Rows of a sequence's hidden states this TP rank stores / passes to a sharded restore."""
        from ...models.transformer import shard_rows
        lo, hi, _ = shard_rows(n_tokens, self._model.tp_size, self._model.tp_rank)
        return lo, hi

    # ------------------------------------------------------------------------------------------------
    # scheduling helpers
    # ------------------------------------------------------------------------------------------------
    def query(self, uid: int, max_request_tokens: int, max_request_blocks) -> Tuple[int, int]:
        """Tokens and KV blocks that scheduling up to ``max_request_tokens`` for ``uid`` would take."""
        seq = self._state_manager.get_sequence(uid)
        if seq is None:
            if self._state_manager.n_tracked_sequences == self._config.state_manager.max_tracked_sequences:
                return (0, 0)
            seq = PlaceholderSequenceDescriptor()
        return self._model.get_kv_requirements(seq, max_request_tokens, max_request_blocks)

    def can_schedule(self, uids: Iterable[int], lengths: Iterable[int]) -> SchedulingResult:
        """Dry-run a batch (ref: engine_v2.py:217-264)."""
        uids, lengths = list(uids), list(lengths)
        cfg = self._config.state_manager
        cur_seqs = self._state_manager.n_tracked_sequences
        free_blocks = int(self._state_manager.free_blocks[0])
        batch_len = 0
        if len(uids) > cfg.max_ragged_sequence_count:
            return SchedulingResult.BatchSequenceLimitExceeded
        for uid, length in zip(uids, lengths):
            seq = self._state_manager.get_sequence(uid)
            if seq is None:
                cur_seqs += 1
                seq = PlaceholderSequenceDescriptor()
            if seq.seen_tokens + length > cfg.max_context:
                return SchedulingResult.SequenceTokenLimitExceeded
            sched_len, sched_blocks = self._model.get_kv_requirements(seq, length, free_blocks)
            if sched_len != length:
                return SchedulingResult.KVCacheLimitExceeded
            batch_len += length
            free_blocks -= sched_blocks
        if cur_seqs > cfg.max_tracked_sequences:
            return SchedulingResult.EngineSequenceLimitExceeded
        if batch_len > cfg.max_ragged_batch_size:
            return SchedulingResult.BatchTokenLimitExceeded
        return SchedulingResult.Success

    def get_remaining_block_capacity(self, uid: int) -> int:
        seq = self._state_manager.get_sequence(uid)
        if seq is None:
            return 0
        return self._model.get_remaining_block_capacity(seq)

    def flush(self, uid: int) -> None:
        """Remove all state of a sequence and release its KV blocks."""
        self._model.finish_pending_restore()  # a deferred restore may still target these blocks
        self._state_manager.flush_sequence(uid)

    # ------------------------------------------------------------------------------------------------
    # serialisation (ref: engine_v2.py:284-301 + model_implementations/flat_model_helpers.py:59-282)
    # ------------------------------------------------------------------------------------------------
    def serialize(self, save_path: str) -> None:
        from ...models.flat_model import flatten_model_weights, make_metadata_filename, make_param_filename
        os.makedirs(save_path, exist_ok=True)
        rank, size = self._model.tp_rank, self._model.tp_size
        flat, meta = flatten_model_weights(self._model.w)
        torch.save(flat.cpu(), make_param_filename(save_path, rank, size))
        with open(make_metadata_filename(save_path, rank, size), "w") as f:
            json.dump(meta, f)
        if rank == 0:
            with open(os.path.join(save_path, "ds_model_config.pkl"), "wb") as f:
                pickle.dump({"arch": self._model.spec, "dtype": str(self._model.dtype).replace("torch.", "")}, f)
