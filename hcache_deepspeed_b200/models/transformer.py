"""The single model executor: runs any :class:`ArchSpec` over a ragged batch.

Plays the role of the reference's ``DSTransformerModelBase`` + per-family ``model.py``
(``deepspeed/inference/v2/model_implementations/inference_transformer_base.py:48-623``,
``llama_v2/model.py:118-252``, ``opt/model.py:101-197`` ...). Design points:

* All activations live in buffers sized once for ``max_ragged_batch_size`` tokens.
* Kernel count per layer (dense, no TP): norm -> QKV(+RoPE+KV append) -> attention -> O-proj ->
  norm -> up(+act/gate) -> down. Residual adds never run as kernels of their own: a row-parallel
  GEMM leaves its result as a *pending delta* (16-bit, or fp32 partial sums when split-K /
  tensor-parallel) and the next norm kernel folds it into the residual stream while normalising.
* HCache capture is free: with capture on, the attention-norm kernel of layer ``l`` writes its output
  straight into ``latent_buf[l]`` — the QKV GEMM reads it from there — so "saving the hidden states"
  is one async D2H of that buffer (the reference allocates a fresh ``[L, T, H]`` tensor per forward,
  copies every layer into it and then blocks on ``.cpu()``: ``llama_v2/model.py:203-220``).
* HCache restore projects ONLY K and V (a row-slice view of ``w_qkv``; the reference runs the full QKV
  GEMM and discards Q, ``llama_v2/model.py:239``), fused with RoPE and the paged write, and is
  layer-pipelined against the H2D stream with a ring of staging slots and reused events.
"""
from __future__ import annotations

import os
from typing import List, Optional, Sequence, Tuple

import torch

from ..inference.v2.ragged import DSStateManager, KVCacheConfig, RaggedBatchWrapper
from ..inference.v2.ragged.manager_configs import DSStateManagerConfig
from ..ops import reference as ref
from ..ops.backend import MAX_DECODE_SPLITS, get_ops, pick_decode_splits
from ..ops.fp8 import Fp8Weight
from ..ops.kernels import ACT_CODES
from .arch import ArchSpec
from .weights import ModelWeights, shard_info

#: row-parallel GEMMs switch from fp32 partial-sum accumulation (split-K) to plain 16-bit stores above this
ACCUM_MAX_TOKENS = 256
FP8_ACCUM_MAX_TOKENS = 64   # fp8 weights: the fp32-accumulate epilogue exists on the decode (feature-major) kernel only
#: staging slots of the restore pipeline (H2D of layer l+1.. overlaps the projection of layer l)
RESTORE_SLOTS = 4
#: tensor parallel: up to this many tokens the collective after a row-parallel GEMM is a one-shot all-reduce inside the
#: norm kernel (one flag exchange, every rank reduces every row); above it reduce-scatter + all-gather (each rank reduces
#: only its rows: tp x fewer NVLink bytes, one more flag exchange)
TP_ONESHOT_MAX_TOKENS = int(os.environ.get("HDS_TP_ONESHOT_TOKENS", "16"))


def ceil_div(a: int, b: int) -> int:
    return -(-a // b)


def shard_rows(n_tokens: int, tp_size: int, tp_rank: int) -> Tuple[int, int, int]:
    """Row range ``[start, end)`` of a sequence's hidden states owned by a TP rank, and the per-rank stride.
    Strides are multiples of 128 (one GEMM m-tile) so a tile never straddles two ranks."""
    per = ceil_div(ceil_div(n_tokens, tp_size), 128) * 128
    start = min(n_tokens, tp_rank * per)
    return start, min(n_tokens, start + per), per


class TPComm:
    """How row-parallel partial sums meet across tensor-parallel ranks (see parallel/tp_comm.py)."""
    fused = False
    mode = "nccl"

    def __init__(self, group=None, size: int = 1, rank: int = 0):
        self.group, self.size, self.rank = group, size, rank

    def all_reduce_delta(self, delta32: torch.Tensor) -> None:
        if self.size > 1:
            import torch.distributed as dist
            dist.all_reduce(delta32, group=self.group)

    def all_gather_logits(self, local: torch.Tensor, out_full: torch.Tensor, vocab_ranges) -> torch.Tensor:
        """Vocabulary-sharded logits -> full ``[n, vocab]`` on every rank (ref: llama_v2/model.py:195-197).
        Uneven vocabulary shards are padded to the widest one for the collective."""
        import torch.distributed as dist
        widths = [b - a for a, b in vocab_ranges]
        wmax = max(widths)
        n = local.shape[0]
        src = local
        if local.shape[1] != wmax:
            src = torch.zeros((n, wmax), dtype=local.dtype, device=local.device)
            src[:, :local.shape[1]] = local
        parts = torch.empty((self.size, n, wmax), dtype=local.dtype, device=local.device)
        if local.device.type == "cuda":
            dist.all_gather_into_tensor(parts, src.contiguous(), group=self.group)
        else:  # gloo wants a list of per-rank outputs
            dist.all_gather(list(parts.unbind(0)), src.contiguous(), group=self.group)
        if all(w == wmax for w in widths):
            out_full.copy_(parts.permute(1, 0, 2).reshape(n, -1))
        else:
            for r, (a, b) in enumerate(vocab_ranges):
                out_full[:, a:b] = parts[r, :, :b - a]
        return out_full


class RaggedTransformer:

    def __init__(self, spec: ArchSpec, weights: ModelWeights, state_config: DSStateManagerConfig,
                 dtype: torch.dtype = torch.bfloat16, device: Optional[torch.device] = None, tp_rank: int = 0,
                 tp_size: int = 1, comm: Optional[TPComm] = None, capture_latents: bool = True) -> None:
        self.spec, self.w, self.dtype = spec, weights, dtype
        self.device = device if device is not None else weights.embed.device
        self.tp_rank, self.tp_size = tp_rank, tp_size
        self.comm = comm if comm is not None else TPComm(size=tp_size, rank=tp_rank)
        self.info = shard_info(spec, tp_rank, tp_size)
        self.ops = get_ops(self.device)
        self._state_config = state_config
        self._state_manager: Optional[DSStateManager] = None
        self._pending_restore = None
        self._pending_kv = None      # deferred layer-pipelined reload of offloaded KV pages
        self._kv_stage = None
        self.capture_enabled = capture_latents
        self.n_q, self.n_kv, self.d = self.info.n_q, self.info.n_kv, spec.head_dim
        self.scale = spec.attn_scale if spec.attn_scale is not None else ref.attention_scale(spec.head_dim)
        self.act = ACT_CODES.get(spec.act, None)
        if self.act is None:
            raise ValueError(f"unsupported activation {spec.act}")
        self.inv_freq = None
        if spec.pos_emb == "rope":
            self.inv_freq = self._make_inv_freq().to(self.device)
        self.alibi = None
        if spec.pos_emb == "alibi":      # slopes of THIS rank's query heads; the kernels add slope * (key - query position)
            slopes = ref.alibi_slopes(spec.n_heads) * float(spec.alibi_scale)
            self.alibi = slopes[torch.tensor(self.info.q_heads)].contiguous().to(self.device)
        self._alloc_buffers()

    # ------------------------------------------------------------------------------------------------
    # construction helpers
    # ------------------------------------------------------------------------------------------------
    def _make_inv_freq(self) -> torch.Tensor:
        s = self.spec
        inv = ref.default_inv_freq(s.rot_dim, s.rope_theta)
        rs = s.rope_scaling or {}
        kind = rs.get("rope_type", rs.get("type"))
        if kind == "linear":
            inv = inv / float(rs.get("factor", 1.0))
        elif kind == "llama3":
            import math
            factor, lo, hi = rs.get("factor", 8.0), rs.get("low_freq_factor", 1.0), rs.get("high_freq_factor", 4.0)
            old = rs.get("original_max_position_embeddings", 8192)
            wavelen = 2 * math.pi / inv
            smooth = ((old / wavelen) - lo) / (hi - lo)
            scaled = torch.where(wavelen > old / lo, inv / factor, inv)
            mid = (wavelen <= old / lo) & (wavelen >= old / hi)
            inv = torch.where(mid, (1 - smooth) * inv / factor + smooth * inv, scaled)
        return inv.float()

    @property
    def kv_block_size(self) -> int:
        return 128 if self.d <= 64 else 64

    def _alloc_buffers(self) -> None:
        s, c = self.spec, self._state_config
        T, S, H = c.max_ragged_batch_size, c.max_ragged_sequence_count, s.hidden
        dev, dt = self.device, self.dtype

        def buf(*shape, dtype=dt):
            return torch.zeros(shape, dtype=dtype, device=dev)

        inter = max(self.info.inter, self.info.moe_inter_range[1] - self.info.moe_inter_range[0],
                    self.info.shared_inter_range[1] - self.info.shared_inter_range[0])
        qkv_cols = (self.n_q + 2 * self.n_kv) * self.d
        self._tp_fused = self.tp_size > 1 and bool(getattr(self.comm, "fused", False))
        self._tp_wait_next = False   # the latest norm output was all-gathered by the peers: its first reader must wait
        #: reduce-scatter + all-gather norms need every rank to write into every rank's activation buffers
        self._tp_scatter_ok = self._tp_fused and not s.is_moe and not s.parallel_residual and s.final_norm
        sym = (lambda *shape, dtype=dt: self.comm.alloc_symmetric(shape, dtype)) if self._tp_scatter_ok else buf
        self.residual = buf(T, H)
        self.embed_tmp = buf(T, H) if s.embed_norm else None     # raw embeddings before their LayerNorm (Bloom)
        self.xn = sym(T, H)
        self.xn2 = sym(T, H)
        self.latent_buf = sym(s.n_layers, T, H) if self.capture_enabled else None
        #: 16-bit partial sums of the store-epilogue (prefill-size) row-parallel GEMMs, reduced by the next norm kernel
        self.partial16 = self.comm.alloc_symmetric((T, H), dt) if self._tp_fused else None
        self.qkv_buf = buf(T, qkv_cols)
        self.q_buf = buf(T, self.n_q * self.d)
        self.attn_buf = buf(T, self.n_q * self.d)
        self.act_buf = buf(T, inter)
        # fp32 pending-delta buffers; two alternate under tensor parallelism (symmetric memory when fused) so a
        # fast rank's next reducing GEMM never lands in a buffer a slow rank is still consuming
        if self._tp_fused:
            self._delta_bufs = self.comm.alloc_delta(T, H, 2)
        else:
            self._delta_bufs = [buf(T, H, dtype=torch.float32) for _ in range(2 if self.tp_size > 1 else 1)]
        self._delta_idx = 0
        self.delta32 = self._delta_bufs[0]
        self.delta16 = buf(T, H)
        self.hidden_last = buf(S, H)
        self.logits_local = self.comm.alloc_symmetric((S, self.info.vocab), dt) if self._tp_fused else buf(S, self.info.vocab)
        self.logits_full = buf(S, s.vocab) if self.tp_size > 1 else None
        self.kv_scratch = None
        self.ws_gemm = self.ws_small = self.ws_large = None
        self.io_stream = None
        if dev.type == "cuda":
            k = self.ops.k
            n_max = max(qkv_cols, (2 if s.gated_mlp else 1) * inter, self.info.vocab, H)
            self.ws_gemm = torch.zeros(k.gemm_splitk_workspace_bytes(128, n_max), dtype=torch.uint8, device=dev)
            G = max(1, self.n_q // max(1, self.n_kv))
            rows = min(8, max(1, (8 // G) * G))
            worst = 16
            for n in sorted({1, 2, 4, 8, 16, 32, 64, 128, 256, S}):
                if n <= S:
                    worst = max(worst, k.splitkv_workspace_bytes(n, self.n_q, self.n_kv, self.d,
                                                                 pick_decode_splits(n, self.n_kv, 1 << 30), rows))
            worst = max(worst, k.splitkv_workspace_bytes(S, self.n_q, self.n_kv, self.d, 1, rows))
            self.ws_small = torch.zeros(worst, dtype=torch.uint8, device=dev)
            worst_l = 16
            for n in range(1, 149):
                units = n * self.n_q
                if units >= 148:
                    break
                sp = min(16, -(-148 // units))
                worst_l = max(worst_l, k.prefill_workspace_bytes(n, self.n_q, self.d, sp))
            from ..ops.backend import pick_packed_splits, use_packed_decode
            if use_packed_decode(self.n_q, self.n_kv, self.d):   # head-packed GQA decode parks its split-KV partials here too
                for n in range(1, S + 1):
                    sp = pick_packed_splits(n, self.n_kv, 1 << 30)
                    if sp > 1:
                        worst_l = max(worst_l, k.prefill_workspace_bytes(n, self.n_kv, self.d, sp))
            self.ws_large = torch.zeros(worst_l, dtype=torch.uint8, device=dev)
            self.io_stream = torch.cuda.Stream(device=dev)
            self.restore_stage = buf(RESTORE_SLOTS, T, H)
            self.restore_stage_sym = None
            self._slot_peer_ptrs = {}
            if self.tp_size > 1 and getattr(self.comm, "fused", False):
                # per-rank slice of the staging ring in symmetric memory: peers' restore kernels read it over NVLink
                # "copy" gather: full-size slots, each rank fills its own rows and pulls the peers' rows (one P2P read of
                # every byte); "kernel" gather: per-rank slices that the projection kernel reads through peer tensor maps
                import os
                self.tp_gather = os.environ.get("HDS_TP_GATHER", "copy")
                self._sym_rows = T if self.tp_gather == "copy" else ceil_div(ceil_div(T, self.tp_size), 128) * 128
                self.restore_stage_sym = self.comm.alloc_symmetric((RESTORE_SLOTS, self._sym_rows, H), dt)
            self.restore_pos = torch.zeros(T, dtype=torch.int32, device=dev)
            self.restore_slot = torch.zeros(T, dtype=torch.int32, device=dev)
            self.ev_loaded = [torch.cuda.Event() for _ in range(RESTORE_SLOTS)]
            self.ev_consumed = [torch.cuda.Event() for _ in range(RESTORE_SLOTS)]
            if not (self.d in (64, 128) and (self.inv_freq is None or s.rot_dim == self.d)):
                self.kv_scratch = buf(T, 2 * self.n_kv * self.d)

    # ------------------------------------------------------------------------------------------------
    # DSInferenceModelBase-style KV accounting (ref: inference_transformer_base.py:336-387)
    # ------------------------------------------------------------------------------------------------
    def kv_cache_config(self) -> Tuple[KVCacheConfig, ...]:
        bs = self.kv_block_size
        max_blocks = ceil_div(self._state_config.max_context, bs)
        return (KVCacheConfig(block_size=bs, cache_shape=(self.spec.n_layers, self.n_kv, self.d),
                              cache_dtype=self.dtype, max_blocks_per_allocation_group=max_blocks), )

    def set_state_manager(self, state_manager: DSStateManager) -> None:
        self._state_manager = state_manager

    @property
    def state_manager(self) -> DSStateManager:
        return self._state_manager

    def get_kv_requirements(self, sequence, max_new_tokens: int, max_new_blocks) -> Tuple[int, int]:
        max_new_blocks = int(max_new_blocks.reshape(-1)[0]) if isinstance(max_new_blocks, torch.Tensor) else int(max_new_blocks)
        bs = self.kv_block_size
        total = sequence.seen_tokens + max_new_tokens
        need = ceil_div(total, bs) - sequence.cur_allocated_blocks
        if need <= max_new_blocks:
            return max_new_tokens, max(need, 0)
        capacity = (max_new_blocks + sequence.cur_allocated_blocks) * bs - sequence.seen_tokens
        return max(capacity, 0), max_new_blocks

    def get_remaining_block_capacity(self, sequence) -> int:
        return (-sequence.seen_tokens) % self.kv_block_size

    def maybe_allocate_kv(self, sequence, n_new_tokens: int) -> None:
        need = ceil_div(sequence.seen_tokens + n_new_tokens, self.kv_block_size) - sequence.cur_allocated_blocks
        if need <= 0:   # the common decode step: the current block still has room
            return
        free = self._state_manager.free_blocks
        _, n_blocks = self.get_kv_requirements(sequence, n_new_tokens, free)
        if n_blocks > 0:
            sequence.extend_kv_cache(self._state_manager.allocate_blocks(n_blocks))

    def maybe_free_kv(self, sequence) -> None:
        """Dense attention keeps every block until the sequence is flushed."""
        return

    def prepare_batch(self, batch: RaggedBatchWrapper) -> None:
        """Attention work items are produced by ``batch.finalize`` (native); nothing left to do here."""
        return

    # ------------------------------------------------------------------------------------------------
    # W8A8 (fp8) weights
    # ------------------------------------------------------------------------------------------------
    def quantize_weights_fp8(self, fmt: str = "e4m3") -> int:
        """``quantization_mode = "wf8af8"`` (``fmt="e4m3"``) or ``"wf6af16"`` (``fmt="e3m2"``: FP6 weights, 0.75 byte each in
        HBM): convert every layer's QKV / attention-out / MLP weights to the tensor-core float format with
        per-output-channel scales (embedding, LM head and norms stay 16-bit); activations are quantised to E4M3 per token on
        the fly and the GEMMs run on ``tcgen05.mma.kind::f8f6f4``. FP6 needs an input dimension that is a multiple of 128
        (a TMA rule of the packed 6-bit type); a weight that does not qualify (odd tensor-parallel shards) stays E4M3.
        Returns the bytes saved."""
        s = self.spec
        if s.is_moe or s.parallel_residual or (self.tp_size > 1 and not self._tp_fused):
            raise NotImplementedError("wf8af8 / wf6af16 cover dense, sequential-residual models (tensor parallel: with the "
                                      "fused flag-protocol collectives; mixture-of-experts and parallel-residual blocks "
                                      "reduce through fp32 partial sums)")
        saved = 0
        for lw in self.w.layers:
            for name in ("w_qkv", "w_o", "w_up", "w_down"):
                w = getattr(lw, name)
                if isinstance(w, torch.Tensor):
                    q = Fp8Weight.from_tensor(w, fmt if w.shape[1] % 128 == 0 else "e4m3")
                    saved += w.numel() * w.element_size() - q.nbytes()
                    setattr(lw, name, q)
        ks = set()
        for lw in self.w.layers:
            ks |= {int(lw.w_qkv.shape[1]), int(lw.w_o.shape[1]), int(lw.w_up.shape[1]), int(lw.w_down.shape[1])}
        self.ops.reserve_fp8(self._state_config.max_ragged_batch_size, sorted(ks), self.device)
        self.fp8 = True
        return saved

    # ------------------------------------------------------------------------------------------------
    # forward
    # ------------------------------------------------------------------------------------------------
    def _tp_scatter(self, T: int) -> bool:
        """Shape of the collective inside the norm kernels of a T-token forward (see TP_ONESHOT_MAX_TOKENS)."""
        return self._tp_scatter_ok and T > TP_ONESHOT_MAX_TOKENS

    def _tp_rows(self, T: int) -> Tuple[int, int]:
        """(first row, row count) of the residual stream this rank owns under reduce-scatter + all-gather."""
        per = ceil_div(T, self.tp_size)
        r0 = min(T, self.tp_rank * per)
        return r0, min(T, r0 + per) - r0

    def _norm(self, x, gamma, out, beta=None, delta=None, res_out=None) -> torch.Tensor:
        """Norm that folds the pending delta into the residual stream. Under fused tensor parallelism the delta is this
        rank's partial-sum buffer and the kernel performs the collective over NVLink (``hds_norm_tp``): one-shot
        all-reduce for small batches, reduce-scatter + all-gather (``out`` written on every rank, flag signalled, first
        reader must wait) otherwise. fp32 partial sums alternate between two buffers; the idle one is cleared here."""
        s = self.spec
        if delta is not None and self._tp_fused:
            T = delta.shape[0]
            zero_buf = None
            if delta.dtype == torch.float32:
                i = 0 if delta.data_ptr() == self._delta_bufs[0].data_ptr() else 1
                zero_buf = self._delta_bufs[1 - i][:T]
            scatter = None
            if self._tp_scatter(T):
                scatter = self._tp_rows(T)
                self._tp_wait_next = True
            return self.ops.norm(x, gamma, out, beta=beta, delta=delta, res_out=res_out, eps=s.norm_eps,
                                 layer_norm=s.norm == "layer", comm=self.comm, zero_buf=zero_buf, scatter_rows=scatter)
        return self.ops.norm(x, gamma, out, beta=beta, delta=delta, zero_delta=True, res_out=res_out, eps=s.norm_eps,
                             layer_norm=s.norm == "layer")

    def _take_tp_wait(self) -> bool:
        """True once after a reduce-scatter + all-gather norm: the caller passes it to the first reader of that output."""
        w, self._tp_wait_next = self._tp_wait_next, False
        return w

    def _row_parallel(self, x: torch.Tensor, w: torch.Tensor, b: Optional[torch.Tensor], T: int, force_accum: bool,
                      last_into_delta: bool = True) -> torch.Tensor:
        """y = x @ w.T (+b), left as the pending delta: fp32 partial sums when the GEMM is split over K (decode-size
        batches, parallel-residual blocks), else a 16-bit store. Tensor parallel (fused): the buffers are symmetric
        memory, the GEMM's last CTA signals the peers, the next norm kernel reduces over the ranks."""
        tp = self._tp_fused
        fp8 = isinstance(w, Fp8Weight)
        accum = force_accum or T <= (FP8_ACCUM_MAX_TOKENS if fp8 else ACCUM_MAX_TOKENS) or (self.tp_size > 1 and not tp)
        if not accum:
            out16 = self.partial16[:T] if tp else self.delta16[:T]
            return self.ops.gemm(x, w, out16, bias=b, comm=self.comm, tp_signal=tp)
        d32 = self._delta_bufs[self._delta_idx][:T]
        self.ops.gemm_accum(x, w, d32, bias=b, comm=self.comm, tp_signal=tp and last_into_delta)
        if last_into_delta:  # this delta is complete: the next reducing GEMM uses the other buffer
            if self.tp_size > 1 and not tp:
                self.comm.all_reduce_delta(d32)
            self._delta_idx = (self._delta_idx + 1) % len(self._delta_bufs)
        return d32

    def _mlp(self, x: torch.Tensor, lw, T: int, force_accum: bool) -> torch.Tensor:
        s = self.spec
        if s.is_moe:
            return self._moe(x, lw, T)
        n_act = lw.w_up.shape[0] // (2 if s.gated_mlp else 1)
        act = self.act_buf[:T, :n_act]
        self.ops.gemm(x, lw.w_up, act, bias=lw.b_up, act=self.act, gated=s.gated_mlp, workspace=self.ws_gemm,
                      comm=self.comm, tp_wait=self._take_tp_wait())
        return self._row_parallel(act, lw.w_down, lw.b_down, T, force_accum)

    def _moe_state(self, lw, E: int, kk: int) -> dict:
        """Persistent routing buffers (shared by all layers) + this layer's stacked expert weights."""
        if getattr(self, "_moe_bufs", None) is None:
            dev, dt, Tm = self.device, self.dtype, self._state_config.max_ragged_batch_size
            rows = self.ops.k.moe_padded_rows(Tm, E, kk)
            inter = lw.experts[0].w_down.shape[1]
            self._moe_bufs = dict(
                logits=torch.empty(Tm, E, dtype=dt, device=dev), scores=torch.empty(Tm, kk, dtype=torch.float32, device=dev),
                expert_ids=torch.empty(Tm, kk, dtype=torch.int32, device=dev),
                offsets=torch.empty(Tm, kk, dtype=torch.int32, device=dev),
                mapped_slots=torch.empty(Tm, kk, dtype=torch.int32, device=dev),
                counts=torch.zeros(E, dtype=torch.int32, device=dev), pad_off=torch.zeros(E + 1, dtype=torch.int32, device=dev),
                xs=torch.zeros(rows, self.spec.hidden, dtype=dt, device=dev), hs=torch.zeros(rows, inter, dtype=dt, device=dev),
                ys=torch.zeros(rows, self.spec.hidden, dtype=dt, device=dev))
        stacked = self.__dict__.setdefault("_moe_stacked", {})
        if id(lw) not in stacked:
            # stack the experts once; the per-expert tensors become views of the stacked storage
            up = torch.cat([e.w_up for e in lw.experts], dim=0).contiguous()
            down = torch.cat([e.w_down for e in lw.experts], dim=0).contiguous()
            nu, nd = lw.experts[0].w_up.shape[0], lw.experts[0].w_down.shape[0]
            for i, e in enumerate(lw.experts):
                e.w_up, e.w_down = up[i * nu:(i + 1) * nu], down[i * nd:(i + 1) * nd]
            stacked[id(lw)] = (up, down)
        return dict(self._moe_bufs, w_up=stacked[id(lw)][0], w_down=stacked[id(lw)][1])

    def _moe(self, x: torch.Tensor, lw, T: int) -> torch.Tensor:
        """Top-k mixture of experts (tensor-parallel experts, no expert parallelism — as the reference's v2:
        ``inference_transformer_base.py:588``). On the GPU: gate GEMM -> top-k gating kernel -> scatter -> ONE grouped
        tcgen05 GEMM for all experts' MLP-1 (gated activation in the epilogue) -> ONE grouped GEMM for MLP-2 -> weighted
        gather straight into the fp32 pending delta; nothing is read back to the host.
        Ref: modules/implementations/moe/cutlass_multi_gemm.py:28-249, ragged_ops/{top_k_gating,moe_scatter,moe_gather}."""
        s = self.spec
        d32 = self._delta_bufs[self._delta_idx][:T]
        mult = 2 if s.gated_mlp else 1
        if self.device.type == "cuda":
            k = self.ops.k
            E, kk = len(lw.experts), s.top_k
            st = self._moe_state(lw, E, kk)
            logits = k.gemm(x, lw.w_router, out=st["logits"][:T])
            k.moe_topk_gating(logits, kk, s.norm_topk_prob, st["scores"][:T], st["expert_ids"][:T], st["offsets"][:T],
                              st["counts"])
            rows = k.moe_padded_rows(T, E, kk)
            xs, hs, ys = st["xs"][:rows], st["hs"][:rows], st["ys"][:rows]
            k.moe_scatter(x, xs, st["expert_ids"][:T], st["offsets"][:T], st["counts"], st["pad_off"],
                          st["mapped_slots"][:T], kk)
            k.moe_gemm(xs, st["w_up"], st["pad_off"], E, hs, act=self.act, gated=s.gated_mlp)
            k.moe_gemm(hs, st["w_down"], st["pad_off"], E, ys)
            k.moe_gather(ys, st["mapped_slots"][:T], st["scores"][:T], d32)
        else:
            scores = ref.gemm(x, lw.w_router, out_dtype=torch.float32)
            probs = torch.softmax(scores, dim=-1)
            topv, topi = torch.topk(probs, s.top_k, dim=-1)
            if s.norm_topk_prob:
                topv = topv / topv.sum(-1, keepdim=True)
            for e, ew in enumerate(lw.experts):
                sel = (topi == e)
                rows = sel.any(-1).nonzero(as_tuple=False).flatten()
                if rows.numel() == 0:
                    continue
                gate_w = (topv * sel).sum(-1)[rows]
                xe = x.index_select(0, rows)
                he = torch.empty((rows.numel(), ew.w_up.shape[0] // mult), dtype=x.dtype, device=x.device)
                self.ops.gemm(xe, ew.w_up, he, act=self.act, gated=s.gated_mlp)
                ye = torch.empty((rows.numel(), s.hidden), dtype=x.dtype, device=x.device)
                self.ops.gemm(he, ew.w_down, ye)
                d32.index_add_(0, rows, ye.float() * gate_w[:, None])
        if lw.shared_expert is not None:
            se = lw.shared_expert
            hs = torch.empty((T, se.w_up.shape[0] // mult), dtype=x.dtype, device=x.device)
            self.ops.gemm(x, se.w_up, hs, act=self.act, gated=s.gated_mlp)
            ys = torch.empty((T, s.hidden), dtype=x.dtype, device=x.device)
            self.ops.gemm(hs, se.w_down, ys)
            g = torch.sigmoid(ref.gemm(x, lw.w_shared_gate, out_dtype=torch.float32))
            d32 += ys.float() * g
        if self.tp_size > 1:
            if self._tp_fused:
                self.comm.signal()               # partial sums complete on this rank: the next norm kernel reduces them
            else:
                self.comm.all_reduce_delta(d32)  # NCCL all-reduce
        self._delta_idx = (self._delta_idx + 1) % len(self._delta_bufs)
        return d32

    def forward(self, batch: RaggedBatchWrapper, capture: bool = False, all_logits: bool = False,
                n_layers_limit: Optional[int] = None,
                restore_interleave: float = 0.0) -> Tuple[torch.Tensor, Optional[torch.Tensor]]:
        """One ragged forward. Returns ``(logits [n_seqs, vocab], latents)`` where latents is the device view
        ``[n_layers, n_tokens, hidden]`` of the captured QKV inputs (None when capture is off).
        ``all_logits``: unembed every token instead of each sequence's last one (``[n_tokens, vocab]``; v1 API).
        ``n_layers_limit``: run only the first layers (writing their K/V) and return ``(None, latents)`` — the
        recomputed prefix of a mixed HCache restore plan. With a deferred restore pending, ``restore_interleave``
        saved layers are projected after each recomputed layer (their transfers complete at that rate), so the
        copy engine never waits for a free staging slot while the GPU recomputes."""
        s, w, ops = self.spec, self.w, self.ops
        T, S = batch.current_tokens, batch.current_sequences
        if capture and self.latent_buf is None:
            raise RuntimeError("hidden-state capture was disabled when the model was built")
        pos, slot = batch.tok_pos(), batch.tok_slot()
        resid = self.residual[:T]
        if w.embed_norm_g is not None:   # Bloom: the residual stream starts from the normalised embeddings
            raw = self.embed_tmp[:T]
            ops.embed(batch.input_ids(), w.embed, raw, pos_weight=w.pos_embed, tok_pos=pos, pos_offset=s.learned_pos_offset)
            ops.norm(raw, w.embed_norm_g, resid, beta=w.embed_norm_b, eps=s.norm_eps, layer_norm=True)
        else:
            ops.embed(batch.input_ids(), w.embed, resid, pos_weight=w.pos_embed, tok_pos=pos, pos_offset=s.learned_pos_offset)
        ln = s.norm == "layer"
        pending = None
        kv = self._state_manager
        bs = kv.kv_block_size
        two_norms = (not s.parallel_residual) or s.parallel_two_norms
        n_run = s.n_layers if n_layers_limit is None else min(n_layers_limit, s.n_layers)
        for li in range(n_run):
            lw = w.layers[li]
            xn = (self.latent_buf[li] if capture else self.xn)[:T]
            self._norm(resid, lw.norm1_g, xn, beta=lw.norm1_b, delta=pending,
                       res_out=resid if pending is not None else None)
            kv_layer = kv.get_cache(li)
            if self._pending_kv is not None:
                self._kv_step(li)       # deferred reload of this layer's offloaded KV pages
            if self._pending_restore is not None:
                self._restore_step(li)  # deferred HCache restore: this layer's K/V must land before its attention
            ops.qkv_rope_kv_append(xn, lw.w_qkv, lw.b_qkv, self.qkv_buf[:T], self.q_buf[:T], kv_layer, pos, slot,
                                   self.inv_freq, self.n_q, self.n_kv, self.d, s.rot_dim, bs, workspace=self.ws_gemm,
                                   comm=self.comm, tp_wait=self._take_tp_wait())
            attn = self.attn_buf[:T]
            ops.attention(self.q_buf[:T], attn, kv_layer, batch, self.n_q, self.n_kv, self.d, self.scale,
                          ws_small=self.ws_small, ws_large=self.ws_large, alibi=self.alibi, window=s.window_of(li))
            if s.parallel_residual:
                xm = xn
                if two_norms:
                    xm = self._norm(resid, lw.norm2_g, self.xn2[:T], beta=lw.norm2_b)
                self._row_parallel(attn, lw.w_o, lw.b_o, T, force_accum=True, last_into_delta=False)
                pending = self._mlp(xm, lw, T, force_accum=True)
            else:
                d1 = self._row_parallel(attn, lw.w_o, lw.b_o, T, force_accum=False)
                x2 = self._norm(resid, lw.norm2_g, self.xn2[:T], beta=lw.norm2_b, delta=d1, res_out=resid)
                pending = self._mlp(x2, lw, T, force_accum=False)
            if n_layers_limit is not None and restore_interleave > 0 and self._pending_restore is not None:
                st = self._pending_restore
                target = min(len(st["layers"]), int((li + 1) * restore_interleave))
                while st["done"] < target:
                    self._restore_project(st, st["done"])
                if st["done"] >= len(st["layers"]):
                    self._pending_restore = None
        if n_layers_limit is not None:
            if self._take_tp_wait():
                self.comm.wait()
            if pending is not None and pending.dtype == torch.float32:
                # the partial sums of the last executed layer are not consumed by a norm; under pull-mode TP the
                # previously consumed buffer is still dirty as well (every rank is past the barrier that follows
                # the last GEMM, so nobody reads either buffer any more)
                for b in self._delta_bufs:
                    b[:T].zero_()
            return None, (self.latent_buf[:n_run, :T] if capture else None)
        if self._pending_restore is not None or self._pending_kv is not None:
            self.finish_pending_restore()
        # ---- last-token logits ----
        fin = self.xn2[:T]
        if s.final_norm:
            self._norm(resid, w.final_norm_g, fin, beta=w.final_norm_b, delta=pending)
        else:
            # no final norm: only the residual add (and, tensor parallel, the reduction) is wanted -> res_out; `fin` must hold
            # every row on every rank, so this one always takes the all-reduce shape
            ones = torch.ones(s.hidden, dtype=self.dtype, device=self.device)
            if pending is not None and self._tp_fused:
                zb = None
                if pending.dtype == torch.float32:
                    i = 0 if pending.data_ptr() == self._delta_bufs[0].data_ptr() else 1
                    zb = self._delta_bufs[1 - i][:T]
                ops.norm(resid, ones, self.xn[:T], delta=pending, res_out=fin, eps=s.norm_eps, comm=self.comm, zero_buf=zb)
            else:
                ops.norm(resid, ones, self.xn[:T], delta=pending, zero_delta=True, res_out=fin, eps=s.norm_eps)
        if self._take_tp_wait():
            self.comm.wait()   # the peers' rows of `fin` have landed
        if all_logits:
            out_local = torch.empty((T, self.info.vocab), dtype=self.dtype, device=self.device)
            logits = ops.gemm(fin, w.lm_head, out_local, bias=w.lm_head_b, workspace=self.ws_gemm)
            n_rows = T
        else:
            hidden = ops.gather_rows(fin, batch.last_rows(), self.hidden_last[:S])
            logits = ops.gemm(hidden, w.lm_head, self.logits_local[:S], bias=w.lm_head_b, workspace=self.ws_gemm,
                              comm=self.comm, tp_signal=self._tp_fused)
            n_rows = S
        if self.tp_size > 1:
            ranges = [shard_info(s, r, self.tp_size).vocab_range for r in range(self.tp_size)]
            if self._tp_fused and not all_logits:
                # vocabulary shards land straight in the final [S, vocab] layout: one pull kernel behind the GEMMs' flags
                logits = self.comm.gather_logits(logits, self.logits_full[:S], ranges)
            else:
                full = self.logits_full[:S] if not all_logits else torch.empty((n_rows, s.vocab), dtype=self.dtype,
                                                                               device=self.device)
                logits = self.comm.all_gather_logits(logits, full, ranges)
        return logits, (self.latent_buf[:, :T] if capture else None)

    # ------------------------------------------------------------------------------------------------
    # HCache restore
    # ------------------------------------------------------------------------------------------------
    def kv_weight(self, li: int) -> Tuple[torch.Tensor, Optional[torch.Tensor]]:
        """[W_k; W_v] (and bias) of a layer: row-slice views of the fused QKV parameters (no copy)."""
        lw = self.w.layers[li]
        r0 = self.n_q * self.d
        return lw.w_qkv[r0:], (lw.b_qkv[r0:] if lw.b_qkv is not None else None)

    def restore_kv(self, batch: RaggedBatchWrapper, latents: Sequence[torch.Tensor],
                   layers: Optional[Sequence[int]] = None, defer: bool = False, sharded: bool = False,
                   layer_ready=None) -> None:
        """Rebuild the paged K/V of the batch's tokens from saved hidden states.

        ``latents``: one host tensor per sequence of the batch, ``[n_saved_layers, n_tokens_i, hidden]``, whose
        leading axis follows ``layers`` (default: every layer). Pinned tensors stream asynchronously; pageable
        ones go through the driver's staging path. Returns without waiting for the projections; kernels of the
        next ``put`` queue behind them on the same stream.

        ``defer=True`` only starts the host-to-device stream; the projection of layer ``l`` is then issued by the
        next ``forward`` immediately before its own layer ``l`` — the new turn's prefill fills the gaps in which
        the compute stream would otherwise wait for PCIe (restore is IO-bound on B200), so time-to-first-token
        approaches the pure transfer time.

        ``layer_ready(i)``: optional host callback invoked right before saved layer ``i`` (index into ``layers``) is
        copied to the device; it blocks until that layer's bytes are in ``latents`` — the hook that pipelines NVMe reads
        (issued up front, completed layer by layer) against the host-to-device stream and the projections.
        """
        s, ops = self.spec, self.ops
        layers = list(range(s.n_layers)) if layers is None else list(layers)
        T = batch.current_tokens
        if sharded:
            if self.restore_stage_sym is None or len(latents) != 1 or batch.current_sequences != 1:
                raise ValueError("sharded restore needs fused tensor parallelism and exactly one sequence per call")
            lo, hi, _ = shard_rows(T, self.tp_size, self.tp_rank)
            if int(latents[0].shape[1]) != hi - lo:
                raise ValueError(f"rank {self.tp_rank} must pass rows [{lo}, {hi}) of the hidden states")
        elif sum(int(l.shape[1]) for l in latents) != T:
            raise ValueError("latents do not cover the tokens of the restore batch")
        pos, slot = batch.tok_pos(), batch.tok_slot()
        kv = self._state_manager
        bs = kv.kv_block_size
        if self.device.type != "cuda":
            for i, li in enumerate(layers):
                if layer_ready is not None:
                    layer_ready(i)
                lat = torch.cat([l[i] for l in latents], dim=0).to(self.dtype)
                wk, bk = self.kv_weight(li)
                ops.restore_kv_proj(lat, wk, bk, kv.get_cache(li), pos, slot, self.inv_freq, self.n_kv, self.d, s.rot_dim, bs)
            return
        self.finish_pending_restore()
        comp = torch.cuda.current_stream()
        io = self.io_stream
        io.wait_stream(comp)  # staging slots may still feed kernels of an earlier restore
        # private copy of the per-token metadata: the batch wrapper's buffers are rewritten by the next forward
        self.restore_pos[:T].copy_(pos, non_blocking=True)
        self.restore_slot[:T].copy_(slot, non_blocking=True)
        st = {"T": T, "layers": layers, "latents": list(latents), "issued": 0, "done": 0, "by_layer": {},
              "sharded": bool(sharded), "layer_ready": layer_ready}
        for i, li in enumerate(layers):
            st["by_layer"].setdefault(li, []).append(i)
        self._pending_restore = st
        for _ in range(min(RESTORE_SLOTS, len(layers))):
            self._restore_issue_h2d(st)
        if not defer:
            self.finish_pending_restore()

    def _restore_issue_h2d(self, st) -> None:
        """Queue the H2D of the next saved layer into its staging slot (IO stream)."""
        i = st["issued"]
        k = i % RESTORE_SLOTS
        io = self.io_stream
        if st.get("layer_ready") is not None:
            st["layer_ready"](i)      # e.g. wait for the NVMe read of this layer (later layers keep reading meanwhile)
        stage = self.restore_stage_sym[k] if st["sharded"] else self.restore_stage[k]
        with torch.cuda.stream(io):
            if i >= RESTORE_SLOTS:
                io.wait_event(self.ev_consumed[k])
            off = 0
            if st["sharded"] and self.tp_gather == "copy":
                off = shard_rows(st["T"], self.tp_size, self.tp_rank)[0]  # my rows sit at their global position
            for lat in st["latents"]:
                n = int(lat.shape[1])
                stage[off:off + n].copy_(lat[i], non_blocking=True)
                off += n
            self.ev_loaded[k].record(io)
        st["issued"] = i + 1

    def _restore_project(self, st, i: int) -> None:
        """Project saved layer index ``i`` (compute stream) and keep the IO stream RESTORE_SLOTS layers ahead."""
        s, kv = self.spec, self._state_manager
        li = st["layers"][i]
        k = i % RESTORE_SLOTS
        T = st["T"]
        comp = torch.cuda.current_stream()
        comp.wait_event(self.ev_loaded[k])
        wk, bk = self.kv_weight(li)
        if st["sharded"]:
            # My slice of this layer has landed in my HBM (the stream waited for the copy): tell the peers, then gather
            # every peer's rows over NVLink behind their flags. Slot reuse needs no second exchange: a peer issues the
            # "landed" signal of layer i+1 only after its gather of layer i, so once MY gather of layer i+1 has passed its
            # wait every rank is done reading slot k(i) — the copy into that slot waits for an event recorded there.
            _, _, per = shard_rows(T, self.tp_size, self.tp_rank)
            comm = self.comm
            comm.signal()
            if self.tp_gather == "copy":
                # pull every peer's rows once over NVLink into my full-size slot, then project locally
                slot_t = self.restore_stage_sym[k]
                if k not in self._slot_peer_ptrs:
                    self._slot_peer_ptrs[k] = comm.peer_ptrs_raw(slot_t)
                comm.arm(wait=True)
                self.ops.k.p2p_gather_rows(slot_t, self._slot_peer_ptrs[k], per, T, self.tp_rank)
                if i >= 1:
                    self.ev_consumed[(i - 1) % RESTORE_SLOTS].record(comp)
                self.ops.restore_kv_proj(slot_t[:T], wk, bk, kv.get_cache(li), self.restore_pos[:T], self.restore_slot[:T],
                                         self.inv_freq, self.n_kv, self.d, s.rot_dim, kv.kv_block_size,
                                         scratch=self.kv_scratch[:T] if self.kv_scratch is not None else None)
            else:
                comm.wait()
                if i >= 1:
                    self.ev_consumed[(i - 1) % RESTORE_SLOTS].record(comp)
                ptrs = comm.peer_ptrs_raw(self.restore_stage_sym, byte_offset=k * self._sym_rows * s.hidden * 2)
                self.ops.k.restore_kv_proj_peers(ptrs, per, T, s.hidden, self.dtype, wk, kv.get_cache(li),
                                                 self.restore_pos[:T], self.restore_slot[:T], self.inv_freq, self.n_kv,
                                                 self.d, kv.kv_block_size, bias_kv=bk)
            st["done"] = i + 1
            if st["done"] >= len(st["layers"]):
                comm.barrier()   # every rank has finished reading every slot before the next restore refills them
            elif i >= 1 and st["issued"] < len(st["layers"]):
                self._restore_issue_h2d(st)
            return
        self.ops.restore_kv_proj(self.restore_stage[k][:T], wk, bk, kv.get_cache(li), self.restore_pos[:T],
                                 self.restore_slot[:T], self.inv_freq, self.n_kv, self.d, s.rot_dim, kv.kv_block_size,
                                 scratch=self.kv_scratch[:T] if self.kv_scratch is not None else None)
        self.ev_consumed[k].record(comp)
        st["done"] = i + 1
        if st["issued"] < len(st["layers"]):
            self._restore_issue_h2d(st)

    def _restore_step(self, li: int) -> None:
        """Called by ``forward`` right before layer ``li`` touches its KV cache."""
        st = self._pending_restore
        if li not in st["by_layer"]:
            return
        for i in st["by_layer"].get(li, ()):
            while st["done"] <= i:  # saved layers are consumed in order
                self._restore_project(st, st["done"])
        if st["done"] >= len(st["layers"]):
            self._pending_restore = None

    def finish_pending_restore(self) -> None:
        """Run whatever is left of a deferred restore now (compute stream)."""
        kvst = self._pending_kv
        if kvst is not None:
            while kvst["done"] < len(kvst["layers"]):
                self._kv_apply(kvst, kvst["done"])
            self._pending_kv = None
        st = self._pending_restore
        if st is None:
            return
        while st["done"] < len(st["layers"]):
            self._restore_project(st, st["done"])
        self._pending_restore = None

    # ---- layer-pipelined reload of offloaded KV pages ----------------------------------------------------------
    def _kv_issue(self, st) -> None:
        """Queue the H2D of the next offloaded layer into its staging slot (IO stream)."""
        j = st["issued"]
        k = j % RESTORE_SLOTS
        io = self.io_stream
        with torch.cuda.stream(io):
            if j >= RESTORE_SLOTS:
                io.wait_event(self._kv_ev_consumed[k])
            self._kv_stage[k][:st["n_blocks"]].copy_(st["kv"][j], non_blocking=True)
            self._kv_ev_loaded[k].record(io)
        st["issued"] = j + 1

    def _kv_apply(self, st, j: int) -> None:
        """Scatter staged layer ``j`` into its KV blocks (compute stream) and keep the IO stream ahead."""
        k = j % RESTORE_SLOTS
        li = st["layers"][j]
        comp = torch.cuda.current_stream()
        comp.wait_event(self._kv_ev_loaded[k])
        cache = self._state_manager.kv_cache.full_cache
        self.ops.k.kv_blocks_copy(cache[li:li + 1], self._kv_stage[k][:st["n_blocks"]].unsqueeze(1), st["ids"], to_staging=False)
        self._kv_ev_consumed[k].record(comp)
        st["done"] = j + 1
        if st["issued"] < len(st["layers"]):
            self._kv_issue(st)

    def _kv_step(self, li: int) -> None:
        """Called by ``forward`` right before layer ``li`` reads its KV cache."""
        st = self._pending_kv
        j = st["by_layer"].get(li)
        if j is not None:
            while st["done"] <= j:
                self._kv_apply(st, st["done"])
        if st["done"] >= len(st["layers"]):
            self._pending_kv = None

    def restore_kv_pages(self, block_ids: torch.Tensor, kv_host: torch.Tensor, layers: Sequence[int], defer: bool = False,
                         layer_major: bool = False) -> None:
        """KV reload for the layers the planner assigned to KV offload, written into ``block_ids`` of each listed layer.
        ``kv_host`` is ``[n_blocks, len(layers), 2, kvh, bs, d]`` (one bulk transfer), or with ``layer_major``
        ``[len(layers), n_blocks, 2, kvh, bs, d]``: every layer is then one contiguous transfer, streamed through
        ``RESTORE_SLOTS`` staging slots; with ``defer`` layer ``l`` is scattered by the next ``forward`` right before its
        own layer ``l``, so the new turn's prefill overlaps the transfers exactly as it does for hidden states."""
        kv = self._state_manager.kv_cache
        cache = kv.full_cache
        ids = block_ids.to(torch.int32)
        if self.device.type != "cuda":
            for j, li in enumerate(layers):
                cache[li, ids.long()] = (kv_host[j] if layer_major else kv_host[:, j]).to(cache.dtype)
            return
        if layer_major:
            if self._pending_kv is not None:        # (a pending HIDDEN-state restore of the same plan stays deferred)
                prev = self._pending_kv
                while prev["done"] < len(prev["layers"]):
                    self._kv_apply(prev, prev["done"])
                self._pending_kv = None
            n_blocks = int(kv_host.shape[1])
            shape = tuple(kv_host.shape[2:])
            if self._kv_stage is None or self._kv_stage.shape[1] < n_blocks or tuple(self._kv_stage.shape[2:]) != shape:
                self._kv_stage = torch.empty((RESTORE_SLOTS, n_blocks) + shape, dtype=cache.dtype, device=self.device)
                self._kv_ev_loaded = [torch.cuda.Event() for _ in range(RESTORE_SLOTS)]
                self._kv_ev_consumed = [torch.cuda.Event() for _ in range(RESTORE_SLOTS)]
            self.io_stream.wait_stream(torch.cuda.current_stream())   # slots may still feed an earlier restore's copies
            st = {"layers": list(layers), "kv": kv_host, "ids": ids.to(self.device, non_blocking=True), "n_blocks": n_blocks,
                  "issued": 0, "done": 0, "by_layer": {li: j for j, li in enumerate(layers)}}
            self._pending_kv = st
            for _ in range(min(RESTORE_SLOTS, len(st["layers"]))):
                self._kv_issue(st)
            if not defer:
                self.finish_pending_restore()
            return
        ids_dev = ids.to(self.device, non_blocking=True)
        comp = torch.cuda.current_stream()
        io = self.io_stream
        io.wait_stream(comp)
        with torch.cuda.stream(io):
            staged = kv_host.to(self.device, non_blocking=True)
            done = torch.cuda.Event()
            done.record(io)
        comp.wait_event(done)
        staged.record_stream(comp)
        for j, li in enumerate(layers):
            self.ops.k.kv_blocks_copy(cache[li:li + 1], staged[:, j:j + 1].contiguous() if len(layers) > 1 else staged,
                                      ids_dev, to_staging=False)

    # ------------------------------------------------------------------------------------------------
    def n_params_local(self) -> int:
        return sum(t.numel() for t in self.w.named_tensors().values())
