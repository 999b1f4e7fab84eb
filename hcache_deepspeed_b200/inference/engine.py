"""``init_inference`` engine (the v1 API).

Counterpart of ``deepspeed/inference/engine.py:40-625``. Modes:

* **kernel injection** (``replace_with_kernel_inject=True`` or ``hcache.enabled``): the module's weights are
  re-laid-out once and the transformer stack runs on this framework's executor (sm_100a kernels on a GPU, the
  fp32 restatement on CPU). In the reference, injection swaps each HF layer for a fused ``DeepSpeedTransformerInference``
  module (``module_inject/replace_module.py:183-265``); here the whole stack is swapped, which also gives the v1
  API paged KV, ragged batches and HCache.
* **no injection**: the module runs unchanged (dtype conversion, device placement, optional CUDA-graph replay),
  as in the reference.

New relative to the reference: sessions (``session_ids=``) and hidden-state save / restore on the v1 API
(``hidden_states`` / ``restore_hidden_states``) — BASELINE.json config 1.
"""
from __future__ import annotations

import os
import time
from typing import Dict, List, Optional, Sequence

import torch

from ..utils.logging import log_dist, logger
from .config import DeepSpeedInferenceConfig
from .v2.config_v2 import RaggedInferenceEngineConfig

INFERENCE_MODEL_TIMER = "model-forward-inference"


class _Output(dict):
    """Minimal ``CausalLMOutput``-like result (attribute and key access, tuple-style indexing)."""

    def __getattr__(self, k):
        try:
            return self[k]
        except KeyError as e:
            raise AttributeError(k) from e

    def __getitem__(self, k):
        if isinstance(k, int):
            return list(self.values())[k]
        return super().__getitem__(k)


class InferenceEngine(torch.nn.Module):
    inference_mp_group = None
    inference_ep_group = None
    expert_mp_group = None

    def __init__(self, model: torch.nn.Module, config: DeepSpeedInferenceConfig):
        super().__init__()
        self.module = model
        self._config = config
        self.dtype = config.dtype
        self.mp_world_size = config.tensor_parallel.tp_size
        self.mpu = config.tensor_parallel.mpu
        self.model_profile_enabled = False
        self._model_times: List[float] = []
        self._is_compiled = False
        self._cuda_graph = None
        self._graph_inputs = None
        self._graph_output = None
        self._sessions_latents: Dict[int, List[torch.Tensor]] = {}
        self._sessions_tokens: Dict[int, List[torch.Tensor]] = {}
        self._next_ephemeral_uid = 1 << 40
        self._engine = None
        if config.enable_cuda_graph and config.tensor_parallel.tp_size > 1:
            raise AssertionError("CUDA graph is not supported for model-parallel inference on the v1 API")
        on_gpu = torch.cuda.is_available() and not config.keep_module_on_host
        self.device = torch.device("cuda", int(os.getenv("LOCAL_RANK", 0)) % max(1, torch.cuda.device_count())) if on_gpu \
            else torch.device("cpu")
        if config.tensor_parallel.tp_size > 1:
            self._create_model_parallel_group(config)
        if config.checkpoint is not None:
            self._load_checkpoint(config.checkpoint, base_dir=config.base_dir)
        self.injected = bool(config.replace_with_kernel_inject or config.hcache.enabled)
        self.fused_layers = 0
        if self.injected and not config.hcache.enabled and not self._has_v2_policy():
            # families outside the ragged executor: per-layer injection of the v1 fused layer
            self.injected = False
            self._apply_v1_injection(config)
        elif self.injected:
            self._apply_injection()
        else:
            if config.tensor_parallel.tp_size > 1:
                from ..module_inject.auto_tp import AutoTP
                AutoTP.apply(self.module, self.mp_world_size, torch.distributed.get_rank(self.inference_mp_group),
                             self.inference_mp_group, injection_policy=config.injection_policy)
            self._convert_to_dtype(config)
            self.module.to(self.device)
        if config.save_mp_checkpoint_path:
            from ..module_inject.load_checkpoint import save_mp_checkpoint
            tp = config.tensor_parallel.tp_size
            rank = torch.distributed.get_rank(self.inference_mp_group) if tp > 1 else 0
            save_mp_checkpoint(self.module, config.save_mp_checkpoint_path, tp, rank,
                               model_type=getattr(getattr(self.module, "config", None), "model_type", "ds_model"),
                               dtype=config.dtype)
        if hasattr(self.module, "generate"):
            self._orig_generate = self.module.generate
        if config.tensor_parallel.tp_size > 1 and on_gpu:
            # every TP rank must draw the same random numbers when sampling (ref: engine.py:174-177)
            state = torch.cuda.get_rng_state().to(self.device)
            torch.distributed.broadcast(state, 0, group=self.inference_mp_group)
            torch.cuda.set_rng_state(state.cpu())

    # ---- construction -------------------------------------------------------------------------------
    def _create_model_parallel_group(self, config) -> None:
        from ..comm import comm as dist
        if InferenceEngine.inference_mp_group is None:
            dist.init_distributed()
            if config.tensor_parallel.tp_group is not None:
                InferenceEngine.inference_mp_group = config.tensor_parallel.tp_group
            elif self.mpu is not None:
                InferenceEngine.inference_mp_group = self.mpu.get_model_parallel_group()
            else:
                InferenceEngine.inference_mp_group = dist.new_group(list(range(config.tensor_parallel.tp_size)))
        self.mp_group = InferenceEngine.inference_mp_group

    def _convert_to_dtype(self, config) -> None:
        if config.dtype in (torch.half, torch.bfloat16, torch.float32):
            self.module.to(config.dtype)

    def _load_checkpoint(self, load_dir, base_dir: str = "", load_module_strict: bool = False, tag=None) -> None:
        """``init_inference(model, checkpoint=...)``: a checkpoint json (path or dict), a folder of weight files, or a
        DeepSpeed / Megatron ``mp_rank_*`` folder (merged or split to this TP degree); meta-device modules are
        materialised by the load (ref ``inference/engine.py:411-468``)."""
        from ..module_inject.load_checkpoint import load_checkpoint_into
        tp = self._config.tensor_parallel.tp_size
        rank = torch.distributed.get_rank(self.inference_mp_group) if tp > 1 and self.inference_mp_group is not None else 0
        n = load_checkpoint_into(self.module, load_dir, base_dir or "", tp_size=tp, tp_rank=rank, strict=load_module_strict, tag=tag)
        log_dist(f"init_inference: loaded {n} checkpoint file(s)", [0])

    def _has_v2_policy(self) -> bool:
        """True when the module's architecture is one the ragged executor runs. Variants a family's spec builder rejects
        (ALiBi Falcon, GPT-2 with per-layer attention scaling, ...) stay on the per-block container path."""
        from ..models.arch import arch_from_hf_config
        from .v2.model_implementations import MODEL_TYPE_TO_POLICY
        cfg = getattr(self.module, "config", None)
        if getattr(cfg, "model_type", None) not in MODEL_TYPE_TO_POLICY:
            return False
        try:
            arch_from_hf_config(cfg)
        except (ValueError, TypeError) as e:
            log_dist(f"init_inference: {cfg.model_type} variant is outside the ragged executor ({e}); using block injection", [0])
            return False
        return True

    def _apply_v1_injection(self, config) -> None:
        """``replace_with_kernel_inject`` for a family without a ragged-executor policy: swap every block for the fused
        layer through its container policy (``module_inject/containers``; ref ``replace_module.py:210-265``)."""
        from ..module_inject.containers import inject_fused_layers
        self._convert_to_dtype(config)
        tp = config.tensor_parallel.tp_size
        rank = torch.distributed.get_rank(self.inference_mp_group) if tp > 1 else 0
        self.fused_layers = inject_fused_layers(self.module, dtype=config.dtype, mp_size=tp, mp_rank=rank,
                                                mp_group=self.inference_mp_group if tp > 1 else None,
                                                max_out_tokens=getattr(config, "max_out_tokens", 1024))
        if self.fused_layers == 0:
            mt = getattr(getattr(self.module, "config", None), "model_type", None)
            raise ValueError(f"kernel injection is not available for model type {mt}: no ragged-executor policy and no "
                             f"container policy matches its blocks")
        self.module.to(self.device)
        log_dist(f"init_inference: {self.fused_layers} transformer blocks replaced by the fused inference layer", [0])

    def _apply_injection(self) -> None:
        from .v2.checkpoint import InMemoryModelEngine
        from .v2.engine_v2 import InferenceEngineV2
        from .v2.model_implementations import MODEL_TYPE_TO_POLICY
        cfg, hc = self._config, self._config.hcache
        mt = getattr(getattr(self.module, "config", None), "model_type", None)
        if mt not in MODEL_TYPE_TO_POLICY:
            raise ValueError(f"kernel injection is not available for model type {mt}; supported: {sorted(MODEL_TYPE_TO_POLICY)}")
        dtype = cfg.dtype
        quant = {}
        if dtype == torch.int8:
            # the reference's 8-bit serving (dtype=torch.int8: int8 weights + the int8 GEMMs of its v1 op). Blackwell's
            # 8-bit tensor-core path is FP8: weights are quantised to E4M3 per output channel, activations per token
            # (the executor's "wf8af8" mode); everything else runs in bf16 (fp32 on the CPU back end)
            quant = {"quantization_mode": "wf8af8"}
            dtype = torch.bfloat16 if self.device.type == "cuda" else torch.float32
            log_dist("init_inference: dtype=torch.int8 -> 8-bit serving through the E4M3 tensor-core path (wf8af8)", [0])
        if self.device.type == "cuda" and dtype not in (torch.float16, torch.bfloat16):
            raise ValueError("the sm_100a kernels take fp16 or bf16; pass dtype=torch.bfloat16")
        if cfg.use_triton:
            log_dist("init_inference: use_triton is ignored (no Triton kernels in this build; the native sm_100a kernels run)", [0])
        state = {"max_tracked_sequences": max(hc.max_sessions, 8), "max_ragged_batch_size": hc.max_batch_tokens,
                 "max_ragged_sequence_count": min(max(hc.max_sessions, 8), hc.max_batch_tokens, 512),
                 "max_context": hc.max_context, "memory_config": {"mode": "allocate", "size": hc.kv_memory_bytes}}
        ecfg = RaggedInferenceEngineConfig(tensor_parallel={"tp_size": cfg.tensor_parallel.tp_size}, state_manager=state,
                                           hcache={"enabled": True, "capture": "sync" if hc.enabled else "off"},
                                           **({"quantization": quant} if quant else {}))
        policy = MODEL_TYPE_TO_POLICY[mt](self.module.config, checkpoint_engine=InMemoryModelEngine(self.module), dtype=dtype,
                                          device=self.device)
        self._engine = InferenceEngineV2(policy, ecfg)
        log_dist(f"init_inference: transformer stack of {mt} now runs on the {self._engine.model().ops.name} executor", [0])

    # ---- profiling (ref: engine.py:193-245, 523-537) ---------------------------------------------------
    def profile_model_time(self, use_cuda_events: bool = True) -> None:
        self.model_profile_enabled = True
        self.use_cuda_events = use_cuda_events and self.device.type == "cuda"

    def model_times(self) -> List[float]:
        assert self.model_profile_enabled, "model profiling is not enabled"
        out, self._model_times = self._model_times, []
        return out

    # ---- sessions / HCache on the v1 API -------------------------------------------------------------
    @property
    def ragged_engine(self):
        """The underlying :class:`InferenceEngineV2` (None without kernel injection)."""
        return self._engine

    def hidden_states(self, session_id: int) -> torch.Tensor:
        """Saved hidden states of a session, ``[n_layers, n_tokens_so_far, hidden]`` (host tensor)."""
        parts = self._sessions_latents.get(session_id)
        if not parts:
            raise KeyError(f"no hidden states recorded for session {session_id} (hcache.enabled must be set)")
        return torch.cat(parts, dim=1) if len(parts) > 1 else parts[0]

    def session_tokens(self, session_id: int) -> torch.Tensor:
        return torch.cat(self._sessions_tokens[session_id])

    def restore_hidden_states(self, session_id: int, tokens: torch.Tensor, latents: torch.Tensor) -> None:
        """Rebuild a flushed session's KV cache from its hidden states instead of recomputing its tokens."""
        self._require_injected()
        self._engine.restore_kv([session_id], [tokens.reshape(-1)], [latents])
        self._sessions_latents[session_id] = [latents]
        self._sessions_tokens[session_id] = [tokens.reshape(-1).cpu()]

    def flush(self, session_id: int, drop_hidden_states: bool = False) -> None:
        """Drop a session's KV cache (its saved hidden states survive unless ``drop_hidden_states``)."""
        self._require_injected()
        self._engine.flush(session_id)
        if drop_hidden_states:
            self._sessions_latents.pop(session_id, None)
            self._sessions_tokens.pop(session_id, None)

    def _require_injected(self) -> None:
        if self._engine is None:
            raise RuntimeError("this call needs init_inference(..., replace_with_kernel_inject=True) or hcache.enabled")

    # ---- forward ---------------------------------------------------------------------------------------
    def _injected_forward(self, input_ids: torch.Tensor, session_ids: Optional[Sequence[int]] = None,
                          last_token_only: bool = False, attention_mask: Optional[torch.Tensor] = None, **_):
        if input_ids.dim() == 1:
            input_ids = input_ids[None]
        Bsz, T = input_ids.shape
        ephemeral = session_ids is None
        if ephemeral:
            session_ids = [self._next_ephemeral_uid + i for i in range(Bsz)]
            self._next_ephemeral_uid += Bsz
        rows = []
        for b in range(Bsz):
            row = input_ids[b]
            if attention_mask is not None:  # left/right padding is dropped: the ragged batch needs no pad tokens
                row = row[attention_mask[b].bool()]
            rows.append(row.cpu())
        logits, latents = self._engine.put(list(session_ids), rows, all_logits=not last_token_only)
        if self._config.hcache.enabled and not ephemeral:
            for sid, lat, row in zip(session_ids, latents, rows):
                if lat is not None:
                    self._sessions_latents.setdefault(sid, []).append(lat.clone())
                    self._sessions_tokens.setdefault(sid, []).append(row)
        if ephemeral:
            for sid in session_ids:
                self._engine.flush(sid)
        if not last_token_only:
            if all(r.numel() == T for r in rows):
                logits = logits.view(Bsz, T, -1)
            else:  # ragged rows: pad back to [B, T, V] (left-aligned)
                out = logits.new_zeros((Bsz, T, logits.shape[-1]))
                off = 0
                for b, r in enumerate(rows):
                    out[b, :r.numel()] = logits[off:off + r.numel()]
                    off += r.numel()
                logits = out
        return _Output(logits=logits)

    def forward(self, *inputs, **kwargs):
        start = None
        if self.model_profile_enabled:
            if getattr(self, "use_cuda_events", False):
                start, end = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                start.record()
            else:
                t0 = time.perf_counter()
        if self.injected:
            ids = kwargs.pop("input_ids", inputs[0] if inputs else None)
            out = self._injected_forward(ids, **kwargs)
        elif self._config.enable_cuda_graph and self.device.type == "cuda":
            out = self._graph_replay(*inputs, **kwargs) if self._cuda_graph is not None else \
                self._create_cuda_graph(*inputs, **kwargs)
        else:
            out = self.module(*inputs, **kwargs)
        if self.model_profile_enabled:
            if start is not None:
                end.record()
                end.synchronize()
                self._model_times.append(start.elapsed_time(end))
            else:
                self._model_times.append((time.perf_counter() - t0) * 1e3)
        return out

    # ---- model-specific fix-ups under tensor parallelism (ref: engine.py:207-228) ------------------------------------
    def _tp_rank_size(self):
        tp = self._config.tensor_parallel.tp_size
        rank = torch.distributed.get_rank(self.inference_mp_group) if tp > 1 and self.inference_mp_group is not None else 0
        return rank, tp

    def remove_mask_prepare_for_bloom(self) -> None:
        """Older BLOOM implementations expand the attention mask in ``_prepare_attn_mask``; the fused path wants it as given."""
        tr = getattr(self.module, "transformer", None)
        if tr is not None and hasattr(tr, "_prepare_attn_mask"):
            tr._prepare_attn_mask = lambda attention_mask, *args, **kwargs: attention_mask

    def build_alibi_tensor(self) -> None:
        """Make the model's ALiBi builders return this rank's heads only: BLOOM (``build_alibi_tensor``), MPT
        (``build_mpt_alibi_tensor``) and Baichuan-style (``get_alibi_mask``) layouts."""
        from ..module_inject.auto_tp_model_utils import build_bloom_alibi_tensor, slice_heads
        rank, tp = self._tp_rank_size()
        tr = getattr(self.module, "transformer", None)
        if tr is not None and hasattr(tr, "build_alibi_tensor"):
            tr.build_alibi_tensor = lambda attention_mask, num_heads, dtype: build_bloom_alibi_tensor(
                attention_mask, num_heads, dtype, mp_size=tp, mp_rank=rank)
        if tr is not None and hasattr(tr, "build_mpt_alibi_tensor") and not hasattr(tr, "build_mpt_alibi_tensor_orig"):
            tr.build_mpt_alibi_tensor_orig = tr.build_mpt_alibi_tensor
            tr.build_mpt_alibi_tensor = lambda *a, **k: slice_heads(tr.build_mpt_alibi_tensor_orig(*a, **k), 1, rank, tp)
        inner = getattr(self.module, "model", None)
        if inner is not None and hasattr(inner, "get_alibi_mask") and not hasattr(inner, "get_alibi_mask_orig"):
            inner.get_alibi_mask_orig = inner.get_alibi_mask
            inner.get_alibi_mask = lambda *a, **k: slice_heads(inner.get_alibi_mask_orig(*a, **k), 0, rank, tp)

    def build_attn_bias(self) -> None:
        """MPT builds one attention bias ``[1, heads, 1, seq]`` for all heads: keep this rank's heads."""
        from ..module_inject.auto_tp_model_utils import slice_heads
        rank, tp = self._tp_rank_size()
        tr = getattr(self.module, "transformer", None)
        if tr is not None and hasattr(tr, "_attn_bias") and not hasattr(tr, "_attn_bias_orig"):
            tr._attn_bias_orig = tr._attn_bias

            def sliced(*a, **k):
                bias, mask = tr._attn_bias_orig(*a, **k)
                return (slice_heads(bias, 1, rank, tp) if bias is not None else None), mask

            tr._attn_bias = sliced

    def load_model_with_checkpoint(self, r_module: Optional[torch.nn.Module] = None) -> None:
        """Fill ``r_module`` (default: the wrapped module; may be on the meta device) from the checkpoint description given
        at ``init_inference(checkpoint=...)``, slicing to this tensor-parallel rank (ref: engine.py:301-380)."""
        from ..module_inject.load_checkpoint import load_checkpoint_into
        if self._config.checkpoint is None:
            raise ValueError("load_model_with_checkpoint needs init_inference(..., checkpoint=<json | dir>)")
        rank, tp = self._tp_rank_size()
        load_checkpoint_into(r_module if r_module is not None else self.module, self._config.checkpoint,
                             self._config.base_dir or "", tp_size=tp, tp_rank=rank, strict=False)

    def compile(self, backend=None, compile_kwargs=None) -> None:
        """The reference hands the module to ``torch.compile`` (engine.py:608). Here the launch-bound decode loop is
        removed by CUDA-graph capture instead: after this call the un-injected module's forward is captured on its first
        GPU call and replayed afterwards (the injected path already runs graph-captured steps)."""
        if self._is_compiled:
            return
        if self._config.tensor_parallel.tp_size == 1:
            self._config.enable_cuda_graph = True
        self._is_compiled = True

    @property
    def is_compiled(self) -> bool:
        return self._is_compiled

    # ---- CUDA graph for the un-injected module (ref: engine.py:494-520) ------------------------------------
    def _create_cuda_graph(self, *inputs, **kwargs):
        side = torch.cuda.Stream()
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):
            for _ in range(3):
                self.module(*inputs, **kwargs)
        torch.cuda.current_stream().wait_stream(side)
        from ..utils.cuda_graph import capture
        self._graph_inputs, self._graph_kwargs = inputs, kwargs
        self._cuda_graph, self._graph_output = capture(lambda: self.module(*self._graph_inputs, **self._graph_kwargs))
        self._cuda_graph.replay()
        return self._graph_output

    def _graph_replay(self, *inputs, **kwargs):
        for i, t in enumerate(inputs):
            if torch.is_tensor(t):
                self._graph_inputs[i].copy_(t)
        for k, t in kwargs.items():
            if torch.is_tensor(t):
                self._graph_kwargs[k].copy_(t)
        self._cuda_graph.replay()
        return self._graph_output

    # ---- generation ----------------------------------------------------------------------------------------
    @torch.no_grad()
    def generate(self, *inputs, **kwargs):
        """Injected: greedy / temperature / top-k / top-p sampling over sessions. Otherwise HF ``generate``
        (``num_beams > 1`` is rejected exactly as in the reference, engine.py:583-606)."""
        num_beams = kwargs.get("num_beams", 1)
        gc = kwargs.get("generation_config", None)
        if gc is not None:
            num_beams = getattr(gc, "num_beams", num_beams)
        if num_beams > 1:
            raise NotImplementedError("DeepSpeed does not support `num_beams` > 1, if this is important to you please "
                                      "add your request to: https://github.com/deepspeedai/DeepSpeed/issues/2506")
        if not self.injected:
            return self._orig_generate(*inputs, **kwargs)
        ids = kwargs.pop("input_ids", inputs[0] if inputs else None)
        if ids.dim() == 1:
            ids = ids[None]
        max_new = kwargs.get("max_new_tokens", None)
        if max_new is None:
            max_new = max(1, kwargs.get("max_length", ids.shape[1] + 20) - ids.shape[1])
        if ids.shape[1] + max_new > self._config.hcache.max_context:
            raise RuntimeError(f"Input with size {ids.shape[1]} plus {max_new} new tokens exceeds max_context "
                               f"{self._config.hcache.max_context}")
        do_sample = kwargs.get("do_sample", False)
        temperature, top_k, top_p = kwargs.get("temperature", 1.0), kwargs.get("top_k", 0), kwargs.get("top_p", 1.0)
        eos = kwargs.get("eos_token_id", getattr(getattr(self.module, "config", None), "eos_token_id", None))
        eos_set = set(eos if isinstance(eos, (list, tuple)) else [eos]) if eos is not None else set()
        session_ids = kwargs.get("session_ids", None)
        ephemeral = session_ids is None
        Bsz = ids.shape[0]
        if ephemeral:
            session_ids = [self._next_ephemeral_uid + i for i in range(Bsz)]
            self._next_ephemeral_uid += Bsz
        from .sampling import sample_next
        out = self._injected_forward(ids, session_ids=session_ids, last_token_only=True,
                                     attention_mask=kwargs.get("attention_mask"))
        seqs = [ids[b].tolist() for b in range(Bsz)]
        alive = list(range(Bsz))
        for step in range(max_new):
            nxt = sample_next(out.logits, do_sample, temperature, top_k, top_p).cpu()
            still = []
            for j, b in enumerate(alive):
                tok = int(nxt[j])
                seqs[b].append(tok)
                if tok not in eos_set:
                    still.append((b, tok))
            if not still or step == max_new - 1:
                break
            alive = [b for b, _ in still]
            step_ids = torch.tensor([[t] for _, t in still])
            out = self._injected_forward(step_ids, session_ids=[session_ids[b] for b in alive], last_token_only=True)
        if ephemeral:
            for sid in session_ids:
                self._engine.flush(sid)
        width = max(len(s) for s in seqs)
        pad = getattr(getattr(self.module, "config", None), "pad_token_id", None)
        pad = pad if pad is not None else (next(iter(eos_set)) if eos_set else 0)
        return torch.tensor([s + [pad] * (width - len(s)) for s in seqs], dtype=ids.dtype)

    def destroy(self) -> None:
        InferenceEngine.inference_mp_group = None
        InferenceEngine.inference_ep_group = None
        InferenceEngine.expert_mp_group = None
