"""DeepSpeedHybridEngine (RLHF): one engine that TRAINS with ZeRO and GENERATES with the inference executor on the
same weights. Reference: ``deepspeed/runtime/hybrid_engine.py:30-445``.

``generate()``:
  1. gathers the (possibly ZeRO-3 partitioned) parameters;
  2. folds every LoRA adapter into its base weight (``fuse_lora_weight``) so that the inference copy sees W + s * B A;
  3. refreshes the inference engine's weight copy when the weights moved since the last refresh (optimizer step counter,
     or ``mark_weights_dirty()``), un-fuses the adapters again, releases the gathered parameters;
  4. runs the sm_100a ragged engine (families with an executor policy) or the HF module's own ``generate``.
The reference keeps per-layer inference containers that alias the training tensors; here the executor owns a re-laid-out
copy (fused QKV, interleaved gate/up, TP slices), so "share the weights" becomes "refresh the copy once per step".

LoRA layouts recognised: DeepSpeed-Chat's ``LinearLayer_LoRA`` (``weight`` [out, in], ``lora_right_weight`` [in, r],
``lora_left_weight`` [r, out], ``lora_scaling``) and this package's ``LoRAOptimizedLinear`` without base-weight sharding.
"""
from __future__ import annotations

import contextlib
import time
from typing import Iterator, List, Tuple

import torch

from ..utils.logging import log_dist
from .engine import DeepSpeedEngine


class DeepSpeedHybridEngine(DeepSpeedEngine):

    def __init__(self, *args, **kwargs):
        super().__init__(*args, **kwargs)
        he = self._config.hybrid_engine
        self.max_out_tokens = he.get("max_out_tokens", 512)
        self.inference_tp_size = he.get("inference_tp_size", 1)
        self.release_inference_cache = he.get("release_inference_cache", False)
        self.pin_parameters = he.get("pin_parameters", True)
        self.tp_gather_partition_size = he.get("tp_gather_partition_size", 8)
        self._inference_engine = None
        self._weights_step = -1
        self._weights_dirty = False
        self.is_lora_fused = False
        self._lora_layers: List[Tuple[torch.nn.Module, str]] = list(self._find_lora_layers())
        # latency accounting (printed by eval() like the reference's RLHF loop does)
        self._generate_latency = self._training_latency = self._gather_latency = 0.0
        self._iters = 0
        self._training_start_time = None
        self._t_start = None
        self._total_batch_size = None
        self._total_latency = 0.0
        if self._lora_layers:
            log_dist(f"hybrid engine: {len(self._lora_layers)} LoRA layer(s) will be fused for generation", [0])

    # ---- LoRA ------------------------------------------------------------------------------------------------------
    def _find_lora_layers(self) -> Iterator[Tuple[torch.nn.Module, str]]:
        from ..linear.optimized_linear import LoRAOptimizedLinear
        for m in self.module.modules():
            if isinstance(m, LoRAOptimizedLinear):
                yield m, "optimized"
            elif all(hasattr(m, a) for a in ("lora_right_weight", "lora_left_weight", "lora_scaling", "weight")):
                yield m, "chat"

    @staticmethod
    def _lora_delta(m: torch.nn.Module, kind: str) -> torch.Tensor:
        if kind == "chat":   # y = x W^T + s * (x R) L   ->   W += s * (R L)^T
            return (float(m.lora_scaling) * (m.lora_right_weight.float() @ m.lora_left_weight.float()).t()).to(m.weight.dtype)
        if m.zero_shards != 1:
            raise NotImplementedError("LoRAOptimizedLinear with base_weight_sharding > 1 cannot be fused in place")
        return (float(m.lora_scaling_factor) * (m.lora_weight_2.weight.float() @ m.lora_weight_1.weight.float())).to(m.weight.dtype)

    @torch.no_grad()
    def fuse_lora_weight(self) -> None:
        if self.is_lora_fused:
            return
        for m, kind in self._lora_layers:
            m.weight.data.add_(self._lora_delta(m, kind))
        self.is_lora_fused = True

    @torch.no_grad()
    def unfuse_lora_weight(self) -> None:
        if not self.is_lora_fused:
            return
        for m, kind in self._lora_layers:
            m.weight.data.sub_(self._lora_delta(m, kind))
        self.is_lora_fused = False

    unfuse_lora_weight_non_pinned = unfuse_lora_weight

    # ---- inference copy --------------------------------------------------------------------------------------------------
    def mark_weights_dirty(self) -> None:
        """Call after changing parameters outside ``step()`` (e.g. loading a checkpoint into the module by hand)."""
        self._weights_dirty = True

    def _gathered_state_dict(self):
        if self.zero_optimization_stage() == 3:
            return self._zero3_consolidated_16bit_state_dict()
        return {k: v.detach() for k, v in self.module.state_dict().items()}

    def _gather_context(self):
        if self.zero_optimization_stage() == 3:
            from .zero.partition_parameters import GatheredParameters
            return GatheredParameters(list(self.module.parameters()))
        return contextlib.nullcontext()

    def _can_inject(self) -> bool:
        from ..inference.v2.model_implementations import MODEL_TYPE_TO_POLICY
        mt = getattr(getattr(self.module, "config", None), "model_type", None)
        if mt not in MODEL_TYPE_TO_POLICY:
            return False
        if any(kind == "optimized" and m.zero_shards != 1 for m, kind in self._lora_layers):
            return False                # sharded base weights: the module's own forward assembles them
        return self.device.type != "cuda" or next(self.module.parameters()).dtype in (torch.float16, torch.bfloat16)

    def _refresh_inference_engine(self) -> None:
        fresh = self._inference_engine is not None and self._weights_step == self.global_steps and not self._weights_dirty
        if fresh:
            return
        from .. import init_inference
        inject = self._can_inject()
        t0 = time.time()
        with self._gather_context():
            dtype = next(self.module.parameters()).dtype
            if inject:
                self.fuse_lora_weight()
                kw = {"tensor_parallel": {"tp_size": self.inference_tp_size}} if self.inference_tp_size > 1 else {}
                try:
                    self._inference_engine = init_inference(self.module, dtype=dtype, replace_with_kernel_inject=True,
                                                            hcache={"max_context": max(2048, self.max_out_tokens * 2)}, **kw)
                finally:
                    self.unfuse_lora_weight()
            else:
                self._inference_engine = init_inference(self.module, dtype=dtype)
        self._gather_latency = time.time() - t0
        self._weights_step, self._weights_dirty = self.global_steps, False
        log_dist(f"hybrid engine: inference weights refreshed at step {self.global_steps} (kernel_inject={inject}, "
                 f"lora_layers={len(self._lora_layers)})", [0])

    def retake_inference_cache(self) -> None:
        """Make sure an inference engine (and its KV-cache arena) exists; after ``release_inference_cache`` dropped it
        this rebuilds it from the current weights."""
        self._refresh_inference_engine()

    def release_inference_engine(self) -> None:
        """Free the inference copy of the weights and the KV-cache arena (memory goes back to training)."""
        self._inference_engine = None
        if self.device.type == "cuda":
            torch.cuda.empty_cache()

    # ---- generation ---------------------------------------------------------------------------------------------------
    @torch.no_grad()
    def generate(self, *inputs, **kwargs):
        t0 = time.time()
        was_training = self.module.training
        self.module.eval()
        self._gather_latency = 0.0
        self._refresh_inference_engine()
        injected = bool(getattr(self._inference_engine, "injected", False))
        if not injected and self._lora_layers and not any(k == "optimized" and m.zero_shards != 1 for m, k in self._lora_layers):
            # HF path: fused adapters save the two skinny GEMMs per layer and token
            with self._gather_context():
                self.fuse_lora_weight()
                try:
                    out = self._inference_engine.generate(*inputs, **kwargs)
                finally:
                    self.unfuse_lora_weight()
        elif not injected:
            with self._gather_context():
                out = self._inference_engine.generate(*inputs, **kwargs)
        else:
            out = self._inference_engine.generate(*inputs, **kwargs)
        if self.release_inference_cache:
            self.release_inference_engine()
        self.module.train(was_training)
        self._generate_latency = time.time() - t0 - self._gather_latency
        first = inputs[0] if inputs else kwargs.get("input_ids")
        if first is not None and hasattr(first, "shape"):
            self._total_batch_size = first.shape[0] if first.dim() > 1 else 1
        return out

    # ---- mode switches + accounting -------------------------------------------------------------------------------------
    def eval(self):
        if self._t_start is not None:
            self._training_latency = time.time() - self._t_start
            self._total_latency += self._generate_latency + self._gather_latency + self._training_latency
            self._iters += 1
            if self._iters > 0 and self.global_rank == 0 and self._generate_latency > 0:
                e2e = self._generate_latency + self._gather_latency + self._training_latency
                log_dist(f"|E2E latency={e2e:.2f}s |Gather latency={self._gather_latency:.2f}s ({100 * self._gather_latency / e2e:.1f}%) "
                         f"|Generate time={self._generate_latency:.2f}s ({100 * self._generate_latency / e2e:.1f}%) "
                         f"|Training time={self._training_latency:.2f}s ({100 * self._training_latency / e2e:.1f}%) "
                         f"|Others={0.0:.2f} |CurSamplesPerSec={(self._total_batch_size or 0) / e2e:.2f} "
                         f"|AvgSamplesPerSec={(self._total_batch_size or 0) * self._iters / max(self._total_latency, 1e-9):.2f}", [0])
            self._t_start = None
        super().eval()
        return self

    def train(self, mode: bool = True):
        if mode:
            self._t_start = time.time()
        super().train(mode)
        return self

    def step(self, lr_kwargs=None):
        out = super().step(lr_kwargs=lr_kwargs) if lr_kwargs is not None else super().step()
        return out

    def latency_stats(self):
        return {"generate_s": self._generate_latency, "gather_s": self._gather_latency, "training_s": self._training_latency,
                "iterations": self._iters}
