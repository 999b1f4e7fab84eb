"""DeepSpeedHybridEngine (RLHF): one engine that TRAINS with ZeRO and GENERATES with the inference executor on the
same weights. ``generate()`` gathers the (possibly ZeRO-3 partitioned) parameters, refreshes the inference engine's
weight copy when the training step counter moved, and runs the sm_100a ragged engine (or the HF module when the
architecture has no policy). Reference: ``deepspeed/runtime/hybrid_engine.py:30-445``."""
from __future__ import annotations

import time

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
        self._inference_engine = None
        self._weights_step = -1
        self._generate_latency = self._training_latency = 0.0
        self._total_batch_size = None
        self._t_start = None

    def _gathered_state_dict(self):
        if self.zero_optimization_stage() == 3:
            return self._zero3_consolidated_16bit_state_dict()
        return {k: v.detach() for k, v in self.module.state_dict().items()}

    def _refresh_inference_engine(self):
        if self._inference_engine is not None and self._weights_step == self.global_steps:
            return
        from .. import init_inference
        mt = getattr(getattr(self.module, "config", None), "model_type", None)
        from ..inference.v2.model_implementations import MODEL_TYPE_TO_POLICY
        inject = mt in MODEL_TYPE_TO_POLICY and (self.device.type != "cuda" or next(self.module.parameters()).dtype in
                                                 (torch.float16, torch.bfloat16))
        if self.zero_optimization_stage() == 3:
            from .zero.partition_parameters import GatheredParameters
            ctx = GatheredParameters(list(self.module.parameters()))
        else:
            import contextlib
            ctx = contextlib.nullcontext()
        with ctx:
            dtype = next(self.module.parameters()).dtype
            self._inference_engine = init_inference(self.module, dtype=dtype, replace_with_kernel_inject=inject,
                                                    hcache={"max_context": max(2048, self.max_out_tokens * 2)}) if inject \
                else init_inference(self.module, dtype=dtype)
        self._weights_step = self.global_steps
        log_dist(f"hybrid engine: inference weights refreshed at step {self.global_steps} (kernel_inject={inject})", [0])

    @torch.no_grad()
    def generate(self, *inputs, **kwargs):
        t0 = time.time()
        was_training = self.module.training
        self.module.eval()
        self._refresh_inference_engine()
        out = self._inference_engine.generate(*inputs, **kwargs)
        if self.release_inference_cache and getattr(self._inference_engine, "ragged_engine", None) is not None:
            self._inference_engine = None
        self.module.train(was_training)
        self._generate_latency = time.time() - t0
        return out

    def eval(self):
        if self._t_start is not None:
            self._training_latency = time.time() - self._t_start
        super().eval()
        return self

    def train(self, mode: bool = True):
        self._t_start = time.time()
        super().train(mode)
        return self
