"""Turns each compression technique on when training reaches its ``schedule_offset`` (and off again at
``schedule_offset_end`` where given). Reference: ``deepspeed/compression/scheduler.py:10-180``."""
from __future__ import annotations

TECHNIQUES = ("weight_quantization", "activation_quantization", "sparse_pruning", "row_pruning", "head_pruning", "channel_pruning")
_FLAG = {"weight_quantization": "weight_quantization_enabled", "activation_quantization": "activation_quantization_enabled",
         "sparse_pruning": "sparse_pruning_enabled", "row_pruning": "row_pruning_enabled", "head_pruning": "head_pruning_enabled",
         "channel_pruning": "channel_pruning_enabled"}


class compression_scheduler:

    def __init__(self, model, compression_config):
        self.model = model.module if hasattr(model, "module") and not hasattr(model, "_compression_plan") else model
        self.compression_config = compression_config
        self.training_steps = 0
        self.verbose = {t: False for t in TECHNIQUES}
        self.plan = getattr(self.model, "_compression_plan", [])

    def _check(self, tech: str):
        sec = self.compression_config.get(tech, {})
        shared = sec.get("shared_parameters", {})
        if not shared.get("enabled", False):
            return
        on = self.training_steps >= shared.get("schedule_offset", 0)
        end = shared.get("schedule_offset_end")
        if end is not None and self.training_steps > end:
            on = False
        mods = dict(self.model.named_modules())
        for name, t, _, _ in self.plan:
            if t == tech and name in mods and hasattr(mods[name], _FLAG[tech]):
                setattr(mods[name], _FLAG[tech], on)
        if on and not self.verbose[tech]:
            self.verbose[tech] = True

    def step(self, step_zero_check: bool = False):
        if not step_zero_check:
            self.training_steps += 1
        for t in TECHNIQUES:
            self._check(t)
