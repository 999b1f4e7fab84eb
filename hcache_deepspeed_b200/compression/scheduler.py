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

    # ---- the reference's per-technique entry points (scheduler.py:33-170) ------------------------------------------------
    def make_init(self) -> None:
        """(Re)read the plan the model was prepared with: which modules each technique applies to."""
        self.plan = getattr(self.model, "_compression_plan", [])
        self.different_compression_methods = {
            t: {"technique_enabled": bool(self.compression_config.get(t, {}).get("shared_parameters", {}).get("enabled", False)),
                "shared_parameters": self.compression_config.get(t, {}).get("shared_parameters", {}),
                "different_groups": [(name, params) for name, tech, params, _ in self.plan if tech == t]}
            for t in TECHNIQUES}

    def check_weight_quantization(self) -> None:
        self._check("weight_quantization")

    def check_activation_quantization(self) -> None:
        self._check("activation_quantization")

    def check_sparse_pruning(self) -> None:
        self._check("sparse_pruning")

    def check_head_pruning(self) -> None:
        self._check("head_pruning")

    def check_row_pruning(self) -> None:
        self._check("row_pruning")

    def check_channel_pruning(self) -> None:
        self._check("channel_pruning")

    def check_all_modules(self) -> None:
        for t in TECHNIQUES:
            self._check(t)

    def step(self, step_zero_check: bool = False):
        if not step_zero_check:
            self.training_steps += 1
        for t in TECHNIQUES:
            self._check(t)
