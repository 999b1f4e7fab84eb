"""Marker base classes for the framework's optimizer wrappers (``deepspeed/runtime/base_optimizer.py:14-63``).

The concrete ZeRO optimizers (:mod:`.zero.stage_1_and_2`, :mod:`.zero.stage3`, :mod:`.bf16_optimizer`) are registered as
virtual subclasses, so ``isinstance(opt, ZeROOptimizer)`` is true for them without a common implementation base."""
from __future__ import annotations

from abc import ABCMeta


class DeepSpeedOptimizer(metaclass=ABCMeta):
    pass


class ZeROOptimizer(DeepSpeedOptimizer):

    def load_hp_checkpoint_state_from_checkpoint_dir(self, lp_groups_name: str, checkpoint_dir: str) -> None:
        """Fill this rank's fp32 master slices and Adam moments from the per-parameter files of a universal checkpoint.
        ``lp_groups_name`` names the attribute holding the low-precision parameter groups (``bit16_groups`` ...)."""
        from ..checkpoint.universal_checkpoint import load_universal_into_optimizer
        load_universal_into_optimizer(self, checkpoint_dir, lp_groups_name=lp_groups_name)


def _register() -> None:
    from .zero.stage_1_and_2 import DeepSpeedZeroOptimizer
    from .zero.stage3 import DeepSpeedZeroOptimizer_Stage3
    for cls in (DeepSpeedZeroOptimizer, DeepSpeedZeroOptimizer_Stage3):
        ZeROOptimizer.register(cls)
        if not hasattr(cls, "load_hp_checkpoint_state_from_checkpoint_dir"):
            cls.load_hp_checkpoint_state_from_checkpoint_dir = ZeROOptimizer.load_hp_checkpoint_state_from_checkpoint_dir
    from .fp16.fused_optimizer import FP16_Optimizer
    DeepSpeedOptimizer.register(FP16_Optimizer)


_register()
