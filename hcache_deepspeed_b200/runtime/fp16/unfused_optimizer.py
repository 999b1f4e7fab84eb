"""``deepspeed/runtime/fp16/unfused_optimizer.py`` name; both fp16 wrappers live in :mod:`.fused_optimizer`."""
from .fused_optimizer import FP16_UnfusedOptimizer  # noqa: F401
