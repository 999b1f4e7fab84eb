"""``deepspeed/ops/lion/fused_lion.py`` name."""
from .lion import FusedLion  # noqa: F401
