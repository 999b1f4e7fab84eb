"""``deepspeed/ops/lion/cpu_lion.py`` name."""
from .lion import DeepSpeedCPULion  # noqa: F401
