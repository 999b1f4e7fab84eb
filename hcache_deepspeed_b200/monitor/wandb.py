"""``deepspeed/monitor/wandb.py`` name; implementation in :mod:`.monitor`."""
from .monitor import WandbMonitor  # noqa: F401
