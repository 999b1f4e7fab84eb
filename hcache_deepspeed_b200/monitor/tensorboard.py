"""``deepspeed/monitor/tensorboard.py`` name; implementation in :mod:`.monitor`."""
from .monitor import TensorBoardMonitor  # noqa: F401
