"""``deepspeed/monitor/comet.py`` name; implementation in :mod:`.monitor`."""
from .monitor import CometMonitor, EventsLogScheduler  # noqa: F401
