"""``deepspeed/monitor/csv_monitor.py`` name; implementation in :mod:`.monitor`."""
from .monitor import CSVMonitor as csvMonitor  # noqa: F401
from .monitor import CSVMonitor  # noqa: F401
