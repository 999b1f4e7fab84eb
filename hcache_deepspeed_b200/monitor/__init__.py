from .monitor import CometMonitor, CSVMonitor, EventsLogScheduler, MonitorMaster, TensorBoardMonitor, WandbMonitor  # noqa: F401
from .config import DeepSpeedMonitorConfig  # noqa: F401
