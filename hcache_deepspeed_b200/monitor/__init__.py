from .monitor import CSVMonitor, MonitorMaster, TensorBoardMonitor, WandbMonitor  # noqa: F401
from .config import DeepSpeedMonitorConfig  # noqa: F401
