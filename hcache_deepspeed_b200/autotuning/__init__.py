from .autotuner import Autotuner  # noqa: F401
from .tuner import GridSearchTuner, ModelBasedTuner, RandomTuner  # noqa: F401
from .config import DeepSpeedAutotuningConfig, get_autotuning_config  # noqa: F401
from .scheduler import Job, Node, Reservation, ResourceManager  # noqa: F401
