from .autotuner import Autotuner  # noqa: F401
from .tuner import GridSearchTuner, ModelBasedTuner, RandomTuner  # noqa: F401
