from .config import DeepSpeedInferenceConfig  # noqa: F401
from .ops import InferenceOps  # noqa: F401
