from .config import DeepSpeedInferenceConfig  # noqa: F401
from .ops import InferenceOps  # noqa: F401
from .moe_inference import DeepSpeedMoEInference, DeepSpeedMoEInferenceConfig, DeepSpeedMoEMLP  # noqa: F401,E402
