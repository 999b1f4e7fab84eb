"""``deepspeed/inference/v2/model_implementations/qwen_v2`` import path."""
from .model import Qwen2InferenceModel  # noqa: F401
from .policy import Qwen2Policy  # noqa: F401
