"""``deepspeed/inference/v2/model_implementations/qwen`` import path."""
from .model import QwenInferenceModel  # noqa: F401
from .policy import QwenPolicy  # noqa: F401
