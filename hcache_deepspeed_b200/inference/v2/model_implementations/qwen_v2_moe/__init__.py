"""``deepspeed/inference/v2/model_implementations/qwen_v2_moe`` import path."""
from .model import Qwen2MoeInferenceModel  # noqa: F401
from .policy import Qwen2MoePolicy  # noqa: F401
