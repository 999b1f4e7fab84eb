"""``deepspeed/inference/v2/model_implementations/mistral`` import path."""
from .model import MistralInferenceModel  # noqa: F401
from .policy import MistralPolicy  # noqa: F401
