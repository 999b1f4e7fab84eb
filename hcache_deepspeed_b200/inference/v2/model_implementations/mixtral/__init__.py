"""``deepspeed/inference/v2/model_implementations/mixtral`` import path."""
from .model import MixtralInferenceModel  # noqa: F401
from .policy import MixtralPolicy  # noqa: F401
