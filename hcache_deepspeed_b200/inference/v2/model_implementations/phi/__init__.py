"""``deepspeed/inference/v2/model_implementations/phi`` import path."""
from .model import PhiInferenceModel  # noqa: F401
from .policy import PhiPolicy  # noqa: F401
