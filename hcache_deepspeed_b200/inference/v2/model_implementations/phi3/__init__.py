"""``deepspeed/inference/v2/model_implementations/phi3`` import path."""
from .model import Phi3InferenceModel  # noqa: F401
from .policy import Phi3Policy  # noqa: F401
