"""``deepspeed/inference/v2/model_implementations/opt`` import path."""
from .model import OPTInferenceModel  # noqa: F401
from .policy import OPTPolicy  # noqa: F401
