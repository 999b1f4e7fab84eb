"""``deepspeed/inference/v2/model_implementations/falcon`` import path."""
from .model import FalconInferenceModel  # noqa: F401
from .policy import FalconPolicy  # noqa: F401
