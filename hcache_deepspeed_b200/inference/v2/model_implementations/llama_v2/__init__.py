"""``deepspeed/inference/v2/model_implementations/llama_v2`` import path."""
from .model import Llama2InferenceModel  # noqa: F401
from .policy import Llama2Policy  # noqa: F401
