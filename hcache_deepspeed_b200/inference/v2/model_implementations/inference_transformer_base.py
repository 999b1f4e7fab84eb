"""``DSTransformerModelBase`` under its reference module name
(``deepspeed/inference/v2/model_implementations/inference_transformer_base.py``)."""
from .inference_model_base import DSInferenceModelBase, DSTransformerModelBase  # noqa: F401
