"""``DeepSpeedMLP`` under its v1 path (``ops/transformer/inference/ds_mlp.py``)."""
from ....model_implementations.transformers.ds_transformer import DeepSpeedMLP  # noqa: F401
