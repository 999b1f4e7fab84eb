"""``DeepSpeedSelfAttention`` / ``BloomSelfAttention`` under their v1 path (``ops/transformer/inference/ds_attention.py``)."""
from ....model_implementations.transformers.ds_transformer import DeepSpeedSelfAttention  # noqa: F401

BloomSelfAttention = DeepSpeedSelfAttention      # ALiBi is a config flag of the one attention class here
