"""Checkpoint-name mapping of the ``qwen2`` family: which checkpoint tensors feed which fused weights. The table
lives with the architecture description (``models/families.py`` / ``models/weights.py``)."""
from .....models.families import get_mapper  # noqa: F401
from .....models.weights import WeightAssembler, load_weights  # noqa: F401

MODEL_TYPE = "qwen2"
MAPPER = get_mapper(MODEL_TYPE)
