"""Model class of the ``qwen2`` family. The reference has one hand-written forward per family; here every family runs
on the one ragged executor, parameterised by the family's ``ArchSpec`` (``models/families.py``)."""
from .....models.transformer import RaggedTransformer


class Qwen2InferenceModel(RaggedTransformer):
    model_type = "qwen2"
