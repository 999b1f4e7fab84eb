"""Model class of the ``qwen2_moe`` family. The reference has one hand-written forward per family; here every family runs
on the one ragged executor, parameterised by the family's ``ArchSpec`` (``models/families.py``)."""
from .....models.transformer import RaggedTransformer


class Qwen2MoeInferenceModel(RaggedTransformer):
    model_type = "qwen2_moe"
