"""Model class of the ``llama`` family. The reference has one hand-written forward per family; here every family runs
on the one ragged executor, parameterised by the family's ``ArchSpec`` (``models/families.py``)."""
from .....models.transformer import RaggedTransformer


class Llama2InferenceModel(RaggedTransformer):
    model_type = "llama"
