"""Model class of the ``mixtral`` family. The reference has one hand-written forward per family; here every family runs
on the one ragged executor, parameterised by the family's ``ArchSpec`` (``models/families.py``)."""
from .....models.transformer import RaggedTransformer


class MixtralInferenceModel(RaggedTransformer):
    model_type = "mixtral"
