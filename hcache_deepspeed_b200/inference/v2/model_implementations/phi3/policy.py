"""Policy of the ``phi3`` family; the implementation is shared (:class:`..InferenceV2Policy`)."""
from .. import Phi3Policy  # noqa: F401

MODEL_TYPE = "phi3"
