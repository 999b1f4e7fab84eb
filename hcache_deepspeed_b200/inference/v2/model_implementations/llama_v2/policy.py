"""Policy of the ``llama`` family; the implementation is shared (:class:`..InferenceV2Policy`)."""
from .. import Llama2Policy  # noqa: F401

MODEL_TYPE = "llama"
