"""Policy of the ``qwen2`` family; the implementation is shared (:class:`..InferenceV2Policy`)."""
from .. import Qwen2Policy  # noqa: F401

MODEL_TYPE = "qwen2"
