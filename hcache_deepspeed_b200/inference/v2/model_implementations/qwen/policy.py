"""Policy of the ``qwen`` family; the implementation is shared (:class:`..InferenceV2Policy`)."""
from .. import QwenPolicy  # noqa: F401

MODEL_TYPE = "qwen"
