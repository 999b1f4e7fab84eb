"""Policy of the ``qwen2_moe`` family; the implementation is shared (:class:`..InferenceV2Policy`)."""
from .. import Qwen2MoePolicy  # noqa: F401

MODEL_TYPE = "qwen2_moe"
