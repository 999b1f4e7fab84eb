"""Policy of the ``mistral`` family; the implementation is shared (:class:`..InferenceV2Policy`)."""
from .. import MistralPolicy  # noqa: F401

MODEL_TYPE = "mistral"
