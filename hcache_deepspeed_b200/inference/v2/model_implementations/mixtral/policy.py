"""Policy of the ``mixtral`` family; the implementation is shared (:class:`..InferenceV2Policy`)."""
from .. import MixtralPolicy  # noqa: F401

MODEL_TYPE = "mixtral"
