"""Policy of the ``phi`` family; the implementation is shared (:class:`..InferenceV2Policy`)."""
from .. import PhiPolicy  # noqa: F401

MODEL_TYPE = "phi"
