"""Policy of the ``falcon`` family; the implementation is shared (:class:`..InferenceV2Policy`)."""
from .. import FalconPolicy  # noqa: F401

MODEL_TYPE = "falcon"
