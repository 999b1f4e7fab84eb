"""Policy of the ``opt`` family; the implementation is shared (:class:`..InferenceV2Policy`)."""
from .. import OPTPolicy  # noqa: F401

MODEL_TYPE = "opt"
