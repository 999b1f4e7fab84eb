from .tp_manager import TpTrainingManager  # noqa: F401
