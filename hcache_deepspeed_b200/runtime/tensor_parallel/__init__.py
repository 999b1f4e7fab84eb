from .config import AUTOTP_MODE, get_tensor_parallel_config  # noqa: F401
from .tp_manager import TpTrainingManager  # noqa: F401

_autotp_mode = AUTOTP_MODE.INFERENCE


def set_autotp_mode(training: bool = False) -> None:
    """Tell AutoTP whether sharded layers must keep autograd support (``tp_model_init`` switches it to TRAINING; the
    sharded layers of this package carry their backward collectives in both modes, so the flag is informational)."""
    global _autotp_mode
    _autotp_mode = AUTOTP_MODE.TRAINING if training else AUTOTP_MODE.INFERENCE


def get_autotp_mode() -> AUTOTP_MODE:
    return _autotp_mode
