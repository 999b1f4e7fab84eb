from . import config as _config
from .config import AUTOTP_MODE, get_tensor_parallel_config  # noqa: F401
from .tp_manager import TpTrainingManager  # noqa: F401


def set_autotp_mode(training: bool = False) -> None:
    """Tell AutoTP whether sharded layers are trained (``tp_model_init`` switches to TRAINING): self-partitioning layers
    (``module_inject.layers.TensorParallel_Layer``) then leave their shards trainable; in INFERENCE mode they are frozen.
    One flag, also reachable as ``module_inject.layers.set_autotp_mode``."""
    _config._CURRENT_MODE = AUTOTP_MODE.TRAINING if training else AUTOTP_MODE.INFERENCE


def get_autotp_mode() -> AUTOTP_MODE:
    return getattr(_config, "_CURRENT_MODE", AUTOTP_MODE.INFERENCE)
