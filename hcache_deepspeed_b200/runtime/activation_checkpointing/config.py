"""``activation_checkpointing`` config section (``deepspeed/runtime/activation_checkpointing/config.py``)."""
from ..config import ActivationCheckpointingConfig as DeepSpeedActivationCheckpointingConfig  # noqa: F401

ACT_CHKPT = "activation_checkpointing"


def get_activation_checkpointing_config(param_dict: dict) -> DeepSpeedActivationCheckpointingConfig:
    return DeepSpeedActivationCheckpointingConfig(**(param_dict.get(ACT_CHKPT) or {}))
