"""``timers`` config section (``deepspeed/utils/config.py``)."""
from ..config.config_utils import DeepSpeedConfigModel

TIMERS, TIMERS_THROUGHPUT = "timers", "throughput"


class DeepSpeedThroughputTimerConfig(DeepSpeedConfigModel):
    enabled: bool = True
    synchronized: bool = True      # device-synchronise around the timed region (accurate, slower)


def get_timers_config(param_dict: dict) -> DeepSpeedThroughputTimerConfig:
    sec = (param_dict.get(TIMERS) or {}).get(TIMERS_THROUGHPUT) or {}
    return DeepSpeedThroughputTimerConfig(**sec)
