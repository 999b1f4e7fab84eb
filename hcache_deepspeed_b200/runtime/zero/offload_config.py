"""``deepspeed/runtime/zero/offload_config.py`` names."""
from .config import *  # noqa: F401,F403
from .offload_states import OffloadStateTypeEnum  # noqa: F401
