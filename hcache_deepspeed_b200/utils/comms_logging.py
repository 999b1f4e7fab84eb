"""``deepspeed/utils/comms_logging.py`` name; the logger lives with the comm layer."""
from ..comm.comms_logging import *  # noqa: F401,F403
from ..comm.comms_logging import CommsLogger  # noqa: F401
