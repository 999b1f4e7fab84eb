"""Package-level constants (``deepspeed/constants.py``)."""
import os
from datetime import timedelta

TORCH_DISTRIBUTED_DEFAULT_PORT = 29500
# process-group timeout; the default matches torch's NCCL default and can be raised for slow first-iteration compiles
default_pg_timeout = timedelta(minutes=int(os.getenv("DEEPSPEED_TIMEOUT", default=30)))
INFERENCE_GENERIC_MODE = "generic"
INFERENCE_SPECIALIZED_MODE = "specialized"
CROSS_RANK, CROSS_SIZE, LOCAL_RANK = "CROSS_RANK", "CROSS_SIZE", "LOCAL_RANK"
