"""Module-name parity with deepspeed/checkpoint/zero_checkpoint.py; the implementation lives in :mod:`.reshape`."""
from .reshape import *  # noqa: F401,F403
from .reshape import meg_2d_parallel_map, model_3d_desc, ZeROCheckpoint  # noqa: F401
