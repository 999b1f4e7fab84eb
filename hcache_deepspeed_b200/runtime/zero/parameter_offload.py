"""Module-name parity with ``deepspeed/runtime/zero/parameter_offload.py``: the hook installer that gathers / releases
ZeRO-3 parameters around every sub-module lives in :mod:`.stage3`."""
from .stage3 import DeepSpeedZeRoOffload  # noqa: F401
