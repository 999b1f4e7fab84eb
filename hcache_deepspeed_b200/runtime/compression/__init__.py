"""Sign-bit packing used by the 1-bit optimizers' compressed all-reduce (``deepspeed/runtime/compression/cupy.py``)."""
from .cupy import CupyBackend  # noqa: F401
