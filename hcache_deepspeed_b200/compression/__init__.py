"""Model compression (``deepspeed.compression``): quantisation-aware training, sparse / row / head / channel pruning,
activation quantisation and layer reduction, driven by the ``compression_training`` config section."""
from .compress import init_compression, redundancy_clean  # noqa: F401
from .scheduler import compression_scheduler  # noqa: F401
from .helper import convert_conv1d_to_linear  # noqa: F401
