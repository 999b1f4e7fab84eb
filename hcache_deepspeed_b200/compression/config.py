"""``compression_training`` section accessor (``deepspeed/compression/config.py``)."""
from .compress import get_compression_config  # noqa: F401
