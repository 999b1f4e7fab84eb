"""``nhwc_bias_add`` under its v1 path (``deepspeed/ops/transformer/inference/bias_add.py``)."""
from ...spatial import nhwc_bias_add  # noqa: F401
