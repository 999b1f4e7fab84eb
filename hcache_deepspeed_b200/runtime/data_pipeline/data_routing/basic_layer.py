"""``deepspeed/runtime/data_pipeline/data_routing/basic_layer.py`` name."""
from ..random_ltd import RandomLayerTokenDrop  # noqa: F401
