"""``deepspeed/runtime/checkpoint_engine/torch_checkpoint_engine.py`` name."""
from .checkpoint_engine import TorchCheckpointEngine  # noqa: F401
