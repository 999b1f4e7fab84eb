"""``deepspeed/runtime/checkpoint_engine/nebula_checkpoint_engine.py`` name."""
from .checkpoint_engine import NebulaCheckpointEngine  # noqa: F401
