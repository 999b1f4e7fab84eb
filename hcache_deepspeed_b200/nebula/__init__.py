"""Nebula (Azure asynchronous checkpoint service) configuration (``deepspeed/nebula/{config,constants}.py``). The
service itself is hosted; :class:`~hcache_deepspeed_b200.runtime.checkpoint_engine.NebulaCheckpointEngine` falls back
to synchronous torch checkpoints when it is not reachable."""
from .config import DeepSpeedNebulaConfig  # noqa: F401
from .constants import *  # noqa: F401,F403
