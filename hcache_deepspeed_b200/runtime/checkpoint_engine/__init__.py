from .checkpoint_engine import CheckpointEngine, NebulaCheckpointEngine, TorchCheckpointEngine  # noqa: F401
