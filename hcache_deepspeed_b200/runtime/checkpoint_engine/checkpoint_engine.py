"""Checkpoint engines: the storage back end behind ``save_checkpoint`` / ``load_checkpoint``.
Reference: ``deepspeed/runtime/checkpoint_engine/{checkpoint_engine,torch_checkpoint_engine,nebula_checkpoint_engine}.py``."""
from __future__ import annotations

import os

import torch

from ...utils.logging import logger


class CheckpointEngine:
    def __init__(self, config_params=None):
        self.config_params = config_params

    def create(self, tag):
        """Called once per checkpoint tag before the first save."""

    def makedirs(self, path, exist_ok=False):
        os.makedirs(path, exist_ok=exist_ok)

    def save(self, state_dict, path: str):
        raise NotImplementedError

    def load(self, path: str, map_location=None):
        raise NotImplementedError

    def commit(self, tag) -> bool:
        """All files of ``tag`` are written: make the checkpoint visible."""
        return True


class TorchCheckpointEngine(CheckpointEngine):
    def save(self, state_dict, path: str):
        tmp = path + ".tmp"
        torch.save(state_dict, tmp)
        os.replace(tmp, path)  # a crash never leaves a truncated file under the final name

    def load(self, path: str, map_location=None):
        return torch.load(path, map_location=map_location, weights_only=False)


class NebulaCheckpointEngine(TorchCheckpointEngine):
    """Azure Nebula asynchronous snapshots are a hosted service; without it this engine persists synchronously
    through the torch engine and keeps the ``commit`` protocol."""

    def __init__(self, config_params=None):
        super().__init__(config_params)
        logger.warning("Nebula service is not available here: falling back to synchronous torch checkpoints")
