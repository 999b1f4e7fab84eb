"""Checkpoint engine contract (``deepspeed/inference/v2/checkpoint/base_engine.py:16-41``)."""
from abc import ABC, abstractmethod
from typing import Iterable, Tuple

import torch

MEGATRON = "megatron"
HUGGINGFACE = "huggingface"


class CheckpointEngineBase(ABC):
    """Streams ``(parameter_name, tensor)`` pairs; names are the ones of the source framework (HuggingFace),
    which the per-family mappers in :mod:`hcache_deepspeed_b200.models.families` understand."""

    @abstractmethod
    def parameters(self) -> Iterable[Tuple[str, torch.Tensor]]:
        ...
