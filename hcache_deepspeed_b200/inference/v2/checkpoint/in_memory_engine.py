"""Checkpoint engine over a live ``torch.nn.Module`` (``deepspeed/inference/v2/checkpoint/in_memory_engine.py:13-40``)."""
from typing import Iterable, Tuple

import torch

from .base_engine import CheckpointEngineBase


class InMemoryModelEngine(CheckpointEngineBase):

    def __init__(self, model: torch.nn.Module) -> None:
        super().__init__()
        self.model = model
        self.model_config = getattr(model, "config", None)

    def parameters(self) -> Iterable[Tuple[str, torch.Tensor]]:
        # state_dict (not named_parameters) so tied / buffer-held weights such as lm_head are visited too
        for name, tensor in self.model.state_dict().items():
            yield name, tensor
