"""Parameter offload (reference ``deepspeed/compile/passes/offload_parameters.py``): the ZeRO-3 slices of a module live
in pinned host memory and are copied in right before their gather. Converts the slices of an already-partitioned
module in place."""
from __future__ import annotations

import torch

from ...runtime.zero.partition_parameters import is_zero_param

NAME = "offload_parameters"


def offload_parameter_slices(module: torch.nn.Module, pin_memory: bool = True) -> int:
    """Move every ZeRO-3 slice to the host; returns the bytes moved. Gathers copy the slice back (``ds_tensor.to``)."""
    moved = 0
    for p in module.parameters():
        if not is_zero_param(p) or p.ds_tensor.device.type == "cpu":
            continue
        host = p.ds_tensor.detach().to("cpu")
        if pin_memory and torch.cuda.is_available():
            host = host.pin_memory()
        p.ds_tensor = host
        moved += host.numel() * host.element_size()
    return moved
