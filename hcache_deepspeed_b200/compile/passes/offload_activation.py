"""Activation offload (reference ``deepspeed/compile/passes/offload_activation.py``): checkpointed activations go to
host memory between forward and backward. The mechanism is the activation-checkpointing module's ``cpu_checkpointing``;
this pass only switches it on for an engine."""
from __future__ import annotations

NAME = "offload_activation"


def offload_activation(enabled: bool = True) -> None:
    from ...runtime.activation_checkpointing import checkpointing
    checkpointing.configure(getattr(checkpointing, "mpu", None), checkpoint_in_cpu=enabled)
