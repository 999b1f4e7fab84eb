"""``init_z1`` (reference ``deepspeed/compile/init_z1.py``): ZeRO-1/2 needs no parameter schedule; the gradient
reduce-scatter is issued per bucket from gradient hooks. Optimizer-state offload is the one pass that applies."""
from __future__ import annotations

import torch

from .config import CompileConfig
from .passes import offload_adam_states, zero1_compile


def init_z1(engine, backend=None, compile_config=None, compile_kwargs=None, schedule=None, use_z2: bool = False):
    cfg = CompileConfig(**{**(engine._config.compile_config or {}), "deepcompile": True})
    zero1_compile.add_z1_reduce(engine.optimizer)

    def run():
        report = {}
        if cfg.offload_opt_states:
            dev = next((p.device for p in engine.module.parameters()), torch.device("cpu"))
            mover = offload_adam_states.move_opt_states(engine.optimizer, dev, cfg.double_buffer)
            mover.offload()
            report["adam_state_bytes"] = mover._nbytes()
        return report

    return run
