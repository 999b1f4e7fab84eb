"""``init_z3`` (reference ``deepspeed/compile/init_z3.py``): prepare a ZeRO-3 engine for schedule compilation."""
from __future__ import annotations

from .backend import make_backend


def init_z3(engine, backend=None, compile_config=None, compile_kwargs=None, schedule=None):
    """Returns the callable that compiles the schedule; the engine calls it after its first (profiled) step."""
    opt = engine.optimizer
    if not hasattr(opt, "parameter_offload"):
        raise RuntimeError("init_z3 needs a ZeRO stage-3 optimizer")
    return make_backend(engine, schedule)
