"""Schedule compiler: runs the passes over the recorded module trace and applies the result to a live ZeRO engine.

Counterpart of the reference's DeepCompile back end (``deepspeed/compile/backend.py`` ``make_backend`` /
``launch_compile_passes``, ``init_z1.py``, ``init_z3.py``, native part ``csrc/compile/*``). There, ``torch.compile``
captures FX graphs, passes rewrite them at given training steps, and custom ops do the collectives. Here nothing is
traced or code-generated: the ZeRO-3 hooks record the module order of one profiled iteration, the passes compute a
schedule on that order (which gathers to launch when, which parameters to keep resident, when optimizer states
move), and the schedule is written back into the hook installer and the optimizer. Collectives stay NCCL calls on
their own streams; launch-bound regions are CUDA-graphed elsewhere."""
from __future__ import annotations

import time
from typing import Callable, Dict, List, Optional

import torch

from .. import comm as dist
from ..runtime.zero.partition_parameters import all_gather_params
from ..utils.logging import log_dist
from .config import CompileConfig
from .passes import offload_activation as _oa
from .passes import offload_adam_states as _oas
from .passes import offload_parameters as _op
from .passes import prefetch as _pf
from .passes import selective_gather as _sg
from .trace import ModuleTrace

# name -> pass(ctx) ; user passes can be added with register_compile_pass
opt_passes: Dict[str, Callable] = {}
# (pass name, seconds) of every pass run so far - what ``engine.get_compile_time()`` returns (ref: compile/backend.py)
opt_pass_times: List[tuple] = []


def register_compile_pass(name: str, fn: Callable) -> None:
    opt_passes[name] = fn


class CompileContext:
    """What a pass sees: the trace, the hook installer, the optimizer wrapper, the config, and a report dict."""

    def __init__(self, offload, optimizer, config: CompileConfig, group=None) -> None:
        self.offload, self.optimizer, self.config, self.group = offload, optimizer, config, group
        self.trace = ModuleTrace.from_offload(offload)
        self.world_size = dist.get_world_size(group=group) if dist.is_initialized() else 1
        self.report: Dict[str, object] = {"events": len(self.trace.events)}


def _pass_selective_gather(ctx: CompileContext) -> None:
    budget = ctx.config.selective_gather_budget_bytes
    if budget < 0:
        budget = 0
        if torch.cuda.is_available():
            free, _ = torch.cuda.mem_get_info()
            # what the allocator still holds from the profiled step is reusable; leave 20 % head-room
            budget = int(0.8 * max(0, free - (torch.cuda.max_memory_allocated() - torch.cuda.memory_allocated())))
    chosen = _sg.selective_gather(ctx.trace, budget, ctx.world_size)
    for p in chosen:
        p.ds_persist = True
    all_gather_params(chosen)
    ctx.report["persistent_params"] = len(chosen)
    ctx.report["persistent_bytes"] = sum(p.ds_numel * p.element_size() for p in chosen)


def _pass_prefetch(ctx: CompileContext) -> None:
    off = ctx.offload
    off.prefetch_bucket_size = ctx.config.prefetch_bucket_size
    off.prefetch_plan = _pf.schedule_prefetch([(e.params, e.phase) for e in ctx.trace.events], off.prefetch_bucket_size)
    ctx.report["prefetch_launch_points"] = sum(1 for x in off.prefetch_plan if x)


def _pass_offload_adam(ctx: CompileContext) -> None:
    if not ctx.config.offload_opt_states or ctx.optimizer is None:
        return
    dev = next((p.device for p in ctx.offload.module.parameters()), torch.device("cpu"))
    mover = _oas.move_opt_states(ctx.optimizer, dev, ctx.config.double_buffer)
    mover.offload()
    ctx.report["adam_state_bytes"] = mover._nbytes()


def _pass_offload_activation(ctx: CompileContext) -> None:
    if ctx.config.offload_activation:
        _oa.offload_activation(True)
        ctx.report["offload_activation"] = True


def _pass_offload_parameters(ctx: CompileContext) -> None:
    if ctx.config.offload_parameters:
        ctx.report["offloaded_param_bytes"] = _op.offload_parameter_slices(ctx.offload.module)


register_compile_pass(_sg.NAME, _pass_selective_gather)
register_compile_pass(_pf.NAME, _pass_prefetch)
register_compile_pass(_oas.NAME, _pass_offload_adam)
register_compile_pass(_oa.NAME, _pass_offload_activation)
register_compile_pass(_op.NAME, _pass_offload_parameters)

DEFAULT_Z3_PASSES = [_sg.NAME, _pf.NAME, _oas.NAME, _oa.NAME, _op.NAME]


def launch_compile_passes(offload, optimizer=None, config: Optional[CompileConfig] = None, passes: Optional[List[str]] = None,
                          group=None) -> Dict[str, object]:
    """Run ``passes`` (default: all ZeRO-3 ones, selective gather first so prefetch skips what became resident)."""
    ctx = CompileContext(offload, optimizer, config or CompileConfig(deepcompile=True), group)
    for name in passes or DEFAULT_Z3_PASSES:
        t0 = time.perf_counter()
        opt_passes[name](ctx)
        opt_pass_times.append((name, time.perf_counter() - t0))
    if ctx.config.debug_log:
        log_dist("schedule compiler:\n" + ctx.trace.summary() + f"\n{ctx.report}", [0])
    return ctx.report


def make_backend(engine, passes: Optional[List[str]] = None):
    """A callable that compiles the schedule of ``engine`` once its trace is complete (call it after the first step)."""

    def run():
        opt = engine.optimizer
        return launch_compile_passes(opt.parameter_offload, opt, CompileConfig(**{**engine._config.compile_config,
                                                                                  "deepcompile": True}), passes)

    return run
