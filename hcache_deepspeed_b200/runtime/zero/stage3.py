"""ZeRO stage 3: parameters, gradients and optimizer state are all partitioned over the data-parallel group.

* ``DeepSpeedZeRoOffload`` installs module hooks: before a (sub)module's forward / backward its own parameters are
  all-gathered, afterwards they are released again (parameters below ``param_persistence_threshold`` stay resident).
  After a module's backward the full gradients are reduce-scattered into fp32 accumulators that cover only the
  owned slice.
* ``DeepSpeedZeroOptimizer_Stage3`` keeps an fp32 master copy of each owned slice (device or host memory), runs the
  wrapped optimizer on those and writes the result back into the 16-bit slices; no parameter all-gather happens at
  step time — the next forward fetches what it needs.

Reference: ``deepspeed/runtime/zero/stage3.py`` (:73 optimizer, step :2040), ``parameter_offload.py``
(DeepSpeedZeRoOffload :87, hooks :239-437), ``partitioned_param_coordinator.py`` (fetch/release/prefetch order).
"""
from __future__ import annotations

import math
from collections import OrderedDict, UserDict
from enum import Enum
from typing import List, Optional

import torch

from ... import comm as dist
from ...utils.logging import logger
from ..fp16.loss_scaler import CreateLossScaler
from .optimizer_api import ZeroOptimizerAPI
from .partition_parameters import (Init, ZeroParamStatus, _partition_param, all_gather_params, is_zero_param, read_partition,
                                   release_partition, set_param_swapper, write_partition)


class ZeRoTraceMode(Enum):
    """Where the fetch-order trace stands: being recorded, complete (prefetch runs from it), or invalidated because the
    model took another path than the recorded one."""
    RECORD = 1
    COMPLETE = 2
    INVALID = 3


class InflightParamRegistry(UserDict):
    """Parameter -> handle of its gather in flight; only INFLIGHT parameters may be entered, and only once."""

    def __setitem__(self, param, handle) -> None:
        if param in self.data:
            raise RuntimeError(f"{param.ds_summary()} already in registry")
        if param.ds_status != ZeroParamStatus.INFLIGHT:
            raise RuntimeError(f"attempted to add non-inflight parameter to registry {param.ds_summary()}")
        self.data[param] = handle


def get_all_parameters(sub_module: torch.nn.Module, recurse: bool = False):
    """(name, parameter) of a module's own parameters followed by the external ones registered on it."""
    import itertools
    ext = getattr(sub_module, "_external_params", {})
    return itertools.chain(sub_module.named_parameters(recurse=recurse), ((f"external_{i}", p) for i, p in enumerate(ext.values())))


def iter_params(module: torch.nn.Module, recurse: bool = False):
    return (p for _, p in get_all_parameters(module, recurse))


FWD_MODULE_STACK: list = []      # modules whose forward is running (innermost last), kept by the stage-3 hooks


class ZeROOrderedDict(OrderedDict):
    """Stands in for a module's ``_parameters`` under ZeRO-3: reading a partitioned parameter of ANOTHER module while a
    forward is running (``self.other.weight`` inside ``forward``, tied embeddings used through their owner, ...) gathers
    it on the spot and registers it as an external parameter of the module that is executing, so it is released with
    that module and fetched by its hooks from then on (ref: parameter_offload.py:40-75)."""

    def __init__(self, parent_module, *args, **kwargs):
        super().__init__(*args, **kwargs)
        self._parent_module = parent_module

    def __reduce__(self):
        r0, _, *rest = super().__reduce__()
        return (r0, (self._parent_module, )) + tuple(rest)

    def __getitem__(self, key):
        param = super().__getitem__(key)
        if param is None or not FWD_MODULE_STACK:
            return param
        if getattr(param, "ds_status", None) == ZeroParamStatus.NOT_AVAILABLE:
            user = FWD_MODULE_STACK[-1]
            from .partition_parameters import register_external_parameter
            register_external_parameter(user, param)
            all_gather_params([param])
            param.ds_active_sub_modules.add(id(user))
        return param


class DeepSpeedZeRoOffload:
    """Fetch / release orchestration through module hooks."""

    def __init__(self, module: torch.nn.Module, timers=None, ds_config=None, overlap_comm: bool = True,
                 prefetch_bucket_size: int = 0, max_reuse_distance: int = 0, max_live_parameters: int = 0,
                 param_persistence_threshold: int = 100000, model_persistence_threshold: int = 2 ** 62,
                 dp_process_group=None, offload_param_config=None, mpu=None, zero_param_parallel_group=None,
                 zero_quantized_weights=False, zero_quantized_nontrainable_weights=False,
                 zero_module_granularity_threshold=0, log_trace_cache_warnings=False, dtype=None):
        self.module = module
        self.dp_process_group = dp_process_group
        remote = None
        pin = False
        self.param_swapper = None
        device_name = str(getattr(offload_param_config, "device", "none")).split(".")[-1] if offload_param_config is not None else "none"
        if device_name in ("cpu", "nvme"):
            remote, pin = device_name, bool(getattr(offload_param_config, "pin_memory", False))
        if device_name == "nvme":
            # ZeRO-Infinity: partitions live in files under nvme_path, staged through a pool of pinned buffers
            from ..swap_tensor.partitioned_param_swapper import AsyncPartitionedParameterSwapper
            first = next((p for p in module.parameters() if p.dtype.is_floating_point), None)
            model_dtype = dtype if dtype is not None else (first.dtype if first is not None else torch.float32)
            aio = getattr(ds_config, "aio_config", None)
            self.param_swapper = AsyncPartitionedParameterSwapper(
                model_dtype=model_dtype, nvme_path=str(offload_param_config.nvme_path),
                buffer_count=int(getattr(offload_param_config, "buffer_count", 5)),
                buffer_size=int(getattr(offload_param_config, "buffer_size", 1e8)), aio_config=dict(aio) if aio else None)
            set_param_swapper(self.param_swapper)
        init = Init(data_parallel_group=dp_process_group, remote_device=remote, pin_memory=pin, dtype=dtype, enabled=True)
        init.persistence_threshold = param_persistence_threshold
        init._convert_module(module)
        self.zero_quantized_weights = bool(zero_quantized_weights)
        if zero_quantized_weights or zero_quantized_nontrainable_weights:
            for p in module.parameters():
                if is_zero_param(p) and (zero_quantized_weights or not p.requires_grad):
                    p.ds_quantized_gather = True
        self.hooks = []
        self.reduce_fn = None  # set by the optimizer: called with a list of params whose .grad is complete
        # trace-driven prefetch (reference: partitioned_param_coordinator.py record_module / fetch_sub_module): the first
        # iteration records the (module, phase) order; later iterations launch the gathers of the next modules, up to
        # ``prefetch_bucket_size`` elements, while the current module computes. A diverging order drops the trace.
        self.prefetch_bucket_size = int(prefetch_bucket_size)
        self.max_live_parameters = int(max_live_parameters) or 2 ** 62
        # a parameter whose next use (by the trace) is fewer than this many fetched elements away is not released in
        # between: the last layers' weights survive from forward into backward instead of being gathered twice
        self.max_reuse_distance = int(max_reuse_distance)
        self._uses = None                       # id(param) -> sorted trace positions; built with the trace
        self._cum_numel: List[int] = []         # elements fetched by trace events [0, i)
        self.trace: List[tuple] = []            # [(module, phase)] of one forward(+backward)
        self.trace_complete = False
        self.prefetch_plan: Optional[List[List[torch.nn.Parameter]]] = None   # set by compile passes; else derived
        self.event_times: List[float] = []      # perf_counter at each recorded event (profile for the passes)
        self._recording: List[tuple] = []
        self._pos = 0
        self.stats = {"prefetched": 0, "trace_invalidations": 0, "kept_for_reuse": 0}
        self._install(module)

    @staticmethod
    def _own_params(m: torch.nn.Module) -> List[torch.nn.Parameter]:
        # a module flagged by utils.z3_leaf_module owns its whole subtree (children carry no hooks of their own)
        ps = [p for p in m.parameters(recurse=bool(getattr(m, "_z3_leaf", False))) if is_zero_param(p)]
        ps += [p for p in getattr(m, "_external_params", {}).values() if is_zero_param(p)]
        return ps

    def _install(self, root: torch.nn.Module):
        self.hooks.append(root.register_forward_pre_hook(self._begin_iteration))
        for m in root.modules():       # external-parameter detection (see ZeROOrderedDict)
            if not isinstance(m._parameters, ZeROOrderedDict):
                m._parameters = ZeROOrderedDict(m, m._parameters)
        under_leaf = set()
        for m in root.modules():
            if id(m) in under_leaf:
                continue
            if getattr(m, "_z3_leaf", False):
                under_leaf.update(id(c) for c in m.modules() if c is not m)
                for p in m.parameters():
                    p.ds_z3_leaf_module = m
            self.hooks.append(m.register_forward_pre_hook(self._push_fwd))
            self.hooks.append(m.register_forward_hook(self._pop_fwd, always_call=True))
            if not self._own_params(m):
                continue
            m._z3_has_param_hooks = True
            self.hooks.append(m.register_forward_pre_hook(self._pre_forward))
            self.hooks.append(m.register_forward_hook(self._post_forward))
            self.hooks.append(m.register_full_backward_pre_hook(self._pre_backward))
        # a parameter's gradient is final when its AccumulateGrad node has run — which can be after the owning module's
        # backward hook — so reduction + release hang on the per-parameter post-accumulate hook
        for p in root.parameters():
            if is_zero_param(p) and p.requires_grad:
                self.hooks.append(p.register_post_accumulate_grad_hook(self._post_accumulate))

    # ---- trace + prefetch ----------------------------------------------------------------------------------------
    def _begin_iteration(self, m, args):
        if not self.trace_complete and self._recording:
            self.trace, self.trace_complete = self._recording, True
            self.prefetch_plan = None
        self._recording, self._pos = [], 0
        if not self.trace_complete:
            self.event_times = []

    def reset_trace(self):
        self.trace, self.trace_complete, self.prefetch_plan, self._recording, self._pos = [], False, None, [], 0
        self._uses = None

    def _index_uses(self):
        uses, cum, total = {}, [0], 0
        for i, (m, _) in enumerate(self.trace):
            for p in self._own_params(m):
                uses.setdefault(id(p), []).append(i)
                if not p.ds_persist:
                    total += p.ds_numel
            cum.append(total)
        self._uses, self._cum_numel = uses, cum

    def _reused_soon(self, p) -> bool:
        """True when the recorded order says ``p`` is needed again within ``max_reuse_distance`` fetched elements of the
        current position (and keeping it does not break the live-parameter cap)."""
        if self.max_reuse_distance <= 0 or not self.trace_complete or not self.trace:
            return False
        if self._uses is None:
            self._index_uses()
        import bisect
        where = self._uses.get(id(p))
        if not where:
            return False
        k = bisect.bisect_left(where, self._pos)
        if k == len(where):
            return False                        # not used again before the iteration ends
        if self.trace[where[k]][1] == "bwd" and not torch.is_grad_enabled():
            return False                        # evaluation: the recorded backward will not come
        if self._cum_numel[where[k]] - self._cum_numel[min(self._pos, len(self.trace))] > self.max_reuse_distance:
            return False
        return self._live_numel() <= self.max_live_parameters

    def _live_numel(self) -> int:
        zp = self.__dict__.get("_zero_params")
        if zp is None:
            zp = self._zero_params = [p for p in self.module.parameters() if is_zero_param(p)]
            self._total_numel = sum(p.ds_numel for p in zp)
        if self.max_live_parameters >= self._total_numel:
            return 0                # the cap can never bind: skip the scan
        return sum(p.ds_numel for p in zp if not p.ds_persist and p.ds_status != ZeroParamStatus.NOT_AVAILABLE)

    def _derive_plan(self) -> List[List[torch.nn.Parameter]]:
        from ...compile.passes.prefetch import schedule_prefetch
        return schedule_prefetch([(self._own_params(m), ph) for m, ph in self.trace], self.prefetch_bucket_size)

    def _on_event(self, m, phase: str):
        if not self.trace_complete:
            import time
            self._recording.append((m, phase))
            self.event_times.append(time.perf_counter())
            return
        pos = self._pos
        if pos >= len(self.trace) or self.trace[pos][0] is not m or self.trace[pos][1] != phase:
            # the model took another path this time: stop trusting the trace and record again from the next iteration
            self.stats["trace_invalidations"] += 1
            self.reset_trace()
            return
        self._pos = pos + 1
        if self.prefetch_bucket_size <= 0:
            return
        if self.prefetch_plan is None:
            self.prefetch_plan = self._derive_plan()
        ahead = self.prefetch_plan[pos]
        if not ahead:
            return
        if phase == "fwd" and pos + 1 < len(self.trace) and self.trace[pos + 1][1] == "bwd" and not torch.is_grad_enabled():
            return       # evaluation: there is no backward to prefetch for
        todo = [p for p in ahead if p.ds_status == ZeroParamStatus.NOT_AVAILABLE]
        if todo and self._live_numel() + sum(p.ds_numel for p in todo) <= self.max_live_parameters:
            all_gather_params(todo, async_op=True)
            self.stats["prefetched"] += len(todo)

    @staticmethod
    def _push_fwd(m, args):
        FWD_MODULE_STACK.append(m)

    def _pop_fwd(self, m, args, out):
        if FWD_MODULE_STACK and FWD_MODULE_STACK[-1] is m:
            FWD_MODULE_STACK.pop()
        ext = getattr(m, "_external_params", None)
        if ext:
            # a module that had no parameters of its own when the hooks were installed gets none of the fetch / release
            # hooks: release what it pulled in through the getter here
            if not getattr(m, "_z3_has_param_hooks", False):
                self._release(m)

    def _pre_forward(self, m, args):
        ps = self._own_params(m)
        all_gather_params(ps)
        for p in ps:
            p.ds_active_sub_modules.add(id(m))
        self._on_event(m, "fwd")

    def _post_forward(self, m, args, out):
        self._release(m)

    def _pre_backward(self, m, grad_out):
        ps = self._own_params(m)
        all_gather_params(ps)
        for p in ps:
            p.ds_active_sub_modules.add(id(m))
        self._on_event(m, "bwd")

    def _post_accumulate(self, p):
        if self.reduce_fn is not None:
            self.reduce_fn([p])
        p.ds_active_sub_modules.clear()
        if not p.ds_persist:
            _partition_param(p, has_been_updated=False)

    def _release(self, m):
        for p in self._own_params(m):
            p.ds_active_sub_modules.discard(id(m))
            if not p.ds_active_sub_modules and not p.ds_persist:
                if self._reused_soon(p):
                    self.stats["kept_for_reuse"] += 1
                    continue
                _partition_param(p, has_been_updated=False)

    def destroy(self):
        for h in self.hooks:
            h.remove()
        self.hooks = []

    def get_param_coordinator(self, training=True):
        return self

    # ---- trace state under the coordinator's names (ref: partitioned_param_coordinator.py:43-50, 150-175) -------------
    def get_trace_mode(self) -> ZeRoTraceMode:
        if self.trace_complete:
            return ZeRoTraceMode.COMPLETE
        return ZeRoTraceMode.INVALID if self.stats.get("trace_invalidations", 0) and not self._recording else ZeRoTraceMode.RECORD

    def is_complete_trace(self) -> bool:
        return self.get_trace_mode() == ZeRoTraceMode.COMPLETE

    def is_record_trace(self) -> bool:
        return self.get_trace_mode() == ZeRoTraceMode.RECORD

    def is_invalid_trace(self) -> bool:
        return self.get_trace_mode() == ZeRoTraceMode.INVALID

    # ---- the reference's public names for the same steps (parameter_offload.py:200-520) --------------------------------
    def setup_zero_stage3_hooks(self) -> None:
        if not self.hooks:
            self._install(self.module)

    def partition_all_parameters(self) -> None:
        """Release every gathered parameter that is not marked persistent."""
        for p in self.module.parameters():
            if is_zero_param(p) and not p.ds_persist and p.ds_status != ZeroParamStatus.NOT_AVAILABLE:
                p.ds_active_sub_modules.clear()
                _partition_param(p)

    def empty_partition_cache(self) -> None:
        self.partition_all_parameters()
        self._pos = 0

    def mark_persistent_parameters(self, param_threshold: int, model_threshold: int):
        """Keep parameters of at most ``param_threshold`` elements resident, smallest-first in module order until
        ``model_threshold`` elements in total are resident; returns the parameters newly made persistent (gathered)."""
        budget = model_threshold - sum(p.ds_numel for p in self.module.parameters() if is_zero_param(p) and p.ds_persist)
        newly = []
        for p in self.module.parameters():
            if not is_zero_param(p) or p.ds_persist or p.ds_numel > param_threshold:
                continue
            if p.ds_numel > budget:
                continue
            budget -= p.ds_numel
            p.ds_persist = True
            newly.append(p)
        if newly:
            all_gather_params(newly)
        return newly

    def pre_sub_module_forward_function(self, sub_module) -> None:
        self._pre_forward(sub_module, None)

    def post_sub_module_forward_function(self, sub_module) -> None:
        self._release(sub_module)

    def pre_sub_module_backward_function(self, sub_module) -> None:
        self._pre_backward(sub_module, None)

    def post_sub_module_backward_function(self, sub_module) -> None:
        self._release(sub_module)


class DeepSpeedZeroOptimizer_Stage3(ZeroOptimizerAPI):

    def __init__(self, module: torch.nn.Module, init_optimizer, param_names=None, timers=None, ds_config=None,
                 static_loss_scale: float = 1.0, dynamic_loss_scale: bool = False, dynamic_loss_args=None,
                 verbose: bool = True, contiguous_gradients: bool = True, reduce_bucket_size: int = 500000000,
                 prefetch_bucket_size: int = 50000000, max_reuse_distance: int = 1000000000,
                 max_live_parameters: int = 1000000000, param_persistence_threshold: int = 100000,
                 model_persistence_threshold: int = 2 ** 62, dp_process_group=None, reduce_scatter: bool = True,
                 overlap_comm: bool = False, offload_optimizer_config=None, offload_param_config=None,
                 sub_group_size: int = 1000000000000, offload_ratio: float = 0.0, mpu=None, clip_grad: float = 0.0,
                 gradient_accumulation_dtype=torch.float32, communication_data_type=torch.float16,
                 postscale_gradients: bool = True, gradient_predivide_factor: float = 1.0,
                 gradient_accumulation_steps: int = 1, elastic_checkpoint: bool = False, aio_config=None,
                 all2all_process_group=None, zero_hpz_partition_size: int = 1, zero_quantized_weights: bool = False,
                 zero_quantized_nontrainable_weights: bool = False, zero_module_granularity_threshold: int = 0,
                 zeropp_loco_param=None, log_trace_cache_warnings=False, **unused):
        self.module = module
        self.optimizer = init_optimizer
        self.zero_quantized_gradients = bool(unused.get("zero_quantized_gradients", False))
        # ZeRO++ LoCo: error feedback on the quantised gradient exchange ({"err_beta": .., "reset_T": ..}); None = plain qgZ
        self.zeropp_loco_param = zeropp_loco_param
        self._loco_err = {}                                    # parameter id -> (error buffer, calls since reset)
        self.qgz_stats = {"calls": 0, "bytes_sent": 0, "bytes_unquantized": 0}
        self.param_names = param_names or {}
        self.dp_process_group = dp_process_group
        self.partition_count = dist.get_world_size(group=dp_process_group) if dist.is_initialized() else 1
        self.rank = dist.get_rank(group=dp_process_group) if dist.is_initialized() else 0
        self.mpu = mpu
        self.clip_grad = clip_grad
        self.gradient_predivide_factor = gradient_predivide_factor
        self.gradient_accumulation_dtype = gradient_accumulation_dtype or torch.float32
        self.overflow = False
        self.skipped_steps = 0
        self._global_grad_norm = 0.0
        self.custom_loss_scaler = False
        self.external_loss_scale = None
        off = offload_optimizer_config
        off_dev = str(getattr(off, "device", "none")).split(".")[-1] if off is not None else "none"
        self.offload_optimizer = off_dev in ("cpu", "nvme")
        self.optimizer_swapper = None
        dtype = next((p.dtype for p in module.parameters()), torch.float32)
        self.parameter_offload = DeepSpeedZeRoOffload(module, dp_process_group=dp_process_group,
                                                      prefetch_bucket_size=prefetch_bucket_size,
                                                      max_live_parameters=max_live_parameters,
                                                      max_reuse_distance=max_reuse_distance,
                                                      param_persistence_threshold=param_persistence_threshold,
                                                      offload_param_config=offload_param_config, mpu=mpu, dtype=None,
                                                      ds_config=ds_config, zero_quantized_weights=zero_quantized_weights,
                                                      zero_quantized_nontrainable_weights=zero_quantized_nontrainable_weights)
        self.parameter_offload.reduce_fn = self._reduce_ready_grads
        self.dtype = dtype
        self.loss_scaler = CreateLossScaler(dtype=dtype, static_loss_scale=static_loss_scale,
                                            dynamic_scaling=dynamic_loss_scale, dynamic_loss_args=dynamic_loss_args)
        self.dynamic_loss_scale = self.loss_scaler.dynamic

        # fp32 master slices, one flat buffer per param group
        self.fp16_groups: List[List[torch.nn.Parameter]] = []
        self.fp32_partitioned_groups_flat: List[torch.Tensor] = []
        self.grad_partitions = {}
        for group in self.optimizer.param_groups:
            params = [p for p in group["params"] if p.requires_grad]
            self.fp16_groups.append(params)
            if not params:
                self.fp32_partitioned_groups_flat.append(torch.empty(0, requires_grad=True))
                group["params"] = []
                continue
            dev = torch.device("cpu") if self.offload_optimizer else params[0].ds_device
            pieces = []
            for p in params:        # NVMe-resident slices are staged one at a time
                pieces.append(read_partition(p).detach().float().view(-1).to(dev).clone())
                release_partition(p)
            master = torch.cat(pieces)
            master.requires_grad = True
            self.fp32_partitioned_groups_flat.append(master)
            group["params"] = [master]
        if off_dev == "nvme":
            # ZeRO-Infinity: master weights + optimizer state of each group live in files between steps
            from ..swap_tensor import PartitionedOptimizerSwapper, PipelinedOptimizerSwapper
            piped = bool(getattr(off, "pipeline_read", False) or getattr(off, "pipeline_write", False))
            masters = [m for m in self.fp32_partitioned_groups_flat if m.numel()]
            self.optimizer_swapper = (PipelinedOptimizerSwapper if piped else PartitionedOptimizerSwapper)(
                off, aio_config if isinstance(aio_config, dict) else None, str(off.nvme_path), self.optimizer,
                max((m.numel() for m in masters), default=0), "cpu", torch.float32, timers)
            self.optimizer_swapper.initialize_parameters(masters, masters)
        if verbose and self.rank == 0:
            n = sum(p.ds_numel for g in self.fp16_groups for p in g)
            logger.info(f"ZeRO stage 3: {n} parameters over {self.partition_count} ranks, optimizer offload="
                        f"{self.offload_optimizer}")

    # ---------------------------------------------------------------------------------------------
    @property
    def param_groups(self):
        return self.optimizer.param_groups

    @property
    def state(self):
        return self.optimizer.state

    @property
    def loss_scale(self):
        return self.external_loss_scale if self.custom_loss_scaler else self.loss_scaler.cur_scale

    cur_scale = loss_scale

    def zero_grad(self, set_to_none=True):
        for g in self.fp16_groups:
            for p in g:
                p.grad = None

    @torch.no_grad()
    def _reduce_ready_grads(self, params):
        """Average ``p.grad`` over the group, keep this rank's slice in the fp32 accumulator, drop the full gradient."""
        ws = self.partition_count
        for p in params:
            if p.grad is None:
                continue
            psize = p.ds_partition_numel
            g = p.grad.detach().reshape(-1)
            buf = torch.zeros(psize * ws, dtype=g.dtype, device=g.device)
            buf[:g.numel()].copy_(g)
            if ws > 1:
                if self.gradient_predivide_factor != 1.0:
                    buf.div_(self.gradient_predivide_factor)
                if self.zero_quantized_gradients and psize >= 256:
                    out = self._quantized_reduce_scatter(buf, psize, ws, key=p.ds_id).to(buf.dtype)
                    out.mul_(self.gradient_predivide_factor)
                else:
                    out = torch.empty(psize, dtype=buf.dtype, device=buf.device)
                    dist.reduce_scatter_tensor(out, buf, group=self.dp_process_group)
                    out.div_(ws / self.gradient_predivide_factor)
            else:
                out = buf
            out = out.to(self.gradient_accumulation_dtype)
            if p.ds_id in self.grad_partitions:
                self.grad_partitions[p.ds_id].add_(out)
            else:
                self.grad_partitions[p.ds_id] = out.clone()
            p.grad = None

    def _quantized_reduce_scatter(self, buf: torch.Tensor, psize: int, ws: int, bits: int = 8, key=None) -> torch.Tensor:
        """ZeRO++ qgZ on one flat group: every rank quantises the ``ws`` slices of its gradient (one scale per 256 values),
        an all-to-all delivers slice r of every rank to rank r, which dequantises and averages them (ref:
        coalesced_collectives.py all_to_all_quant_reduce; the two-hop intra / inter-node form collapses on one NVSwitch node)."""
        from ...ops.quantizer import dequantize_int, quantize_int
        pp = (psize + 255) // 256 * 256
        send = torch.zeros(ws, pp, dtype=torch.float32, device=buf.device)
        send[:, :psize] = buf.view(ws, psize).float()
        loco = self.zeropp_loco_param if key is not None else None
        if loco:
            # error feedback: what quantisation lost last time is (partly) added back before quantising again
            err, age = self._loco_err.get(key, (None, 0))
            if err is None or age >= int(loco.get("reset_T", 1024)) or err.shape != send.shape:
                err, age = torch.zeros_like(send), 0
            send = send + float(loco.get("err_beta", 0.8)) * err
        codes, params = quantize_int(send.reshape(-1), ws * pp // 256, bits, symmetric=True)
        if loco:
            self._loco_err[key] = (send - dequantize_int(codes, params, bits, torch.float32).reshape(ws, pp), age + 1)
        rc, rp = torch.empty_like(codes), torch.empty_like(params)
        dist.all_to_all_single(rc.view(-1), codes.reshape(-1).contiguous(), group=self.dp_process_group)
        dist.all_to_all_single(rp.view(-1), params.reshape(-1).contiguous(), group=self.dp_process_group)
        deq = dequantize_int(rc, rp, bits, torch.float32).reshape(ws, pp)
        self.qgz_stats["calls"] += 1
        self.qgz_stats["bytes_sent"] += codes.numel() * codes.element_size() + params.numel() * params.element_size()
        self.qgz_stats["bytes_unquantized"] += buf.numel() * buf.element_size()
        return deq.mean(dim=0)[:psize]

    def backward(self, loss, retain_graph: bool = False):
        if self.custom_loss_scaler:
            (loss * self.external_loss_scale).backward(retain_graph=retain_graph)
        else:
            self.loss_scaler.backward(loss.float(), retain_graph=retain_graph)
        # sweep: gradients that materialised after their module's hook, and parameters still gathered
        self._reduce_ready_grads([p for g in self.fp16_groups for p in g if p.grad is not None])
        for g in self.fp16_groups:
            for p in g:
                if p.ds_status == ZeroParamStatus.AVAILABLE and not p.ds_persist:
                    p.ds_active_sub_modules.clear()
                    _partition_param(p)

    def _all_reduce_scalar(self, v: float, op) -> float:
        if self.partition_count == 1 and self.mpu is None:
            return v
        dev = next((p.ds_device for g in self.fp16_groups for p in g), torch.device("cpu"))
        t = torch.tensor([float(v)], dtype=torch.float32, device=dev)
        if self.partition_count > 1:
            dist.all_reduce(t, op=op, group=self.dp_process_group)
        if self.mpu is not None:
            dist.all_reduce(t, op=op, group=self.mpu.get_model_parallel_group())
        return float(t[0])

    def get_global_grad_norm(self):
        return self._global_grad_norm

    @torch.no_grad()
    def step(self, closure=None):
        local_bad = 0.0
        sq = 0.0
        for g in self.grad_partitions.values():
            s = float(g.float().sum())
            if s in (float("inf"), -float("inf")) or s != s:
                local_bad = 1.0
            sq += float(g.float().norm(2)) ** 2
        self.overflow = self._all_reduce_scalar(local_bad, dist.ReduceOp.MAX) > 0
        prev = self.loss_scale
        if not self.custom_loss_scaler:
            self.loss_scaler.update_scale(self.overflow)
        if self.overflow:
            self.skipped_steps += 1
            self.grad_partitions = {}
            if self.rank == 0:
                logger.info(f"[deepspeed] OVERFLOW! Skipping step. Attempted loss scale: {prev}, reducing to {self.loss_scale}")
            return
        scaled_norm = math.sqrt(self._all_reduce_scalar(sq, dist.ReduceOp.SUM))
        self._global_grad_norm = scaled_norm / prev
        combined = prev
        if self.clip_grad > 0.0:
            clip = (scaled_norm / prev + 1e-6) / self.clip_grad
            if clip > 1:
                combined = clip * prev
        if self.optimizer_swapper is not None:
            return self._step_swapped(combined)
        for gi, params in enumerate(self.fp16_groups):
            master = self.fp32_partitioned_groups_flat[gi]
            if not params:
                continue
            grads = []
            for p in params:
                g = self.grad_partitions.get(p.ds_id)
                grads.append(g.float().to(master.device) if g is not None else
                             torch.zeros(p.ds_partition_numel, dtype=torch.float32, device=master.device))
            master.grad = torch.cat(grads).div_(combined)
        self.grad_partitions = {}
        self.optimizer.step()
        for gi, params in enumerate(self.fp16_groups):
            master = self.fp32_partitioned_groups_flat[gi]
            master.grad = None
            off = 0
            for p in params:
                n = p.ds_partition_numel
                write_partition(p, master.detach().narrow(0, off, n))
                off += n
                if p.ds_status == ZeroParamStatus.AVAILABLE:  # persistent parameters: refresh the resident full copy
                    p.ds_status = ZeroParamStatus.NOT_AVAILABLE
                    all_gather_params([p])

    def _write_back(self, gi: int):
        master, off = self.fp32_partitioned_groups_flat[gi], 0
        master.grad = None
        for p in self.fp16_groups[gi]:
            n = p.ds_partition_numel
            write_partition(p, master.detach().narrow(0, off, n))
            off += n
            if p.ds_status == ZeroParamStatus.AVAILABLE:  # persistent parameters: refresh the resident full copy
                p.ds_status = ZeroParamStatus.NOT_AVAILABLE
                all_gather_params([p])

    def _step_swapped(self, combined: float):
        """Optimizer step with the NVMe tier: group by group, swap in (the next group's read already in flight) ->
        update -> write back the low-precision slices -> swap out (asynchronously with a pipelined swapper)."""
        sw = self.optimizer_swapper
        live = [gi for gi, params in enumerate(self.fp16_groups) if params]
        all_groups = self.optimizer.param_groups
        try:
            for k, gi in enumerate(live):
                master = self.fp32_partitioned_groups_flat[gi]
                nxt = self.fp32_partitioned_groups_flat[live[k + 1]] if k + 1 < len(live) else None
                sw.swap_in_optimizer_state(master, async_parameter=nxt)
                grads = []
                for p in self.fp16_groups[gi]:
                    g = self.grad_partitions.get(p.ds_id)
                    grads.append(g.float().to(master.device) if g is not None else
                                 torch.zeros(p.ds_partition_numel, dtype=torch.float32, device=master.device))
                master.grad = torch.cat(grads).div_(combined)
                self.optimizer.param_groups = [all_groups[gi]]
                self.optimizer.step()
                self._write_back(gi)
                sw.swap_out_optimizer_state(master, async_swap=True)
        finally:
            self.optimizer.param_groups = all_groups
        self.grad_partitions = {}

    class _Resident:
        """Everything swapped in for the duration (checkpoint save / load, consolidation)."""

        def __init__(self, opt):
            self.opt = opt

        def __enter__(self):
            sw = self.opt.optimizer_swapper
            if sw is not None:
                sw.swap_in_all([m for m in self.opt.fp32_partitioned_groups_flat if id(m) in sw.records])

        def __exit__(self, *exc):
            sw = self.opt.optimizer_swapper
            if sw is not None:
                sw.swap_out_all([m for m in self.opt.fp32_partitioned_groups_flat if id(m) in sw.records])
                sw.purge_state()
            return False

    # ---------------------------------------------------------------------------------------------
    def state_dict(self):
        with self._Resident(self):
            sd = self._state_dict()
            if self.optimizer_swapper is not None:     # the live state tensors are about to be swapped out again
                import copy
                sd["optimizer_state_dict"] = copy.deepcopy(sd["optimizer_state_dict"])
            return sd

    def _state_dict(self):
        return {"loss_scaler": self.loss_scaler, "dynamic_loss_scale": self.dynamic_loss_scale, "overflow": self.overflow,
                "partition_count": self.partition_count, "zero_stage": 3,
                "optimizer_state_dict": self.optimizer.state_dict(),
                "fp32_flat_groups": [m.detach().clone() for m in self.fp32_partitioned_groups_flat],
                "param_shapes": self._param_shapes()}

    def _param_shapes(self):
        return [{self.param_names.get(p, f"group{gi}_param{i}"): torch.Size(p.ds_shape) for i, p in enumerate(params)}
                for gi, params in enumerate(self.fp16_groups)]

    def load_state_dict(self, *args, **kw):
        with self._Resident(self):
            return self._load_state_dict(*args, **kw)

    def _load_state_dict(self, state_dict_list, load_optimizer_states=True, load_from_fp32_weights=False,
                         checkpoint_folder=None, load_serial=None, param_shapes=None):
        sd = state_dict_list[self.rank] if isinstance(state_dict_list, (list, tuple)) else state_dict_list
        if sd.get("partition_count", self.partition_count) != self.partition_count:
            raise RuntimeError("ZeRO-3 checkpoints can only be loaded with the data-parallel size they were saved with")
        self.loss_scaler = sd.get("loss_scaler", self.loss_scaler)
        self.overflow = sd.get("overflow", False)
        if load_optimizer_states:
            self.optimizer.load_state_dict(sd["optimizer_state_dict"])
        for cur, saved in zip(self.fp32_partitioned_groups_flat, sd["fp32_flat_groups"]):
            cur.data.copy_(saved.data.to(cur.device))
        for gi, params in enumerate(self.fp16_groups):
            master, off = self.fp32_partitioned_groups_flat[gi], 0
            for p in params:
                n = p.ds_partition_numel
                write_partition(p, master.detach().narrow(0, off, n))
                off += n
                if p.ds_status == ZeroParamStatus.AVAILABLE:
                    p.ds_status = ZeroParamStatus.NOT_AVAILABLE
                    all_gather_params([p])

    # ---- hooks of ZeroOptimizerAPI + the stage-3 part of the reference's public surface ---------------------------------
    def _api_master_tensors(self):
        return self.fp32_partitioned_groups_flat

    def _api_local_grads(self):
        return list(self.grad_partitions.values())

    def _api_params(self):
        return [p for g in self.fp16_groups for p in g]

    @torch.no_grad()
    def refresh_fp32_params(self) -> None:
        """Low-precision partitions -> fp32 masters (after the partitions were written from outside, e.g. a module-only
        load through ``GatheredParameters``); ref: stage3.py _restore_from_bit16_weights."""
        with self._Resident(self):
            for gi, params in enumerate(self.fp16_groups):
                master, off = self.fp32_partitioned_groups_flat[gi], 0
                for p in params:
                    n = p.ds_partition_numel
                    master.data.narrow(0, off, n).copy_(read_partition(p).detach().float().view(-1).to(master.device))
                    release_partition(p)
                    off += n

    _restore_from_bit16_weights = refresh_fp32_params

    def has_overflow(self, partition_gradients: bool = True) -> bool:
        return self.check_overflow(partition_gradients)

    def get_fp32_grad_partitions(self):
        """``{group: {param index: this rank's fp32 gradient slice}}`` of the gradients reduced so far."""
        out = {}
        for gi, params in enumerate(self.fp16_groups):
            for i, p in enumerate(params):
                g = self.grad_partitions.get(p.ds_id)
                if g is not None:
                    out.setdefault(gi, {})[i] = g.float()
        return out

    def empty_partition_cache(self) -> None:
        """Release every gathered, non-persistent parameter and forget the fetch trace's in-flight state."""
        for g in self.fp16_groups:
            for p in g:
                if p.ds_status != ZeroParamStatus.NOT_AVAILABLE and not p.ds_persist:
                    p.ds_active_sub_modules.clear()
                    _partition_param(p)
        off = self.parameter_offload
        if hasattr(off, "_pos"):
            off._pos = 0

    def offload_states(self, include=None, device="cpu", pin_memory: bool = True, non_blocking: bool = False) -> None:
        from .offload_states import offload_engine_states
        offload_engine_states(self._engine_view(), include, torch.device(str(getattr(device, "value", device))), pin_memory,
                              non_blocking)

    def reload_states(self, non_blocking: bool = False) -> None:
        from .offload_states import reload_engine_states
        reload_engine_states(self._engine_view(), non_blocking)

    def _engine_view(self):
        """The (optimizer, module) pair the state-movement helpers work on, with their bookkeeping kept on this object."""
        view = getattr(self, "_state_view", None)
        if view is None:
            import types
            view = self._state_view = types.SimpleNamespace(optimizer=self, module=self.module)
        return view

    def quantize_nontrainable_params(self) -> None:
        """ZeRO++ for frozen weights: from now on their all-gathers travel as int8 blocks (same wire format as qwZ)."""
        for p in self.module.parameters():
            if is_zero_param(p) and not p.requires_grad:
                p.ds_quantized_gather = True

    def checkpoint_event_prologue(self) -> None:
        """Before a checkpoint is written: NVMe-resident optimizer state is brought into memory."""
        self._ckpt_resident = self._Resident(self)
        self._ckpt_resident.__enter__()

    def checkpoint_event_epilogue(self) -> None:
        res = getattr(self, "_ckpt_resident", None)
        if res is not None:
            res.__exit__(None, None, None)
            self._ckpt_resident = None

    def reset_swap_buffers(self) -> None:
        sw = self.optimizer_swapper
        if sw is not None and hasattr(sw, "reset_swap_buffers"):
            sw.reset_swap_buffers()

    def get_fp32_state_dict_on_rank0(self):
        with self._Resident(self):
            return self._get_fp32_state_dict_on_rank0()

    def _get_fp32_state_dict_on_rank0(self):
        """Gather full fp32 parameters (for ``save_16bit_model`` / consolidation)."""
        out = {}
        for gi, params in enumerate(self.fp16_groups):
            master, off = self.fp32_partitioned_groups_flat[gi], 0
            for i, p in enumerate(params):
                n = p.ds_partition_numel
                local = master.detach().narrow(0, off, n).to(p.ds_device)
                off += n
                if self.partition_count > 1:
                    full = torch.empty(n * self.partition_count, dtype=local.dtype, device=local.device)
                    dist.all_gather_into_tensor(full, local.contiguous(), group=self.dp_process_group)
                else:
                    full = local
                out[self.param_names.get(p, f"group{gi}_param{i}")] = full[:p.ds_numel].view(p.ds_shape).cpu()
        return out


def estimate_zero3_model_states_mem_needs(total_params, largest_layer_params, num_gpus_per_node=1, num_nodes=1,
                                          cpu_offload=True, cpu_offload_params=True, zero_init=True,
                                          additional_buffer_factor=1.5):
    """(cpu bytes, gpu bytes, largest-layer bytes) — reference ``stage3.py:3044``."""
    total_gpus = num_nodes * num_gpus_per_node
    gpus_factor = 1 / num_nodes
    largest_layer_memory = 4 * largest_layer_params
    if cpu_offload:
        if cpu_offload_params:
            gpu_mem = largest_layer_memory
            cpu_mem = total_params * (18 * gpus_factor if zero_init else max(4 * num_gpus_per_node, 18 * gpus_factor)) * additional_buffer_factor
        else:
            gpu_mem = largest_layer_memory + int(2 * total_params / total_gpus)
            cpu_mem = total_params * (16 * gpus_factor if zero_init else max(4 * num_gpus_per_node, 16 * gpus_factor)) * additional_buffer_factor
    else:
        gpu_mem = largest_layer_memory + int(18 * total_params / total_gpus)
        cpu_mem = largest_layer_params * 4 * num_gpus_per_node * additional_buffer_factor if zero_init else \
            total_params * 4 * num_gpus_per_node * additional_buffer_factor
    return int(cpu_mem), int(gpu_mem), largest_layer_memory


import contextlib


@contextlib.contextmanager
def unwrap_model_for_generation(model):
    """ZeRO-3 engines: gather every parameter ONCE and take the fetch / release hooks off for the duration, so a
    ``generate`` loop does not all-gather each layer at every token (ref: stage3.py:77-106). ``model`` is the engine."""
    from .partition_parameters import GatheredParameters
    opt = getattr(model, "optimizer", None)
    offload = getattr(opt, "parameter_offload", None)
    module = getattr(model, "module", model)
    with GatheredParameters(list(module.parameters())):
        if offload is not None:
            offload.destroy()
        try:
            yield model
        finally:
            if offload is not None:
                offload._install(offload.module)
                offload.reset_trace()


# ---- printed estimate tables (ref: stage3.py:3044-3150) ----------------------------------------------------------------------
def model_to_params(model: torch.nn.Module):
    """(parameter count with tied parameters once, parameter count of the largest single module)."""
    total = sum({p.data_ptr(): p.numel() for p in model.parameters()}.values())
    largest = max((sum(p.numel() for p in m.parameters(recurse=False)) for m in model.modules()), default=0)
    return total, largest


def estimate_zero3_model_states_mem_needs_all_cold(total_params: int, largest_layer_params: int, num_gpus_per_node: int = 1,
                                                   num_nodes: int = 1, additional_buffer_factor: float = 1.5):
    """Print the per-CPU / per-GPU memory for params + optimizer states + gradients under ZeRO-3 for every combination of
    optimizer offload, parameter offload and ``zero.Init``; returns the rows."""
    print("Estimated memory needed for params, optim states and gradients for a:\n"
          f"HW: Setup with {num_nodes} {'nodes' if num_nodes > 1 else 'node'}, {num_gpus_per_node} "
          f"{'GPUs' if num_gpus_per_node > 1 else 'GPU'} per node.\n"
          f"SW: Model with {int(total_params / 1e6)}M total params, {int(largest_layer_params / 1e6)}M largest layer params.")
    print("  per CPU  |  per GPU |   Options")
    rows = []
    for cpu_offload in (True, False):
        for cpu_offload_params in (True, False):
            if cpu_offload_params and not cpu_offload:
                continue
            for zero_init in (True, False):
                cpu_mem, gpu_mem, _ = estimate_zero3_model_states_mem_needs(
                    total_params=total_params, largest_layer_params=largest_layer_params, num_gpus_per_node=num_gpus_per_node,
                    num_nodes=num_nodes, cpu_offload=cpu_offload, cpu_offload_params=cpu_offload_params, zero_init=zero_init,
                    additional_buffer_factor=additional_buffer_factor)
                opts = (f"offload_param={'cpu ' if cpu_offload_params else 'none'}, offload_optimizer="
                        f"{'cpu ' if cpu_offload else 'none'}, zero_init={1 if zero_init else 0}")
                print(f" {cpu_mem / 2**30:7.2f}GB | {gpu_mem / 2**30:6.2f}GB | {opts}")
                rows.append((cpu_mem, gpu_mem, opts))
    return rows


def estimate_zero3_model_states_mem_needs_all_live(model: torch.nn.Module, num_gpus_per_node: int = 1, num_nodes: int = 1,
                                                   additional_buffer_factor: float = 1.5):
    total, largest = model_to_params(model)
    return estimate_zero3_model_states_mem_needs_all_cold(total, largest, num_gpus_per_node, num_nodes, additional_buffer_factor)


def move_to_cpu(tensor_list) -> None:
    for t in tensor_list:
        t.data = t.data.cpu()


def lcm(x: int, y: int) -> int:
    return x * y // math.gcd(x, y)


def isclose(a, b, rtol: float = 1e-09, atol: float = 0.0) -> bool:
    return abs(a - b) <= max(rtol * max(abs(a), abs(b)), atol)


def print_rank_0(message, debug: bool = False, force: bool = False) -> None:
    if (debug or force) and (not dist.is_initialized() or dist.get_rank() == 0):
        logger.info(message)
