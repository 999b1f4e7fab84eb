"""ZeRO stage 1 (optimizer-state partitioning) and stage 2 (+ gradient partitioning).

Design (not the reference's hook/bucket machinery; behaviourally equivalent for the user):

* every param group's 16/32-bit parameters are flattened into ONE buffer padded to a multiple of the data-parallel
  world size; the parameters become views into it. Rank r owns the r-th slice and keeps an fp32 master copy of it
  (on the device, or in pinned host memory with ``offload_optimizer``) — the only tensor the wrapped optimizer sees.
* gradients: stage 1 accumulates full gradients locally and all-reduces once per optimizer step; stage 2
  reduce-scatters after EVERY micro-batch into an fp32 (or ``grad_accum_dtype``) accumulator that only covers the
  owned slice, then drops the full gradients, so gradient memory is 1/dp between micro-batches.
* ``step``: global overflow check (MAX all-reduce), unscale + clip by the global norm (partition norms summed over
  the group), optimizer step on the fp32 slice, copy into the flat buffer slice, one ``all_gather_into_tensor``.

Reference behaviour: ``deepspeed/runtime/zero/stage_1_and_2.py`` (DeepSpeedZeroOptimizer :97, step :1815,
reduce paths :912-1482, state_dict :2157, load :2321).
"""
from __future__ import annotations

import math
from typing import Dict, List, Optional

import torch

from ... import comm as dist
from ...utils.logging import logger
from ..fp16.loss_scaler import CreateLossScaler
from ..utils import align_dense_tensors
from .optimizer_api import ZeroOptimizerAPI

SINGLE_PARTITION_OF_FP32_GROUPS = "single_partition_of_fp32_groups"
BASE_OPTIMIZER_STATE = "base_optimizer_state"
PARTITION_COUNT = "partition_count"
ZERO_STAGE = "zero_stage"
LOSS_SCALER = "loss_scaler"
PARAM_SLICE_MAPPINGS = "param_slice_mappings"
DS_VERSION = "ds_version"
CLIP_GRAD = "clip_grad"
GROUP_PADDINGS = "group_paddings"


def _comm_dtype(name):
    return {None: None, "fp16": torch.float16, "bf16": torch.bfloat16, "bfp16": torch.bfloat16,
            "fp32": torch.float32}.get(name, name)


class DeepSpeedZeroOptimizer(ZeroOptimizerAPI):
    """Wraps ``init_optimizer``; see module docstring. ``partition_grads=True`` selects stage 2."""

    def __init__(self, init_optimizer, param_names: Optional[dict] = None, timers=None, static_loss_scale: float = 1.0,
                 dynamic_loss_scale: bool = False, dynamic_loss_args: Optional[dict] = None, verbose: bool = True,
                 contiguous_gradients: bool = True, reduce_bucket_size: int = 500000000,
                 use_multi_rank_bucket_allreduce: bool = True, allgather_bucket_size: int = 5000000000,
                 dp_process_group=None, expert_parallel_group=None, expert_data_parallel_group=None,
                 reduce_scatter: bool = True, overlap_comm: bool = False, offload_optimizer_config=None, mpu=None,
                 clip_grad: float = 0.0, gradient_accumulation_dtype=torch.float32, communication_data_type=None,
                 postscale_gradients: bool = True, gradient_predivide_factor: float = 1.0,
                 gradient_accumulation_steps: int = 1, ignore_unused_parameters: bool = True,
                 partition_grads: bool = True, round_robin_gradients: bool = False, has_moe_layers: bool = False,
                 fp16_master_weights_and_gradients: bool = False, elastic_checkpoint: bool = False,
                 check_grad_overflow: bool = True, **unused):
        self.optimizer = init_optimizer
        self.param_names = param_names or {}
        self.timers = timers
        self.mpu = mpu
        self.dp_process_group = dp_process_group
        self.partition_count = dist.get_world_size(group=dp_process_group) if dist.is_initialized() else 1
        self.rank = dist.get_rank(group=dp_process_group) if dist.is_initialized() else 0
        self.partition_gradients = partition_grads
        self.zero_stage = 2 if partition_grads else 1
        self.contiguous_gradients = contiguous_gradients
        self.reduce_scatter = reduce_scatter
        self.overlap_comm = overlap_comm
        self.clip_grad = clip_grad
        self.gradient_predivide_factor = gradient_predivide_factor
        self.postscale_gradients = postscale_gradients
        self.gradient_accumulation_steps = gradient_accumulation_steps
        self.micro_step_id = 0
        self.communication_data_type = _comm_dtype(communication_data_type)
        self.gradient_accumulation_dtype = gradient_accumulation_dtype or torch.float32
        self.check_grad_overflow = check_grad_overflow
        off = offload_optimizer_config
        self.cpu_offload = bool(off is not None and str(getattr(off, "device", "none")).endswith("cpu"))
        self.cpu_offload_pin_memory = bool(off is not None and getattr(off, "pin_memory", False))
        self.overflow = False
        self._global_grad_norm = 0.0
        self.skipped_steps = 0
        self.custom_loss_scaler = False
        self.external_loss_scale = None

        # ---- flatten + partition ---------------------------------------------------------------------
        self.bit16_groups: List[List[torch.nn.Parameter]] = []
        self.bit16_groups_flat: List[torch.Tensor] = []
        self.groups_padding: List[int] = []
        self.single_partition_of_fp32_groups: List[torch.Tensor] = []
        self.grad_partitions: List[Optional[torch.Tensor]] = []  # stage-2 accumulators (owned slice)
        self.param_slice_mappings: List[Dict[str, tuple]] = []
        align = self.partition_count * 2  # keeps every slice 4-byte aligned for 16-bit dtypes
        for gi, group in enumerate(self.optimizer.param_groups):
            params = [p for p in group["params"] if p.requires_grad]
            self.bit16_groups.append(params)
            if not params:
                self.bit16_groups_flat.append(torch.empty(0))
                self.groups_padding.append(0)
                self.single_partition_of_fp32_groups.append(torch.empty(0, requires_grad=True))
                self.grad_partitions.append(None)
                self.param_slice_mappings.append({})
                group["params"] = []
                continue
            dev, dt = params[0].device, params[0].dtype
            assert all(p.dtype == dt for p in params), "all parameters of a ZeRO param group must share one dtype"
            padded = align_dense_tensors([p.data for p in params], align)
            flat = torch.cat([t.contiguous().view(-1) for t in padded]).to(dev)
            n_real = sum(p.numel() for p in params)
            self.groups_padding.append(flat.numel() - n_real)
            off_ = 0
            for p in params:  # parameters become views of the flat buffer
                p.data = flat.narrow(0, off_, p.numel()).view_as(p.data)
                off_ += p.numel()
            self.bit16_groups_flat.append(flat)
            psize = flat.numel() // self.partition_count
            lo = self.rank * psize
            master = flat.narrow(0, lo, psize).detach().clone().float()
            if self.cpu_offload:
                master = master.cpu()
                if self.cpu_offload_pin_memory and torch.cuda.is_available():
                    master = master.pin_memory()
            master.requires_grad = True
            self.single_partition_of_fp32_groups.append(master)
            group["params"] = [master]
            self.grad_partitions.append(None)
            # which slice of each parameter this rank owns: name -> (start in partition, numel, offset in param)
            mapping, cursor = {}, 0
            for p in params:
                s, e = max(cursor, lo), min(cursor + p.numel(), lo + psize)
                if s < e:
                    mapping[self.param_names.get(p, f"group{gi}_param{len(mapping)}")] = (s - lo, e - s, s - cursor)
                cursor += p.numel()
            self.param_slice_mappings.append(mapping)

        # ---- loss scaling ----------------------------------------------------------------------------
        self.dtype = next((g[0].dtype for g in self.bit16_groups if g), torch.float32)
        self.loss_scaler = CreateLossScaler(dtype=self.dtype, static_loss_scale=static_loss_scale,
                                            dynamic_scaling=dynamic_loss_scale, dynamic_loss_args=dynamic_loss_args)
        self.dynamic_loss_scale = self.loss_scaler.dynamic
        if verbose and self.rank == 0:
            logger.info(f"ZeRO stage {self.zero_stage}: {self.partition_count} partitions, "
                        f"{sum(f.numel() for f in self.bit16_groups_flat)} elements, cpu_offload={self.cpu_offload}")
        self._initialize_optimizer_states()

    # -------------------------------------------------------------------------------------------------
    def _initialize_optimizer_states(self):
        """Run one zero-gradient step so the wrapped optimizer allocates its state now (checkpoint loading and memory
        accounting need it before the first real step). Skipped for optimizers where a zero-grad step is not a no-op."""
        name = type(self.optimizer).__name__.lower()
        if not any(k in name for k in ("adam", "adagrad", "lion")) or "lion" in name:
            return
        saved = [dict(g) for g in self.optimizer.param_groups]
        for g in self.optimizer.param_groups:
            g["weight_decay"] = 0.0
            g["lr"] = 0.0
        for m in self.single_partition_of_fp32_groups:
            if m.numel():
                m.grad = torch.zeros_like(m)
        self.optimizer.step()
        for m in self.single_partition_of_fp32_groups:
            m.grad = None
        for g, s in zip(self.optimizer.param_groups, saved):
            g["weight_decay"], g["lr"] = s["weight_decay"], s["lr"]
            if "step" in g:
                g["step"] = 0
        for st in self.optimizer.state.values():
            if "step" in st:
                st["step"] = st["step"] * 0 if torch.is_tensor(st["step"]) else 0

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

    def zero_grad(self, set_to_none: bool = True):
        for group in self.bit16_groups:
            for p in group:
                if set_to_none:
                    p.grad = None
                elif p.grad is not None:
                    p.grad.detach_().zero_()

    # -------------------------------------------------------------------------------------------------
    # backward / gradient reduction
    # -------------------------------------------------------------------------------------------------
    def backward(self, loss: torch.Tensor, retain_graph: bool = False):
        self.micro_step_id += 1
        if self.custom_loss_scaler:
            (loss * self.external_loss_scale).backward(retain_graph=retain_graph)
        else:
            self.loss_scaler.backward(loss.float(), retain_graph=retain_graph)
        if self.partition_gradients:
            self._reduce_micro_batch()

    def _flat_grads(self, gi: int) -> torch.Tensor:
        flat = self.bit16_groups_flat[gi]
        cd = self.communication_data_type or flat.dtype
        buf = torch.zeros(flat.numel(), dtype=cd, device=flat.device)
        off = 0
        for p in self.bit16_groups[gi]:
            if p.grad is not None:
                buf.narrow(0, off, p.numel()).copy_(p.grad.detach().reshape(-1))
            off += p.numel()
        return buf

    def _reduce_to_partition(self, gi: int) -> torch.Tensor:
        """Average this group's full gradients over the data-parallel group; return the owned slice."""
        buf = self._flat_grads(gi)
        ws = self.partition_count
        psize = buf.numel() // ws
        if ws == 1:
            return buf
        # average = (sum / pre) / post with pre * post = world size. postscale (default): pre is the configured predivide
        # factor (1 unless set) and the rest is divided out after the collective; prescale: everything is divided out BEFORE
        # the collective so a 16-bit sum over many ranks cannot overflow.
        pre = self.gradient_predivide_factor if self.postscale_gradients else float(ws)
        post = ws / pre
        if pre != 1.0:
            buf.div_(pre)
        if self.reduce_scatter and self.partition_gradients:
            out = torch.empty(psize, dtype=buf.dtype, device=buf.device)
            dist.reduce_scatter_tensor(out, buf, group=self.dp_process_group)
        else:
            dist.all_reduce(buf, group=self.dp_process_group)
            out = buf.narrow(0, self.rank * psize, psize)
        return out.div_(post) if post != 1.0 else out

    def _reduce_micro_batch(self):
        for gi, params in enumerate(self.bit16_groups):
            if not params:
                continue
            part = self._reduce_to_partition(gi).to(self.gradient_accumulation_dtype)
            if self.grad_partitions[gi] is None:
                self.grad_partitions[gi] = part.clone()
            else:
                self.grad_partitions[gi].add_(part)
            for p in params:
                p.grad = None

    def reduce_gradients(self, pipeline_parallel: bool = False):
        """Stage 1: called once at the accumulation boundary (stage 2 already reduced per micro-batch)."""
        if self.partition_gradients:
            return
        for gi, params in enumerate(self.bit16_groups):
            if params:
                self.grad_partitions[gi] = self._reduce_to_partition(gi).to(self.gradient_accumulation_dtype).clone()

    def overlapping_partition_gradients_reduce_epilogue(self):
        self.reduce_gradients()

    # -------------------------------------------------------------------------------------------------
    # step
    # -------------------------------------------------------------------------------------------------
    def _all_reduce_scalar(self, v: float, op) -> float:
        if self.partition_count == 1 and self.mpu is None:
            return v
        dev = next((f.device for f in self.bit16_groups_flat if f.numel()), torch.device("cpu"))
        t = torch.tensor([float(v)], dtype=torch.float32, device=dev)
        if self.partition_count > 1:
            dist.all_reduce(t, op=op, group=self.dp_process_group)
        if self.mpu is not None:
            dist.all_reduce(t, op=op, group=self.mpu.get_model_parallel_group())
        return float(t[0])

    def has_overflow(self) -> bool:
        local = 0.0
        for g in self.grad_partitions:
            if g is not None and g.numel():
                s = float(g.float().sum())
                if s in (float("inf"), -float("inf")) or s != s:
                    local = 1.0
                    break
        return self._all_reduce_scalar(local, dist.ReduceOp.MAX) > 0

    def scaled_global_norm(self, norm_type: float = 2) -> float:
        sq = sum(float(g.float().norm(2)) ** 2 for g in self.grad_partitions if g is not None and g.numel())
        total = self._all_reduce_scalar(sq, dist.ReduceOp.SUM)
        n = math.sqrt(total)
        return -1.0 if (n in (float("inf"), ) or n != n) else n

    def get_global_grad_norm(self) -> float:
        return self._global_grad_norm

    def _combined_scale(self, scaled_norm: float, scale: float) -> float:
        """Divisor that removes the loss scale and, when the global norm exceeds ``clip_grad``, clips in the same pass."""
        if self.clip_grad > 0.0:
            clip = (scaled_norm / scale + 1e-6) / self.clip_grad
            if clip > 1:
                scale = clip * scale
        return scale

    def step(self, closure=None):
        self.micro_step_id = 0
        if not self.partition_gradients:
            self.reduce_gradients()
            for params in self.bit16_groups:
                for p in params:
                    p.grad = None
        self.overflow = self.has_overflow() if self.check_grad_overflow else False
        prev_scale = self.loss_scale
        if not self.custom_loss_scaler:
            self.loss_scaler.update_scale(self.overflow)
        if self.overflow:
            self.skipped_steps += 1
            self.grad_partitions = [None] * len(self.grad_partitions)
            if self.rank == 0:
                logger.info(f"[deepspeed] OVERFLOW! Rank {self.rank} Skipping step. Attempted loss scale: {prev_scale}, "
                            f"reducing to {self.loss_scale}")
            return
        scaled_norm = self.scaled_global_norm()
        self._global_grad_norm = scaled_norm / prev_scale
        combined = self._combined_scale(scaled_norm, prev_scale)
        for gi, master in enumerate(self.single_partition_of_fp32_groups):
            g = self.grad_partitions[gi]
            if g is None or not master.numel():
                continue
            master.grad = (g.to(master.device, dtype=torch.float32) / combined)
        self.grad_partitions = [None] * len(self.grad_partitions)
        self.optimizer.step()
        for gi, master in enumerate(self.single_partition_of_fp32_groups):
            master.grad = None
            flat = self.bit16_groups_flat[gi]
            if not flat.numel():
                continue
            psize = flat.numel() // self.partition_count
            flat.narrow(0, self.rank * psize, psize).copy_(master.detach().to(flat.device))
            if self.partition_count > 1:
                dist.all_gather_into_tensor(flat, flat.narrow(0, self.rank * psize, psize).clone(),
                                            group=self.dp_process_group)

    # -------------------------------------------------------------------------------------------------
    # checkpointing
    # -------------------------------------------------------------------------------------------------
    def _param_shapes(self):
        return [{self.param_names.get(p, f"group{gi}_param{i}"): torch.Size(p.shape) for i, p in enumerate(params)}
                for gi, params in enumerate(self.bit16_groups)]

    def state_dict(self):
        return {
            LOSS_SCALER: self.loss_scaler,
            "dynamic_loss_scale": self.dynamic_loss_scale,
            "overflow": self.overflow,
            CLIP_GRAD: self.clip_grad,
            BASE_OPTIMIZER_STATE: self.optimizer.state_dict(),
            SINGLE_PARTITION_OF_FP32_GROUPS: [m.detach().clone() for m in self.single_partition_of_fp32_groups],
            ZERO_STAGE: self.zero_stage,
            GROUP_PADDINGS: list(self.groups_padding),
            PARTITION_COUNT: self.partition_count,
            PARAM_SLICE_MAPPINGS: self.param_slice_mappings,
            "param_shapes": self._param_shapes(),
        }

    def load_state_dict(self, state_dict_list, load_optimizer_states: bool = True, load_from_fp32_weights: bool = False,
                        checkpoint_folder=None, load_serial=None, param_shapes=None):
        sd = state_dict_list[self.rank] if isinstance(state_dict_list, (list, tuple)) else state_dict_list
        if sd.get(PARTITION_COUNT, self.partition_count) != self.partition_count:
            raise RuntimeError(f"checkpoint was written with dp world size {sd[PARTITION_COUNT]}, current is "
                               f"{self.partition_count}; convert it with the universal-checkpoint tools first")
        ls = sd.get(LOSS_SCALER)
        if ls is not None:
            self.loss_scaler = ls
        self.overflow = sd.get("overflow", False)
        self.clip_grad = sd.get(CLIP_GRAD, self.clip_grad)
        if load_optimizer_states:
            self.optimizer.load_state_dict(sd[BASE_OPTIMIZER_STATE])
        if load_from_fp32_weights or True:
            for cur, saved in zip(self.single_partition_of_fp32_groups, sd[SINGLE_PARTITION_OF_FP32_GROUPS]):
                cur.data.copy_(saved.data.to(cur.device))
        # refresh the 16-bit parameters from the restored fp32 slices
        for gi, master in enumerate(self.single_partition_of_fp32_groups):
            flat = self.bit16_groups_flat[gi]
            if not flat.numel():
                continue
            psize = flat.numel() // self.partition_count
            flat.narrow(0, self.rank * psize, psize).copy_(master.detach().to(flat.device))
            if self.partition_count > 1:
                dist.all_gather_into_tensor(flat, flat.narrow(0, self.rank * psize, psize).clone(),
                                            group=self.dp_process_group)

    def refresh_fp32_params(self):
        for gi, master in enumerate(self.single_partition_of_fp32_groups):
            flat = self.bit16_groups_flat[gi]
            if flat.numel():
                psize = flat.numel() // self.partition_count
                master.data.copy_(flat.narrow(0, self.rank * psize, psize).float().to(master.device))

    _restore_from_bit16_weights = refresh_fp32_params

    # ---- hooks of ZeroOptimizerAPI + the stage-1/2 part of the reference's public surface -------------------------------
    def _api_master_tensors(self):
        return self.single_partition_of_fp32_groups

    def _api_local_grads(self):
        if any(g is not None for g in self.grad_partitions):
            return [g for g in self.grad_partitions if g is not None]
        return [p.grad for params in self.bit16_groups for p in params if p.grad is not None]

    def _api_params(self):
        return [p for params in self.bit16_groups for p in params]

    def get_bit16_param_group(self, group_no: int):
        """This rank's slice of group ``group_no``'s flat low-precision buffer."""
        flat = self.bit16_groups_flat[group_no]
        psize = flat.numel() // self.partition_count
        return flat.narrow(0, self.rank * psize, psize)

    def update_lp_params(self) -> None:
        """Master weights -> low-precision parameters on every rank (what ``step`` does after the optimizer update; call
        it after modifying the fp32 master copy by hand)."""
        for gi, master in enumerate(self.single_partition_of_fp32_groups):
            flat = self.bit16_groups_flat[gi]
            if not flat.numel():
                continue
            own = self.get_bit16_param_group(gi)
            own.copy_(master.detach().to(flat.device))
            if self.partition_count > 1:
                dist.all_gather_into_tensor(flat, own.clone(), group=self.dp_process_group)

    def is_moe_group(self, group) -> bool:
        return bool(group.get("moe", False))

    def get_ep_ranks(self, rank: int = 0, group_name: Optional[str] = None):
        from ...utils import groups
        ep = groups._get_expert_parallel_group(group_name)
        return list(range(dist.get_world_size(group=ep))) if ep is not None else [0]

    def backward_prologue(self) -> None:
        self.micro_step_id += 0     # accumulation bookkeeping lives in backward(); kept for callers that bracket it

    def backward_epilogue(self) -> None:
        pass

    def print_rank_0(self, message) -> None:
        if not dist.is_initialized() or dist.get_rank() == 0:
            logger.info(message)


def estimate_zero2_model_states_mem_needs(total_params: int, num_gpus_per_node: int = 1, num_nodes: int = 1,
                                          cpu_offload: bool = True, additional_buffer_factor: float = 1.5):
    """(cpu bytes, gpu bytes) per device for params + grads + optimizer states (reference ``stage_1_and_2.py:2479``)."""
    total_gpus = num_nodes * num_gpus_per_node
    if cpu_offload:
        gpu_mem = 2 * total_params
        cpu_mem = total_params * max(4 * total_gpus, 16) * additional_buffer_factor
    else:
        gpu_mem = 4 * total_params + int(16 * total_params / total_gpus)
        cpu_mem = total_params * 4 * num_gpus_per_node * additional_buffer_factor
    return int(cpu_mem), int(gpu_mem)


# ---- printed estimate tables (ref: stage_1_and_2.py:2462-2540) ------------------------------------------------------------
def model_to_params(model: torch.nn.Module) -> int:
    """Parameter count with tied parameters counted once."""
    return sum({p.data_ptr(): p.numel() for p in model.parameters()}.values())


def _hw_line(num_nodes: int, num_gpus_per_node: int) -> str:
    return (f"HW: Setup with {num_nodes} {'nodes' if num_nodes > 1 else 'node'}, {num_gpus_per_node} "
            f"{'GPUs' if num_gpus_per_node > 1 else 'GPU'} per node.")


def estimate_zero2_model_states_mem_needs_all_cold(total_params: int, num_gpus_per_node: int = 1, num_nodes: int = 1,
                                                   additional_buffer_factor: float = 1.5):
    """Print the per-CPU / per-GPU memory for params + optimizer states + gradients under ZeRO-2, with and without
    optimizer offload, for a model given only by its parameter count; returns the rows."""
    print("Estimated memory needed for params, optim states and gradients for a:\n"
          f"{_hw_line(num_nodes, num_gpus_per_node)}\nSW: Model with {int(total_params / 1e6)}M total params.")
    print("  per CPU  |  per GPU |   Options")
    rows = []
    for cpu_offload in (True, False):
        cpu_mem, gpu_mem = estimate_zero2_model_states_mem_needs(total_params=total_params, num_gpus_per_node=num_gpus_per_node,
                                                                 num_nodes=num_nodes, cpu_offload=cpu_offload,
                                                                 additional_buffer_factor=additional_buffer_factor)
        opts = f"offload_optimizer={'cpu ' if cpu_offload else 'none'}"
        print(f" {cpu_mem / 2**30:7.2f}GB | {gpu_mem / 2**30:6.2f}GB | {opts}")
        rows.append((cpu_mem, gpu_mem, opts))
    return rows


def estimate_zero2_model_states_mem_needs_all_live(model: torch.nn.Module, num_gpus_per_node: int = 1, num_nodes: int = 1,
                                                   additional_buffer_factor: float = 1.5):
    return estimate_zero2_model_states_mem_needs_all_cold(model_to_params(model), num_gpus_per_node, num_nodes,
                                                          additional_buffer_factor)


def get_alignment_padding(tensor_list, alignment: int) -> int:
    """Elements to append so that the total of ``tensor_list`` is a multiple of ``alignment``."""
    rem = sum(t.numel() for t in tensor_list) % alignment
    return (alignment - rem) if rem else 0


def lcm(x: int, y: int) -> int:
    return x * y // math.gcd(x, y)


def isclose(a, b, rtol: float = 1e-09, atol: float = 0.0) -> bool:
    return abs(a - b) <= max(rtol * max(abs(a), abs(b)), atol)


def split_half_float_double(tensors):
    """Group tensors by (device kind, dtype): one bucket per combination present, in a fixed dtype order."""
    order = (torch.float16, torch.float32, torch.float64, torch.bfloat16)
    kind = tensors[0].device.type if tensors else "cpu"
    return [[t for t in tensors if t.dtype == dt and t.device.type == kind] for dt in order
            if any(t.dtype == dt and t.device.type == kind for t in tensors)]


def print_rank_msg(msg) -> None:
    print(f"rank {dist.get_rank() if dist.is_initialized() else 0} - {msg}")
