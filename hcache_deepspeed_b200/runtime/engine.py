"""DeepSpeedEngine — the training engine returned by ``initialize()``.

``loss = engine(batch); engine.backward(loss); engine.step()`` with gradient accumulation, mixed precision
(fp16 + loss scaling, bf16 with fp32 masters), ZeRO stages 0-3 (+ CPU optimizer offload), gradient clipping,
LR schedules, data loader construction, monitors / timers / FLOPS profiler, and checkpoint save / load.

Behavioural reference: ``deepspeed/runtime/engine.py`` (DeepSpeedEngine :183; forward :2006, backward :2165,
step :2352, _take_model_step :2279, allreduce_gradients :2088, save_checkpoint :3230, load_checkpoint :2850,
_configure_optimizer :1370, _configure_zero_optimizer :1660) and ``deepspeed/__init__.py:69-235`` (initialize).
Data-parallel gradient averaging without ZeRO is a bucketed flat all-reduce; everything distributed goes through
:mod:`hcache_deepspeed_b200.comm` (NCCL on GPUs, gloo on CPU).
"""
from __future__ import annotations

import os
from typing import Callable, Dict, Iterable, Optional, Union

import torch

from .. import comm as dist
from ..monitor.monitor import MonitorMaster
from ..utils import groups
from ..utils.logging import log_dist, logger
from ..utils.timer import NoopTimer, SynchronizedWallClockTimer, ThroughputTimer
from . import lr_schedules
from .bf16_optimizer import BF16_Optimizer
from .checkpoint_engine import NebulaCheckpointEngine, TorchCheckpointEngine
from .config import (ADAGRAD_OPTIMIZER, ADAM_OPTIMIZER, ADAMW_OPTIMIZER, LAMB_OPTIMIZER, LION_OPTIMIZER,
                     ONEBIT_ADAM_OPTIMIZER, ONEBIT_LAMB_OPTIMIZER, SGD_OPTIMIZER, ZERO_ONE_ADAM_OPTIMIZER, DeepSpeedConfig)
from .dataloader import DeepSpeedDataLoader
from .engine_accessors import EngineAccessors
from .fp16.fused_optimizer import FP16_Optimizer, FP16_UnfusedOptimizer
from .utils import clip_grad_norm_, flatten_dense_tensors, get_grad_norm, unflatten_dense_tensors
from .zero.stage3 import DeepSpeedZeroOptimizer_Stage3
from .zero.stage_1_and_2 import DeepSpeedZeroOptimizer

FORWARD_MICRO_TIMER, FORWARD_GLOBAL_TIMER = "fwd_microstep", "fwd"
BACKWARD_MICRO_TIMER, BACKWARD_GLOBAL_TIMER = "bwd_microstep", "bwd"
STEP_MICRO_TIMER, STEP_GLOBAL_TIMER = "step_microstep", "step"
MEMORY_OPT_ALLREDUCE_SIZE = 500000000


class EngineTimers:
    """Names of the wall-clock timers the engine drives."""

    def __init__(self, enable_micro_timers: bool, enable_global_timers: bool):
        self.forward_timers = [FORWARD_MICRO_TIMER] if enable_micro_timers else []
        self.backward_timers = [BACKWARD_MICRO_TIMER] if enable_micro_timers else []
        self.step_timers = [STEP_MICRO_TIMER] if enable_micro_timers else []
        self.global_timers = [FORWARD_GLOBAL_TIMER, BACKWARD_GLOBAL_TIMER, STEP_GLOBAL_TIMER] if enable_global_timers else []
        self.micro_timers = self.forward_timers + self.backward_timers + self.step_timers


def _device_for(local_rank: int) -> torch.device:
    if torch.cuda.is_available():
        return torch.device("cuda", local_rank if local_rank >= 0 else 0)
    return torch.device("cpu")


class DeepSpeedEngine(EngineAccessors, torch.nn.Module):

    def __init__(self, args=None, model: torch.nn.Module = None, optimizer=None, model_parameters=None, training_data=None,
                 lr_scheduler=None, mpu=None, dist_init_required: Optional[bool] = None, collate_fn=None, config=None,
                 config_class: Optional[DeepSpeedConfig] = None, mesh_device=None, dont_change_device: bool = False):
        super().__init__()
        assert model is not None, "deepspeed.initialize requires a model"
        self.client_optimizer = optimizer
        self.client_lr_scheduler = lr_scheduler
        self.training_data = training_data
        self.collate_fn = collate_fn
        self.mpu = mpu
        self.global_steps = 0
        self.global_samples = 0
        self.micro_steps = 0
        self.skipped_steps = 0
        self.gradient_average = True
        self.warn_unscaled_loss = True
        self.loaded_checkpoint_mp_world_size = None
        self.loaded_checkpoint_dp_world_size = None
        self.enable_backward_allreduce = True
        self._is_gradient_accumulation_boundary = None
        self.losses = None
        self._global_grad_norm = None
        self.dont_change_device = dont_change_device

        if dist_init_required is None or dist_init_required:
            if not dist.is_initialized() and int(os.environ.get("WORLD_SIZE", "1")) > 1:
                dist.init_distributed()
        self.local_rank = int(os.environ.get("LOCAL_RANK", getattr(args, "local_rank", 0) if args is not None else 0) or 0)
        self.device = _device_for(self.local_rank)
        if self.device.type == "cuda":
            torch.cuda.set_device(self.device)
        self.global_rank = dist.get_rank() if dist.is_initialized() else 0
        self.world_size = dist.get_world_size() if dist.is_initialized() else 1

        if mpu is not None:
            groups.mpu = mpu
        self._config = config_class if config_class is not None else DeepSpeedConfig(
            config if config is not None else getattr(args, "deepspeed_config", None), mpu=mpu)
        self._configure_groups()
        self.dp_world_size = dist.get_world_size(group=self.data_parallel_group) if dist.is_initialized() else 1
        self.mp_world_size = mpu.get_model_parallel_world_size() if mpu is not None else 1
        self.seq_parallel_group = groups._get_sequence_parallel_group()
        self.sequence_parallel_size = groups._get_sequence_parallel_world_size() if dist.is_initialized() else 1

        if self._config.comms_config.enabled:
            dist.configure(self._config)
        self.monitor = MonitorMaster(self._config.monitor_config)
        self.timers = SynchronizedWallClockTimer() if self.wall_clock_breakdown() else NoopTimer()
        self.engine_timers = EngineTimers(self.wall_clock_breakdown(), self.wall_clock_breakdown())
        self.tput_timer = ThroughputTimer(batch_size=self.train_batch_size(), steps_per_output=self.steps_per_print())

        self._configure_module(model)
        self.param_names = {p: n for n, p in self.module.named_parameters()}
        self.optimizer = None
        self.basic_optimizer = None
        self.lr_scheduler = None
        has_optimizer = optimizer is not None or self._config.optimizer_name is not None
        if model_parameters is None:
            model_parameters = [p for p in self.module.parameters() if p.requires_grad]
        if has_optimizer:
            self._configure_optimizer(optimizer, list(model_parameters))
            self._configure_lr_scheduler(lr_scheduler)
        elif self.zero_optimization_stage() == 3:
            from .zero.stage3 import DeepSpeedZeRoOffload
            self.optimizer = None
            self._zero3_offload = DeepSpeedZeRoOffload(self.module, dp_process_group=self.data_parallel_group,
                                                       param_persistence_threshold=self._config.zero_config.param_persistence_threshold,
                                                       offload_param_config=self._config.zero_config.offload_param)
        self.training_dataloader = self.deepspeed_io(training_data) if training_data is not None else None
        self.checkpoint_engine = None
        self._configure_checkpointing()
        # compression training: techniques switch on at their schedule offsets as optimizer steps go by (the model must have
        # been prepared with compression.init_compression; ref: engine.py _configure_compression_scheduler)
        self.compression_scheduler = None
        if self._config.compression_config:
            from ..compression.compress import get_compression_config
            from ..compression.scheduler import compression_scheduler
            self.compression_scheduler = compression_scheduler(self.module, get_compression_config(self._config._param_dict))
            self.compression_scheduler.step(step_zero_check=True)
        self._amp_autocast = False
        if self.amp_enabled():
            if self.fp16_enabled() or self.bfloat16_enabled():
                raise ValueError("amp cannot be combined with the fp16 / bf16 blocks (ref: config.py _do_error_check)")
            if self.zero_optimization_stage() > 0:
                raise ValueError("amp is not compatible with ZeRO: use the bf16 block")
            params = self._config.amp_params or {}
            if str(params.get("dtype", "bf16")).lower() in ("fp16", "float16", "half"):
                raise NotImplementedError('amp with fp16 autocast is not provided: use "fp16": {"enabled": true}')
            self._amp_autocast = str(params.get("opt_level", "O1")).upper() != "O0"
            log_dist(f"amp: fp32 parameters, torch.autocast(bfloat16) {'on' if self._amp_autocast else 'off (O0)'}", ranks=[0])
        self.flops_profiler = None
        if self._config.flops_profiler_config.enabled:
            from ..profiling.flops_profiler.profiler import FlopsProfiler
            self.flops_profiler = FlopsProfiler(self.module, self,
                                                recompute_fwd_factor=self._config.flops_profiler_config.recompute_fwd_factor)
        self.progressive_layer_drop = None
        if self._config.pld_enabled:
            from .progressive_layer_drop import ProgressiveLayerDrop
            self.progressive_layer_drop = ProgressiveLayerDrop(theta=self._config.pld_config.theta,
                                                               gamma=self._config.pld_config.gamma)
        self.random_ltd_scheduler = None
        if self.random_ltd_enabled():
            # random layer-token drop: the scheduler grows the kept-token count; layers wrapped by convert_to_random_ltd
            # read it (ref: engine.py:365-369)
            ltd = self.random_ltd_config()
            ltd["global_batch_size"] = self.train_batch_size()
            ltd["micro_batch_size"] = self.train_micro_batch_size_per_gpu()
            self.random_ltd_scheduler = self._configure_random_ltd_scheduler(ltd)
        self.curriculum_scheduler_legacy = None
        if self._config.curriculum_enabled_legacy:
            from .data_pipeline.curriculum_scheduler import CurriculumScheduler
            self.curriculum_scheduler_legacy = CurriculumScheduler(self._config.curriculum_params_legacy)
        self.eigenvalue = None
        if self._config.eigenvalue_enabled:
            from .eigenvalue import Eigenvalue
            ec = self._config.eigenvalue_config
            self.eigenvalue = Eigenvalue(verbose=ec.verbose, max_iter=ec.max_iter, tol=ec.tol, stability=ec.stability,
                                         gas_boundary_resolution=ec.gas_boundary_resolution, layer_name=ec.layer_name,
                                         layer_num=ec.layer_num)
        if self.global_rank == 0 and self._config.dump_state:
            self._config.print("DeepSpeedEngine configuration")

    # ------------------------------------------------------------------------------------------------
    # configuration accessors (same names as the reference engine)
    # ------------------------------------------------------------------------------------------------
    def train_batch_size(self):
        return self._config.train_batch_size

    def train_micro_batch_size_per_gpu(self):
        return self._config.train_micro_batch_size_per_gpu

    def gradient_accumulation_steps(self):
        return self._config.gradient_accumulation_steps

    def set_train_batch_size(self, train_batch_size: int):
        """Adjust the global batch size by changing the number of accumulation steps (micro-batch size is fixed)."""
        per_step = self.train_micro_batch_size_per_gpu() * self.dp_world_size
        if train_batch_size % per_step != 0:
            raise ValueError("Train batch size must be divisible by micro-batch size * data parallelism")
        self._config.train_batch_size = train_batch_size
        self._config.gradient_accumulation_steps = train_batch_size // per_step

    def set_train_micro_batch_size(self, micro_batch_size: int):
        self._config.train_batch_size = micro_batch_size * self.gradient_accumulation_steps() * self.dp_world_size
        self._config.train_micro_batch_size_per_gpu = micro_batch_size

    def steps_per_print(self):
        return self._config.steps_per_print

    def wall_clock_breakdown(self):
        return self._config.wall_clock_breakdown

    def zero_optimization(self):
        return self._config.zero_enabled

    def zero_optimization_stage(self):
        return self._config.zero_optimization_stage

    def zero_cpu_offload(self):
        o = self._config.zero_config.offload_optimizer
        return o is not None and str(getattr(o.device, "value", o.device)) == "cpu"

    def zero_offload_param(self):
        return self._config.zero_config.offload_param

    def fp16_enabled(self):
        return self._config.fp16_enabled

    def bfloat16_enabled(self):
        return self._config.bfloat16_enabled

    def amp_enabled(self):
        return self._config.amp_enabled

    def gradient_clipping(self):
        return self._config.gradient_clipping

    def dynamic_loss_scale(self):
        return self._config.loss_scale == 0

    def loss_scale(self):
        return self._config.loss_scale

    def optimizer_name(self):
        return self.client_optimizer.__class__.__name__ if self.client_optimizer else self._config.optimizer_name

    def optimizer_params(self):
        return self._config.optimizer_params

    def scheduler_name(self):
        return self._config.scheduler_name

    def scheduler_params(self):
        return self._config.scheduler_params

    def postscale_gradients(self):
        return not self._config.prescale_gradients

    def gradient_predivide_factor(self):
        return self._config.gradient_predivide_factor

    def communication_data_type(self):
        return {None: None, "fp16": torch.float16, "bf16": torch.bfloat16, "fp32": torch.float32}.get(
            self._config.communication_data_type, None)

    def get_data_types(self):
        model_dtype = torch.float16 if self.fp16_enabled() else torch.bfloat16 if self.bfloat16_enabled() else torch.float32
        acc = {None: model_dtype if not self.bfloat16_enabled() else torch.float32, "fp32": torch.float32,
               "fp16": torch.float16, "bf16": torch.bfloat16}[self._config.grad_accum_dtype]
        return model_dtype, acc

    @property
    def config(self):
        return self._config._param_dict

    # ------------------------------------------------------------------------------------------------
    def _configure_groups(self):
        cfg = self._config
        if not dist.is_initialized():
            self.data_parallel_group = None
            return
        if self.mpu is not None:
            self.data_parallel_group = self.mpu.get_data_parallel_group()
        elif cfg.sequence_parallel_size > 1:
            groups._create_model_parallel(1, cfg.sequence_parallel_size)
            self.data_parallel_group = groups._get_sequence_data_parallel_group()
        else:
            self.data_parallel_group = groups._get_data_parallel_group()

    def _configure_module(self, model: torch.nn.Module):
        if self.fp16_enabled():
            model = model.half()
        elif self.bfloat16_enabled():
            model = model.bfloat16()
        if not self.dont_change_device:
            from .zero.partition_parameters import is_zero_param
            if not any(is_zero_param(p) for p in model.parameters()):
                model = model.to(self.device)
        self.module = model
        self._configure_moe()
        self._broadcast_model()

    def _configure_moe(self):
        """Expert-parallel process groups for every MoE layer of the module (ref: engine.py _configure_distributed_model:
        ``module.set_deepspeed_parallelism``), and the bookkeeping the MoE checkpoint layout needs."""
        from ..moe.layer import MoE
        self.has_moe_layers, self.num_experts = False, []
        self._moe_layers = []
        for name, m in self.module.named_modules():
            if isinstance(m, MoE):
                self.has_moe_layers = True
                self.num_experts.append(m.num_experts)
                self._moe_layers.append((name, m))
                m.deepspeed_moe.wall_clock_breakdown = bool(self._config.wall_clock_breakdown)
                if dist.is_initialized():
                    m.set_deepspeed_parallelism(self._config.use_data_before_expert_parallel_)

    def _broadcast_model(self):
        if not dist.is_initialized() or self.dp_world_size_safe() == 1:
            return
        from .zero.partition_parameters import is_zero_param
        src = dist.get_global_rank(self.data_parallel_group, 0)
        for t in list(self.module.parameters()) + list(self.module.buffers()):
            if is_zero_param(t) or not torch.is_tensor(t) or t.numel() == 0:
                continue
            if getattr(t, "allreduce", True) is False:
                # expert parameters are replicated over their expert-data-parallel group only
                g = groups._get_expert_data_parallel_group(t.group_name) if hasattr(t, "group_name") and \
                    t.group_name in groups._get_expert_data_parallel_group_dict() else None
                if g is not None and dist.get_world_size(group=g) > 1:
                    dist.broadcast(t.data, dist.get_global_rank(g, 0), group=g)
                continue
            dist.broadcast(t.data, src, group=self.data_parallel_group)

    def dp_world_size_safe(self):
        return dist.get_world_size(group=self.data_parallel_group) if dist.is_initialized() else 1

    # ------------------------------------------------------------------------------------------------
    # optimizer
    # ------------------------------------------------------------------------------------------------
    def _configure_basic_optimizer(self, model_parameters):
        name = self._config.optimizer_name
        params = dict(self._config.optimizer_params)
        params.pop("max_grad_norm", None)
        if name in (ADAM_OPTIMIZER, ADAMW_OPTIMIZER):
            torch_adam = params.pop("torch_adam", False)
            adam_w_mode = params.pop("adam_w_mode", name == ADAMW_OPTIMIZER)
            if self.zero_cpu_offload() and self._config.zero_force_ds_cpu_optimizer:
                from ..ops.adam import DeepSpeedCPUAdam
                return DeepSpeedCPUAdam(model_parameters, **params, adamw_mode=adam_w_mode)
            if torch_adam:
                cls = torch.optim.AdamW if adam_w_mode else torch.optim.Adam
                return cls(model_parameters, **params)
            from ..ops.adam import FusedAdam
            return FusedAdam(model_parameters, **params, adam_w_mode=adam_w_mode)
        if name == ADAGRAD_OPTIMIZER:
            if self.zero_cpu_offload():
                from ..ops.adagrad import DeepSpeedCPUAdagrad
                return DeepSpeedCPUAdagrad(model_parameters, **params)
            return torch.optim.Adagrad(model_parameters, **params)
        if name == LAMB_OPTIMIZER:
            from ..ops.lamb import FusedLamb
            return FusedLamb(model_parameters, **params)
        if name == LION_OPTIMIZER:
            if self.zero_cpu_offload():
                from ..ops.lion import DeepSpeedCPULion
                return DeepSpeedCPULion(model_parameters, **params)
            from ..ops.lion import FusedLion
            return FusedLion(model_parameters, **params)
        if name == SGD_OPTIMIZER:
            return torch.optim.SGD(model_parameters, **params)
        if name == ONEBIT_ADAM_OPTIMIZER:
            from .fp16.onebit.adam import OnebitAdam
            return OnebitAdam(model_parameters, self, **params)
        if name == ZERO_ONE_ADAM_OPTIMIZER:
            from .fp16.onebit.zoadam import ZeroOneAdam
            return ZeroOneAdam(model_parameters, self, **params)
        if name == ONEBIT_LAMB_OPTIMIZER:
            from .fp16.onebit.lamb import OnebitLamb
            return OnebitLamb(model_parameters, self, **params)
        if isinstance(name, str) and hasattr(torch.optim, name):
            return getattr(torch.optim, name)(model_parameters, **params)
        raise ValueError(f"unknown optimizer type {name!r}")

    def _configure_optimizer(self, client_optimizer, model_parameters):
        if client_optimizer is None:
            basic = self._configure_basic_optimizer(model_parameters)
        elif isinstance(client_optimizer, torch.optim.Optimizer):
            basic = client_optimizer
        elif callable(client_optimizer):
            basic = client_optimizer(model_parameters)
        else:
            raise TypeError("optimizer must be a torch.optim.Optimizer or a callable returning one")
        self.basic_optimizer = basic
        stage = self.zero_optimization_stage()
        zc = self._config.zero_config
        clip = self.gradient_clipping()
        model_dtype, acc_dtype = self.get_data_types()
        if stage in (1, 2):
            self.optimizer = DeepSpeedZeroOptimizer(
                basic, param_names=self.param_names, timers=self.timers, static_loss_scale=self.loss_scale() or 1.0,
                dynamic_loss_scale=self.fp16_enabled() and self.dynamic_loss_scale(),
                dynamic_loss_args=self._config.dynamic_loss_scale_args, contiguous_gradients=zc.contiguous_gradients,
                reduce_bucket_size=zc.reduce_bucket_size, allgather_bucket_size=zc.allgather_bucket_size,
                dp_process_group=self.data_parallel_group, reduce_scatter=zc.reduce_scatter, overlap_comm=zc.overlap_comm,
                offload_optimizer_config=zc.offload_optimizer, mpu=self.mpu, clip_grad=clip,
                gradient_accumulation_dtype=acc_dtype if self.bfloat16_enabled() else torch.float32,
                communication_data_type=self._config.communication_data_type,
                postscale_gradients=self.postscale_gradients(), gradient_predivide_factor=self.gradient_predivide_factor(),
                gradient_accumulation_steps=self.gradient_accumulation_steps(), partition_grads=stage == 2,
                round_robin_gradients=zc.round_robin_gradients, verbose=self.global_rank == 0)
        elif stage == 3:
            cls = DeepSpeedZeroOptimizer_Stage3
            extra = {}
            if zc.mics_shard_size > 0:
                from .zero.mics import MiCS_Optimizer, create_mics_comm_groups
                cls = MiCS_Optimizer
                extra["mics_comm_groups"] = create_mics_comm_groups(zc.mics_shard_size, self.data_parallel_group, mpu=self.mpu)
            self.optimizer = cls(
                self.module, basic, param_names=self.param_names, timers=self.timers, ds_config=self._config,
                static_loss_scale=self.loss_scale() or 1.0,
                dynamic_loss_scale=self.fp16_enabled() and self.dynamic_loss_scale(),
                dynamic_loss_args=self._config.dynamic_loss_scale_args, dp_process_group=self.data_parallel_group,
                param_persistence_threshold=zc.param_persistence_threshold, offload_optimizer_config=zc.offload_optimizer,
                prefetch_bucket_size=zc.prefetch_bucket_size, max_live_parameters=zc.max_live_parameters,
                max_reuse_distance=zc.max_reuse_distance, aio_config=self._config.aio_config,
                offload_param_config=zc.offload_param, mpu=self.mpu, clip_grad=clip,
                zero_quantized_weights=zc.zero_quantized_weights, zero_quantized_gradients=zc.zero_quantized_gradients, zeropp_loco_param=zc.zeropp_loco_param,
                zero_quantized_nontrainable_weights=zc.zero_quantized_nontrainable_weights,
                gradient_predivide_factor=self.gradient_predivide_factor(), verbose=self.global_rank == 0, **extra)
        elif self.fp16_enabled():
            dyn = self.dynamic_loss_scale()
            if type(basic).__name__ in ("FusedLamb", "OnebitLamb"):
                self.optimizer = FP16_UnfusedOptimizer(basic, self, static_loss_scale=self.loss_scale() or 1.0,
                                                       dynamic_loss_scale=dyn,
                                                       dynamic_loss_args=self._config.dynamic_loss_scale_args,
                                                       mpu=self.mpu, clip_grad=clip)
            else:
                self.optimizer = FP16_Optimizer(basic, self, static_loss_scale=self.loss_scale() or 1.0,
                                                dynamic_loss_scale=dyn, initial_dynamic_scale=self._config.initial_dynamic_scale,
                                                dynamic_loss_args=self._config.dynamic_loss_scale_args, mpu=self.mpu,
                                                clip_grad=clip)
        elif self.bfloat16_enabled():
            self.optimizer = BF16_Optimizer(basic, param_names=self.param_names, mpu=self.mpu, clip_grad=clip,
                                            dp_process_group=self.data_parallel_group, timers=self.timers,
                                            grad_acc_dtype=acc_dtype,
                                            immediate_grad_update=self._config.bf16_config.immediate_grad_update)
        else:
            self.optimizer = basic
        for p in self.module.parameters():          # utils.safe_get_full_fp32_param & co find their optimizer through this
            p._ds_optimizer = self.optimizer
        log_dist(f"DeepSpeed optimizer = {type(self.optimizer).__name__} over {type(basic).__name__}", ranks=[0])

    def _configure_lr_scheduler(self, client_lr_scheduler):
        if client_lr_scheduler is not None:
            if callable(client_lr_scheduler) and not hasattr(client_lr_scheduler, "step"):
                self.lr_scheduler = client_lr_scheduler(self.basic_optimizer)
            else:
                self.lr_scheduler = client_lr_scheduler
        elif self._config.scheduler_name is not None:
            cls = lr_schedules.get_lr_schedule_class(self._config.scheduler_name)
            self.lr_scheduler = cls(self.basic_optimizer, **self._config.scheduler_params)
        log_dist(f"DeepSpeed LR scheduler = {type(self.lr_scheduler).__name__ if self.lr_scheduler else None}", ranks=[0])

    def _configure_checkpointing(self):
        self.checkpoint_engine = TorchCheckpointEngine()
        if (self.config.get("nebula") or {}).get("enabled"):
            self.checkpoint_engine = NebulaCheckpointEngine(self.config.get("nebula"))
        dp_rank = dist.get_rank(group=self.data_parallel_group) if dist.is_initialized() else 0
        self.save_non_zero_checkpoint = dp_rank == 0
        self.save_zero_checkpoint = self.zero_optimization() or (self.bfloat16_enabled() and self.optimizer is not None and
                                                                 isinstance(self.optimizer, BF16_Optimizer))

    # ------------------------------------------------------------------------------------------------
    # data
    # ------------------------------------------------------------------------------------------------
    def deepspeed_io(self, dataset, batch_size=None, route="train", pin_memory=True, data_sampler=None, collate_fn=None,
                     num_local_io_workers=None):
        if not isinstance(dataset, torch.utils.data.Dataset):
            raise ValueError("Training data must be a torch Dataset")
        if batch_size is None:
            batch_size = self.train_micro_batch_size_per_gpu()
        dp_ws = self.dp_world_size
        dp_rank = dist.get_rank(group=self.data_parallel_group) if dist.is_initialized() else 0
        loader_cfg = None
        de = self._config.data_efficiency_config or {}
        sampling = (de.get("data_sampling") or {}) if de.get("enabled", False) else {}
        if route == "train" and data_sampler is None and sampling.get("enabled", False):
            # data efficiency: curriculum-ordered sampling (ref: engine.py deepspeed_io + data_sampling/data_sampler.py)
            from .data_pipeline.data_sampler import DeepSpeedDataSampler
            data_sampler = DeepSpeedDataSampler(de, len(dataset), batch_size, dp_rank, dp_ws, self.data_parallel_group,
                                                self.gradient_accumulation_steps(), self.global_rank,
                                                drop_last=True)
            self.training_data_sampler = data_sampler
            loader_cfg = {"curriculum_learning_enabled": True}
        return DeepSpeedDataLoader(dataset=dataset, batch_size=batch_size, pin_memory=pin_memory,
                                   deepspeed_dataloader_config=loader_cfg,
                                   collate_fn=collate_fn or self.collate_fn, local_rank=self.local_rank,
                                   tput_timer=self.tput_timer if route == "train" else None,
                                   num_local_io_workers=num_local_io_workers, data_sampler=data_sampler,
                                   data_parallel_world_size=dp_ws, data_parallel_rank=dp_rank,
                                   dataloader_drop_last=self._config.dataloader_drop_last)

    # ------------------------------------------------------------------------------------------------
    # train loop pieces
    # ------------------------------------------------------------------------------------------------
    def train(self, mode: bool = True):
        self.warn_unscaled_loss = True
        self.module.train(mode)
        return self

    def eval(self):
        self.warn_unscaled_loss = True
        self.module.train(False)
        return self

    def is_gradient_accumulation_boundary(self) -> bool:
        if self._is_gradient_accumulation_boundary is None:
            return (self.micro_steps + 1) % self.gradient_accumulation_steps() == 0
        return self._is_gradient_accumulation_boundary

    def set_gradient_accumulation_boundary(self, is_boundary: bool):
        self._is_gradient_accumulation_boundary = is_boundary
        if hasattr(self.optimizer, "is_gradient_accumulation_boundary"):
            self.optimizer.is_gradient_accumulation_boundary = is_boundary

    def _scale_loss_by_gas(self, loss):
        gas = self.gradient_accumulation_steps()
        if isinstance(loss, torch.Tensor):
            return loss / gas
        if isinstance(loss, (tuple, list)) and isinstance(loss[0], torch.Tensor):
            return type(loss)(l / gas for l in loss)
        if self.warn_unscaled_loss:
            logger.warning(f"DeepSpeed unable to scale loss because of type: {type(loss)}")
            self.warn_unscaled_loss = False
        return loss

    def forward(self, *inputs, **kwargs):
        if self.flops_profiler is not None and self.global_steps == self._config.flops_profiler_config.profile_step and \
                self.global_rank == 0 and self.micro_steps % self.gradient_accumulation_steps() == 0:
            self.flops_profiler.start_profile(ignore_list=None)
        if self.progressive_layer_drop is not None and self.module.training:
            kwargs.update(self.progressive_layer_drop.get_state())
        if self.random_ltd_scheduler is not None and self.module.training:
            self.random_ltd_scheduler.update_seq(self.global_steps)
        if self.curriculum_scheduler_legacy is not None and self.module.training:
            self.curriculum_scheduler_legacy.update_difficulty(self.global_steps + 1)
            if self._config.curriculum_params_legacy.get("curriculum_type") == "seqlen":
                kwargs["curriculum_seqlen"] = self.curriculum_scheduler_legacy.get_current_difficulty()
        self.timers(FORWARD_MICRO_TIMER).start()
        self.timers(FORWARD_GLOBAL_TIMER).start()
        if self.training_dataloader is None and self.module.training:
            self.tput_timer.start()
        if self.amp_enabled():
            # "amp": apex's mixed precision in the reference. Here: fp32 master parameters + torch.autocast to bf16 (the
            # B200 default; no loss scaling needed). opt_level O0 keeps fp32; fp16 autocast is what the "fp16" block is for.
            with torch.autocast(device_type=self.device.type, dtype=torch.bfloat16, enabled=self._amp_autocast):
                loss = self.module(*inputs, **kwargs)
        else:
            loss = self.module(*inputs, **kwargs)
        self.timers(FORWARD_MICRO_TIMER).stop()
        self.timers(FORWARD_GLOBAL_TIMER).stop()
        if self.flops_profiler is not None and self.flops_profiler.started and self.flops_profiler.func_patched:
            self.flops_profiler.stop_profile()      # counting ends with the forward; the report waits for bwd / step times
        return loss

    def _uses_wrapper_backward(self) -> bool:
        return isinstance(self.optimizer, (DeepSpeedZeroOptimizer, DeepSpeedZeroOptimizer_Stage3, FP16_Optimizer,
                                           FP16_UnfusedOptimizer))

    def backward(self, loss, retain_graph: bool = False, scale_wrt_gas: bool = True):
        assert self.optimizer is not None or self.zero_optimization_stage() == 3, \
            "must provide optimizer during init in order to use backward"
        self.timers(BACKWARD_MICRO_TIMER).start()
        self.timers(BACKWARD_GLOBAL_TIMER).start()
        if self.gradient_accumulation_steps() > 1 and scale_wrt_gas:
            loss = self._scale_loss_by_gas(loss.float())
        if self.monitor.enabled and self.is_gradient_accumulation_boundary():
            self.losses = loss.detach() if self.losses is None else self.losses + loss.detach()
        elif self.monitor.enabled:
            self.losses = loss.detach() if self.losses is None else self.losses + loss.detach()
        if self._uses_wrapper_backward():
            self.optimizer.backward(loss, retain_graph=retain_graph)
        else:
            if self.eigenvalue is not None:
                loss.backward(create_graph=True, retain_graph=True)
            else:
                loss.backward(retain_graph=retain_graph)
        if self.enable_backward_allreduce and self.is_gradient_accumulation_boundary():
            self.allreduce_gradients()
        self.timers(BACKWARD_MICRO_TIMER).stop()
        self.timers(BACKWARD_GLOBAL_TIMER).stop()
        return loss

    def allreduce_gradients(self, bucket_size: int = MEMORY_OPT_ALLREDUCE_SIZE):
        """Average gradients over the data-parallel group at the accumulation boundary. ZeRO optimizers own their
        reduction (stage 1 at step, stage 2/3 per micro-batch); this is the plain-DP / fp16 / non-partitioned path."""
        if isinstance(self.optimizer, (DeepSpeedZeroOptimizer, DeepSpeedZeroOptimizer_Stage3)):
            return
        if not dist.is_initialized() or self.dp_world_size == 1:
            return
        self.buffered_allreduce_fallback(elements_per_buffer=bucket_size)

    def buffered_allreduce_fallback(self, grads=None, elements_per_buffer: int = MEMORY_OPT_ALLREDUCE_SIZE):
        by_dtype: Dict[torch.dtype, list] = {}
        expert: Dict[str, list] = {}
        sparse_ids = self._sparse_gradient_param_ids()
        for p in self.module.parameters():
            if not p.requires_grad:
                continue
            if id(p) in sparse_ids:
                # "sparse_gradients": embedding rows touched by this rank travel as (row indices, row values) and are
                # all-gathered; the averaged gradient stays sparse when autograd produced a sparse one
                self._sparse_allreduce_param(p)
                continue
            if p.grad is None:
                p.grad = torch.zeros_like(p)  # every rank must contribute the same buckets
            if getattr(p, "allreduce", True) is False and hasattr(p, "group_name"):
                expert.setdefault(p.group_name, []).append(p.grad.data)
            else:
                by_dtype.setdefault(p.grad.dtype, []).append(p.grad.data)
        for tensors in by_dtype.values():
            self._allreduce_buckets(tensors, self.data_parallel_group, self.dp_world_size, elements_per_buffer)
        for name, tensors in expert.items():  # expert grads average over the expert-data-parallel group only
            g = groups._get_expert_data_parallel_group(name)
            self._allreduce_buckets(tensors, g, dist.get_world_size(group=g), elements_per_buffer)

    def _sparse_gradient_param_ids(self):
        """Weights of ``nn.Embedding`` / ``nn.EmbeddingBag`` modules when ``"sparse_gradients": true`` (ref: engine.py
        sparse_tensor_module_names)."""
        ids = getattr(self, "_sparse_param_ids", None)
        if ids is None:
            ids = set()
            if self._config.sparse_gradients_enabled:
                for m in self.module.modules():
                    if isinstance(m, (torch.nn.Embedding, torch.nn.EmbeddingBag)):
                        ids.add(id(m.weight))
            self._sparse_param_ids = ids
        return ids

    @staticmethod
    def _all_gather_rows(t: torch.Tensor, group, ws: int):
        """All-gather tensors whose first dimension differs per rank: sizes first, then padded payloads."""
        n = torch.tensor([t.shape[0]], dtype=torch.long, device=t.device)
        sizes = [torch.zeros_like(n) for _ in range(ws)]
        dist.all_gather(sizes, n, group=group)
        sizes = [int(x) for x in sizes]
        cap = max(max(sizes), 1)
        padded = t.new_zeros((cap, ) + tuple(t.shape[1:]))
        padded[:t.shape[0]] = t
        parts = [torch.empty_like(padded) for _ in range(ws)]
        dist.all_gather(parts, padded, group=group)
        return [p[:k] for p, k in zip(parts, sizes)]

    def _sparse_allreduce_param(self, p):
        from .sparse_tensor import SparseTensor
        ws, group = self.dp_world_size, self.data_parallel_group
        was_sparse = p.grad is not None and p.grad.is_sparse
        grad = p.grad if p.grad is not None else torch.zeros_like(p)
        st = SparseTensor(grad.data)
        pre = self.gradient_predivide_factor() if self._config.prescale_gradients else ws
        values = st.values / pre
        st.indices = torch.cat(self._all_gather_rows(st.indices, group, ws))
        st.values = torch.cat(self._all_gather_rows(values, group, ws))
        if pre != ws:
            st.values = st.values * (pre / ws)
        self.sparse_allreduce_stats = {"rows_sent": int(st.indices.numel()), "dense_rows": int(p.shape[0])}
        p.grad = st.to_coo_tensor().coalesce() if was_sparse else st.to_dense().to(p.dtype)

    def _allreduce_buckets(self, tensors, group, ws, elements_per_buffer):
        if ws == 1:
            return
        bucket, n = [], 0
        for t in tensors + [None]:
            if t is not None and (n + t.numel() <= elements_per_buffer or not bucket):
                bucket.append(t)
                n += t.numel()
                continue
            flat = flatten_dense_tensors(bucket)
            if self._config.prescale_gradients:
                flat.div_(self.gradient_predivide_factor())
                dist.all_reduce(flat, group=group)
                if self.gradient_predivide_factor() != ws:
                    flat.mul_(self.gradient_predivide_factor() / ws)
            else:
                dist.all_reduce(flat, group=group)
                flat.div_(ws)
            for dst, src in zip(bucket, unflatten_dense_tensors(flat, bucket)):
                dst.copy_(src)
            bucket, n = ([t], t.numel()) if t is not None else ([], 0)

    def clip_fp32_gradients(self):
        clip_grad_norm_(parameters=self.module.parameters(), max_norm=self.gradient_clipping(), mpu=self.mpu)

    def _take_model_step(self, lr_kwargs=None):
        plain = self.optimizer is self.basic_optimizer
        if plain:
            self._global_grad_norm = get_grad_norm(self.module.parameters(), mpu=self.mpu)
            if self.gradient_clipping() > 0.0:
                self.clip_fp32_gradients()
        self.optimizer.step()
        if not plain and hasattr(self.optimizer, "get_global_grad_norm"):
            self._global_grad_norm = self.optimizer.get_global_grad_norm()
        overflow = bool(getattr(self.optimizer, "overflow", False))
        if hasattr(self.optimizer, "zero_grad"):
            self.optimizer.zero_grad()
        if plain:
            for p in self.module.parameters():
                p.grad = None
        if overflow:
            self.skipped_steps += 1
        elif self.lr_scheduler is not None:
            try:
                self.lr_scheduler.step(**(lr_kwargs or {}))
            except TypeError:
                self.lr_scheduler.step()
        self.global_steps += 1
        self.global_samples += self.train_batch_size()
        if self.compression_scheduler is not None:
            self.compression_scheduler.step()
        self._maybe_launch_compile_passes()
        self._autotuning_hook()

    def _autotuning_hook(self):
        """Experiment runs of the autotuner (``"autotuning": {"enabled": true, "metric_path": ...}``, written by
        ``autotuning/autotuner.py``): time the optimizer steps between ``start_profile_step`` and ``end_profile_step``, write
        the metrics file and leave — the tuner only needs the measurement (ref: engine.py autotuning_profile / exit())."""
        at = self._config.autotuning_config or {}
        if not at.get("enabled"):
            return
        import json
        import time
        info_path = at.get("model_info_path")
        if info_path and self.global_steps == 1 and self.global_rank == 0:
            n = sum(getattr(p, "ds_numel", p.numel()) for p in self.module.parameters())
            os.makedirs(os.path.dirname(info_path) or ".", exist_ok=True)
            with open(info_path, "w") as f:
                json.dump({"num_params": int(n), "trainable_params": int(sum(getattr(p, "ds_numel", p.numel())
                                                                            for p in self.module.parameters() if p.requires_grad))}, f)
            if at.get("model_info", {}) and at["model_info"].get("profile"):
                raise SystemExit(0)
        start, end = int(at.get("start_profile_step", 3)), int(at.get("end_profile_step", 5))
        if self.device.type == "cuda":
            torch.cuda.synchronize()
        if self.global_steps == start:
            self._autotune_t0 = time.perf_counter()
        elif self.global_steps >= end and getattr(self, "_autotune_t0", None) is not None:
            dt = (time.perf_counter() - self._autotune_t0) / max(1, end - start)
            if self.global_rank == 0 and at.get("metric_path"):
                os.makedirs(os.path.dirname(at["metric_path"]) or ".", exist_ok=True)
                with open(at["metric_path"], "w") as f:
                    json.dump({"latency": dt * 1e3, "throughput": self.train_batch_size() / dt,
                               "train_micro_batch_size_per_gpu": self.train_micro_batch_size_per_gpu(),
                               "zero_stage": self.zero_optimization_stage()}, f)
            if dist.is_initialized():
                dist.barrier()
            raise SystemExit(0)

    # ---- schedule compiler (the reference's engine.compile / DeepCompile, engine.py:3890+) ---------------------------
    def is_deepcompile_enabled(self) -> bool:
        return bool((self._config.compile_config or {}).get("deepcompile", False))

    @property
    def is_compiled(self) -> bool:
        return getattr(self, "_is_compiled", False)

    def compile(self, backend=None, compile_kwargs=None, schedule=None) -> None:
        """Arm the schedule compiler: after the next optimizer step (one profiled iteration) the passes run once and
        their result (prefetch plan, persistent parameters, optimizer-state offload) is applied to this engine.
        ``backend`` / ``compile_kwargs`` are accepted for signature parity; no tracing compiler is involved."""
        from ..compile import init_z1, init_z3
        if self.is_compiled:
            return
        if self.zero_optimization_stage() == 3:
            self._compile_fn = init_z3(self, backend, self._config.compile_config, compile_kwargs, schedule)
        elif self.zero_optimization_stage() in (1, 2):
            self._compile_fn = init_z1(self, backend, self._config.compile_config, compile_kwargs, schedule,
                                       use_z2=self.zero_optimization_stage() == 2)
        else:
            raise RuntimeError("engine.compile() schedules ZeRO collectives: it needs zero_optimization.stage >= 1")
        self._is_compiled = True
        self.compile_report = None

    def _maybe_launch_compile_passes(self):
        fn = getattr(self, "_compile_fn", None)
        if fn is None:
            return
        off = getattr(self.optimizer, "parameter_offload", None)
        if off is not None:
            off._begin_iteration(None, None)       # close the recorded iteration
            if not off.trace_complete:
                return
        self.compile_report = fn()
        self._compile_fn = None

    def step(self, lr_kwargs=None):
        assert self.optimizer is not None, "must provide optimizer during init in order to use step"
        self.timers(STEP_MICRO_TIMER).start()
        self.timers(STEP_GLOBAL_TIMER).start()
        report = False
        if self.is_gradient_accumulation_boundary():
            self.gas_boundary_ctr = getattr(self, "gas_boundary_ctr", 0) + 1
            if self.progressive_layer_drop is not None:
                self.progressive_layer_drop.update_state(self.global_steps)
            if self.eigenvalue is not None and self.gas_boundary_ctr % self.eigenvalue.gas_boundary_resolution == 0:
                self.block_eigenvalue = self.eigenvalue.compute_eigenvalue(self.module, self.device)
            self._take_model_step(lr_kwargs)
            report = True
        self.tput_timer.stop(global_step=self.is_gradient_accumulation_boundary(), report_speed=report)
        self.timers(STEP_MICRO_TIMER).stop()
        self.timers(STEP_GLOBAL_TIMER).stop()
        if self.flops_profiler is not None and self.flops_profiler.started and report:
            fc = self._config.flops_profiler_config
            self.flops_profiler.print_model_profile(profile_step=fc.profile_step, module_depth=fc.module_depth,
                                                    top_modules=fc.top_modules, detailed=fc.detailed, output_file=fc.output_file)
            self.flops_profiler.end_profile()
        if report:
            if self.monitor.enabled and self.global_rank == 0:
                events = [("Train/Samples/lr", self.get_lr()[0], self.global_samples)]
                if self.losses is not None:
                    events.append(("Train/Samples/train_loss", float(self.losses), self.global_samples))
                if self.fp16_enabled() and hasattr(self.optimizer, "cur_scale"):
                    events.append(("Train/Samples/loss_scale", self.optimizer.cur_scale, self.global_samples))
                self.monitor.write_events(events)
            self.losses = None
            if self.steps_per_print() and self.global_steps % self.steps_per_print() == 0:
                log_dist(f"step={self.global_steps}, skipped={self.skipped_steps}, lr={self.get_lr()}, "
                         f"mom={self.get_mom()}", ranks=[0])
                if self.wall_clock_breakdown():
                    self.timers.log(self.engine_timers.global_timers, normalizer=self.steps_per_print() or 1)
        self.micro_steps += 1

    def no_sync(self):
        """Context manager that suppresses gradient reduction in ``backward`` (ZeRO stages >= 2 cannot skip it)."""
        assert self.zero_optimization_stage() < 2, "no_sync is incompatible with ZeRO stage 2/3 gradient partitioning"
        engine = self

        class _NoSync:
            def __enter__(self_inner):
                self_inner.prev = engine.enable_backward_allreduce
                engine.enable_backward_allreduce = False

            def __exit__(self_inner, *exc):
                engine.enable_backward_allreduce = self_inner.prev
                return False

        return _NoSync()

    def zero_grad(self):
        for p in self.module.parameters():
            p.grad = None

    def get_lr(self):
        return [g["lr"] for g in self.optimizer.param_groups] if self.optimizer is not None else []

    def get_mom(self):
        out = []
        for g in (self.optimizer.param_groups if self.optimizer is not None else []):
            out.append(g["betas"] if "betas" in g else g.get("momentum", 0.0))
        return out

    def get_type(self):
        return [g.get("type") for g in self.optimizer.param_groups]

    def get_global_grad_norm(self):
        return self._global_grad_norm

    def get_loss_scale(self):
        return getattr(self.optimizer, "cur_scale", 1.0)

    def was_step_applied(self) -> bool:
        return not bool(getattr(self.optimizer, "overflow", False))

    def get_batch_info(self):
        return self.train_batch_size(), self.train_micro_batch_size_per_gpu(), self.gradient_accumulation_steps()

    def destroy(self):
        if isinstance(self.optimizer, DeepSpeedZeroOptimizer_Stage3):
            self.optimizer.parameter_offload.destroy()

    # ------------------------------------------------------------------------------------------------
    # checkpointing
    # ------------------------------------------------------------------------------------------------
    def _mp_rank(self) -> int:
        return self.mpu.get_model_parallel_rank() if self.mpu is not None else 0

    def _get_ckpt_name(self, checkpoints_path, tag, mp_placeholder=None):
        mp = mp_placeholder if mp_placeholder is not None else f"{self._mp_rank():02d}"
        if self.zero_optimization_stage() == 3:
            dp_rank = dist.get_rank(group=self.data_parallel_group) if dist.is_initialized() else 0
            return os.path.join(checkpoints_path, str(tag), f"zero_pp_rank_{dp_rank}_mp_rank_{mp}_model_states.pt")
        return os.path.join(checkpoints_path, str(tag), f"mp_rank_{mp}_model_states.pt")

    def _get_zero_ckpt_name(self, checkpoints_path, tag, dp_rank=None):
        if dp_rank is None:
            dp_rank = dist.get_rank(group=self.data_parallel_group) if dist.is_initialized() else 0
        prefix = "bf16_" if isinstance(self.optimizer, BF16_Optimizer) else ""
        return os.path.join(checkpoints_path, str(tag), f"{prefix}zero_pp_rank_{dp_rank}_mp_rank_{self._mp_rank():02d}_optim_states.pt")

    def module_state_dict(self, exclude_frozen_parameters: bool = False):
        sd = self.module.state_dict()
        if self.random_ltd_scheduler is not None:   # checkpoints carry the unwrapped layer names
            from .data_pipeline.random_ltd import remove_random_ltd_state_dict
            sd = remove_random_ltd_state_dict(sd)
        if exclude_frozen_parameters:
            frozen = {n for n, p in self.module.named_parameters() if not p.requires_grad}
            sd = {k: v for k, v in sd.items() if k not in frozen}
        return sd

    def _checkpoint_tag_validation(self, tag):
        if not self._config.checkpoint_tag_validation_enabled or not dist.is_initialized() or self.world_size == 1:
            return
        import hashlib
        h = int(hashlib.sha1(str(tag).encode()).hexdigest()[:12], 16)
        t_min = torch.tensor([float(h)], dtype=torch.float64, device=self.device if self.device.type == "cuda" else "cpu")
        t_max = t_min.clone()
        dist.all_reduce(t_min, op=dist.ReduceOp.MIN)
        dist.all_reduce(t_max, op=dist.ReduceOp.MAX)
        if float(t_min) != float(t_max):
            msg = f"[rank={self.global_rank}] The checkpoint tag name '{tag}' is not consistent across all ranks."
            if self._config.checkpoint_tag_validation_fail:
                raise AssertionError(msg)
            logger.warning(msg)

    def save_checkpoint(self, save_dir, tag=None, client_state=None, save_latest: bool = True,
                        exclude_frozen_parameters: bool = False):
        client_state = client_state or {}
        if tag is None:
            tag = f"global_step{self.global_steps}"
        tag = str(tag)
        self._checkpoint_tag_validation(tag)
        self.checkpoint_engine.create(tag)
        if self.global_rank == 0:
            self.checkpoint_engine.makedirs(os.path.join(save_dir, tag), exist_ok=True)
        if dist.is_initialized():
            dist.barrier()
        if self.has_moe_layers and self.zero_optimization_stage() < 3:
            self._save_moe_experts(os.path.join(save_dir, tag))
            zero_wrapped = isinstance(self.optimizer, (DeepSpeedZeroOptimizer, DeepSpeedZeroOptimizer_Stage3))
            if self.optimizer is not None and not zero_wrapped and self._moe_layers:
                # plain optimizers hold the state of THIS rank's experts: one file per expert-parallel rank of the largest group
                big = max(self._moe_expert_prefixes(), key=lambda t: t[5])
                if big[3]:
                    self.checkpoint_engine.save({"optimizer": self.optimizer.state_dict()},
                                                self._expp_optim_ckpt_name(os.path.join(save_dir, tag), big[2]))
        if self.save_non_zero_checkpoint or self.zero_optimization_stage() == 3:
            self._save_checkpoint(save_dir, tag, client_state, exclude_frozen_parameters)
        if self.save_zero_checkpoint:
            self._save_zero_checkpoint(save_dir, tag)
        self.checkpoint_engine.commit(tag)
        if dist.is_initialized():
            dist.barrier()
        if save_latest and self.global_rank == 0:
            with open(os.path.join(save_dir, "latest"), "w") as f:
                f.write(tag)
        return True

    def _save_checkpoint(self, save_dir, tag, client_state, exclude_frozen_parameters=False):
        zero_wrapped = isinstance(self.optimizer, (DeepSpeedZeroOptimizer, DeepSpeedZeroOptimizer_Stage3))
        frozen_shapes = frozen_fragments = None
        if self.zero_optimization_stage() == 3:
            module_sd = {}  # trainable parameters live in the optimizer's fp32 partitions (zero_pp_rank_*_optim_states.pt)
            buffers = {k: v for k, v in self.module.state_dict().items() if k not in dict(self.module.named_parameters())}
            # frozen parameters (LoRA-style / partial fine-tuning) are in no optimizer group: every rank saves ITS slice of each
            # one next to the shapes, the layout of the reference's zero_pp_rank_*_model_states.pt (engine.py:3508-3560 there)
            if not exclude_frozen_parameters:
                frozen_shapes, frozen_fragments = {}, {}
                for name, p in self.module.named_parameters():
                    if not p.requires_grad and hasattr(p, "ds_tensor"):
                        frozen_shapes[name] = torch.Size(p.ds_shape)
                        from .zero.partition_parameters import read_partition, release_partition
                        frozen_fragments[name] = read_partition(p).detach().cpu().clone()
                        release_partition(p)
        else:
            module_sd = self.module_state_dict(exclude_frozen_parameters)
            if self.has_moe_layers:
                module_sd = self._strip_expert_keys(module_sd)      # experts live in their own files
            buffers = None
        # tied parameters: alias name -> the name under which the tensor is stored
        seen, shared = {}, {}
        for name, p in self.module.named_parameters(remove_duplicate=False):
            key = getattr(p, "ds_id", None) if hasattr(p, "ds_id") else p.data_ptr()
            if key in seen and seen[key] != name:
                shared[name] = seen[key]
            else:
                seen[key] = name
        state = dict(module=module_sd, buffer_names=list(dict(self.module.named_buffers()).keys()), buffers=buffers,
                     optimizer=self.optimizer.state_dict() if (self.optimizer is not None and not zero_wrapped) else None,
                     param_shapes=self.optimizer._param_shapes() if zero_wrapped else None,
                     frozen_param_shapes=frozen_shapes, frozen_param_fragments=frozen_fragments, shared_params=shared,
                     lr_scheduler=self.lr_scheduler.state_dict() if self.lr_scheduler is not None else None,
                     data_sampler=self.training_data_sampler.state_dict() if getattr(self, "training_data_sampler", None) is not None else None,
                     random_ltd=self.random_ltd_scheduler.state_dict() if self.random_ltd_scheduler is not None else None,
                     skipped_steps=self.skipped_steps, global_steps=self.global_steps, global_samples=self.global_samples,
                     dp_world_size=self.dp_world_size, mp_world_size=self.mp_world_size,
                     ds_config=self._config._param_dict, ds_version=_version())
        state.update(client_state)
        self.checkpoint_engine.save(state, self._get_ckpt_name(save_dir, tag))

    # ---- mixture-of-experts checkpoint layout (ref: engine.py _save_moe_checkpoint / load_moe_state_dict) -----------------------
    # layer_<L>_expert_<E>_mp_rank_<MM>_model_states.pt : one file per GLOBAL expert, written by the rank of its expert-data-
    #                                                      parallel group that has rank 0 there
    # expp_rank_<R>_mp_rank_<MM>_optim_states.pt        : the (non-ZeRO) optimizer state of expert-parallel rank R
    # mp_rank_<MM>_model_states.pt                       : everything else (expert tensors removed)
    @staticmethod
    def _expert_ckpt_name(ckpt_dir, layer_id, expert_id, mp_rank=0):
        return os.path.join(ckpt_dir, f"layer_{layer_id}_expert_{expert_id}_mp_rank_{mp_rank:02d}_model_states.pt")

    @staticmethod
    def _expp_optim_ckpt_name(ckpt_dir, expp_rank, mp_rank=0):
        return os.path.join(ckpt_dir, f"expp_rank_{expp_rank}_mp_rank_{mp_rank:02d}_optim_states.pt")

    def _moe_expert_prefixes(self):
        """[(layer id, layer name, ep rank, is writer, ModuleList of local experts, ep size)] for this rank."""
        out = []
        for layer_id, (name, m) in enumerate(self._moe_layers):
            gname = m.expert_group_name
            ep_rank = dist.get_rank(group=groups._get_expert_parallel_group(gname)) if dist.is_initialized() else 0
            edp_rank = dist.get_rank(group=groups._get_expert_data_parallel_group(gname)) if dist.is_initialized() else 0
            out.append((layer_id, name, ep_rank, edp_rank == 0, m.deepspeed_moe.experts.deepspeed_experts, m.ep_size))
        return out

    def _save_moe_experts(self, ckpt_dir):
        mp_rank = self.mpu.get_model_parallel_rank() if self.mpu is not None else 0
        for layer_id, _, ep_rank, writer, experts, _ in self._moe_expert_prefixes():
            if not writer:
                continue
            for local_id, expert in enumerate(experts):
                gid = ep_rank * len(experts) + local_id
                sd = {k: v.detach().cpu().clone() for k, v in expert.state_dict().items()}
                self.checkpoint_engine.save(sd, self._expert_ckpt_name(ckpt_dir, layer_id, gid, mp_rank))

    def _load_moe_experts(self, ckpt_dir, strict: bool = True):
        mp_rank = self.mpu.get_model_parallel_rank() if self.mpu is not None else 0
        for layer_id, name, ep_rank, _, experts, _ in self._moe_expert_prefixes():
            for local_id, expert in enumerate(experts):
                gid = ep_rank * len(experts) + local_id
                path = self._expert_ckpt_name(ckpt_dir, layer_id, gid, mp_rank)
                if not os.path.exists(path):
                    if strict:
                        raise FileNotFoundError(f"expert checkpoint {path} is missing (layer {name}, global expert {gid})")
                    continue
                expert.load_state_dict(self.checkpoint_engine.load(path, map_location="cpu"), strict=strict)

    def _strip_expert_keys(self, module_sd):
        drop = [f"{name}.deepspeed_moe.experts.deepspeed_experts." for name, _ in self._moe_layers]
        return {k: v for k, v in module_sd.items() if not any(k.startswith(d) for d in drop)}

    def _save_zero_checkpoint(self, save_dir, tag):
        sd = dict(optimizer_state_dict=self.optimizer.state_dict(), ds_config=self._config._param_dict, ds_version=_version())
        self.checkpoint_engine.save(sd, self._get_zero_ckpt_name(save_dir, tag))

    def load_checkpoint(self, load_dir, tag=None, load_module_strict: bool = True, load_optimizer_states: bool = True,
                        load_lr_scheduler_states: bool = True, load_module_only: bool = False, custom_load_fn=None):
        if self._config.load_universal_checkpoint:
            return self._load_universal_checkpoint(load_dir, tag, load_optimizer_states, load_lr_scheduler_states, load_module_only)
        if tag is None:
            latest = os.path.join(load_dir, "latest")
            if not os.path.isfile(latest):
                logger.warning(f"Unable to find latest file at {latest}, if trying to load latest checkpoint please "
                               "ensure this file exists or pass an explicit checkpoint tag when loading a checkpoint.")
                return None, None
            with open(latest) as f:
                tag = f.read().strip()
        path = self._get_ckpt_name(load_dir, tag)
        if not os.path.exists(path):
            logger.warning(f"Client provided checkpoint load path: {path} does not exist")
            return None, None
        ckpt = self.checkpoint_engine.load(path, map_location="cpu")
        if self.zero_optimization_stage() != 3:
            if custom_load_fn is not None:
                custom_load_fn(src=ckpt["module"], dst=self.module)
            elif ckpt.get("module") is None and hasattr(self.module, "load_state_dir"):
                self.module.load_state_dir(os.path.join(load_dir, str(tag)), self.checkpoint_engine, strict=load_module_strict)
            elif self.has_moe_layers:
                missing, unexpected = self.module.load_state_dict(ckpt["module"], strict=False)
                expert_keys = set(self.module.state_dict()) - set(self._strip_expert_keys(self.module.state_dict()))
                bad = [k for k in missing if k not in expert_keys] + list(unexpected)
                if bad and load_module_strict:
                    raise RuntimeError(f"checkpoint does not match the module: {bad[:8]}")
                self._load_moe_experts(os.path.join(load_dir, str(tag)), strict=load_module_strict)
            else:
                self.load_module_state_dict(ckpt, strict=load_module_strict)
        else:
            own = dict(self.module.named_buffers())
            for k, v in (ckpt.get("buffers") or {}).items():
                if k in own:
                    own[k].data.copy_(v)
            # frozen parameters: this rank's slice goes back into its partition (and into the live tensor when resident)
            frags = ckpt.get("frozen_param_fragments") or {}
            if frags:
                from .zero.partition_parameters import ZeroParamStatus, all_gather_params
                params = dict(self.module.named_parameters())
                for name, frag in frags.items():
                    p = params.get(name)
                    if p is None or not hasattr(p, "ds_tensor"):
                        if load_module_strict:
                            raise KeyError(f"frozen parameter {name} of the checkpoint is not a ZeRO-3 parameter of this module")
                        continue
                    from .zero.partition_parameters import write_partition
                    write_partition(p, frag)
                    if p.ds_status == ZeroParamStatus.AVAILABLE:
                        p.ds_status = ZeroParamStatus.NOT_AVAILABLE
                        all_gather_params([p])
        self.loaded_checkpoint_dp_world_size = ckpt.get("dp_world_size")
        self.loaded_checkpoint_mp_world_size = ckpt.get("mp_world_size")
        if not load_module_only:
            self.global_steps = ckpt.get("global_steps", 0)
            self.global_samples = ckpt.get("global_samples", self.global_steps * self.train_batch_size())
            self.skipped_steps = ckpt.get("skipped_steps", 0)
            zero_wrapped = isinstance(self.optimizer, (DeepSpeedZeroOptimizer, DeepSpeedZeroOptimizer_Stage3))
            if zero_wrapped:
                zpath = self._get_zero_ckpt_name(load_dir, tag)
                if os.path.exists(zpath):
                    zsd = self.checkpoint_engine.load(zpath, map_location="cpu")
                    self.optimizer.load_state_dict(zsd["optimizer_state_dict"], load_optimizer_states=load_optimizer_states)
                elif load_optimizer_states:
                    logger.warning(f"ZeRO optimizer state {zpath} not found; optimizer state was not restored")
            elif load_optimizer_states and self.optimizer is not None and ckpt.get("optimizer") is not None:
                osd = ckpt["optimizer"]
                if self.has_moe_layers and self._moe_layers:   # this expert-parallel rank's own optimizer state
                    big = max(self._moe_expert_prefixes(), key=lambda t: t[5])
                    own_path = self._expp_optim_ckpt_name(os.path.join(load_dir, str(tag)), big[2])
                    if os.path.exists(own_path):
                        osd = self.checkpoint_engine.load(own_path, map_location="cpu")["optimizer"]
                self.optimizer.load_state_dict(osd)
            if zero_wrapped and not load_optimizer_states and self.zero_optimization_stage() != 3 and \
                    hasattr(self.optimizer, "refresh_fp32_params"):   # (stage 3 restored its fp32 masters from the shard file)
                self.optimizer.refresh_fp32_params()
            if load_lr_scheduler_states and self.lr_scheduler is not None and ckpt.get("lr_scheduler") is not None:
                self.lr_scheduler.load_state_dict(ckpt["lr_scheduler"])
            if ckpt.get("data_sampler") is not None and getattr(self, "training_data_sampler", None) is not None:
                self.training_data_sampler.load_state_dict(ckpt["data_sampler"])
            if ckpt.get("random_ltd") is not None and self.random_ltd_scheduler is not None:
                self.random_ltd_scheduler.load_state_dict(ckpt["random_ltd"])
        elif self.zero_optimization_stage() == 3:
            # module-only load under ZeRO-3: the trainable weights exist only as fp32 partitions in the optimizer shard
            zpath = self._get_zero_ckpt_name(load_dir, tag)
            if not os.path.exists(zpath):
                raise FileNotFoundError(f"load_module_only under ZeRO-3 needs {zpath}: the model-states file holds no trainable weights")
            zsd = self.checkpoint_engine.load(zpath, map_location="cpu")
            self.optimizer.load_state_dict(zsd["optimizer_state_dict"], load_optimizer_states=False)
        elif hasattr(self.optimizer, "refresh_fp32_params"):
            self.optimizer.refresh_fp32_params()
        reserved = {"module", "buffer_names", "buffers", "optimizer", "param_shapes", "frozen_param_shapes",
                    "frozen_param_fragments", "shared_params", "lr_scheduler", "skipped_steps", "data_sampler", "random_ltd",
                    "global_steps", "global_samples", "dp_world_size", "mp_world_size", "ds_config", "ds_version"}
        client_state = {k: v for k, v in ckpt.items() if k not in reserved}
        return os.path.join(load_dir, str(tag)), client_state

    def _load_universal_checkpoint(self, load_dir, tag, load_optimizer_states, load_lr_scheduler_states, load_module_only):
        """``"checkpoint": {"load_universal": true}``: ``load_dir/tag`` (tag defaults to the content of ``latest_universal``)
        is a folder written by ``ds_to_universal``; every rank cuts ITS slices out of the per-parameter files, so the run
        may use another data-parallel degree / ZeRO stage (1, 2 or none) than the one that saved (ref: engine.py
        load_checkpoint -> _load_zero_checkpoint with load_universal_checkpoint())."""
        from ..checkpoint.universal_checkpoint import load_universal_into_engine
        if self.zero_optimization_stage() == 3:
            raise NotImplementedError("universal checkpoints load into ZeRO stage 0 / 1 / 2 (and BF16) optimizers")
        if tag is None:
            latest = os.path.join(load_dir, "latest_universal")
            if not os.path.isfile(latest):
                logger.warning(f"Unable to find {latest}; run ds_to_universal first or pass the universal tag")
                return None, None
            with open(latest) as f:
                tag = f.read().strip()
        folder = os.path.join(load_dir, str(tag))
        if not os.path.isdir(os.path.join(folder, "zero")):
            logger.warning(f"{folder} is not a universal checkpoint (no zero/ folder)")
            return None, None
        n = load_universal_into_engine(self, folder, load_optimizer_states=load_optimizer_states and not load_module_only)
        log_dist(f"universal checkpoint {folder}: {n} parameters restored", ranks=[0])
        client_state = {}
        import glob
        files = sorted(glob.glob(os.path.join(folder, "*model_states.pt")))
        if files:
            ckpt = self.checkpoint_engine.load(files[0], map_location="cpu")
            own = dict(self.module.named_buffers())
            for k in ckpt.get("buffer_names") or []:
                if k in own and k in (ckpt.get("module") or {}):
                    own[k].data.copy_(ckpt["module"][k])
            if not load_module_only:
                self.global_steps = ckpt.get("global_steps", 0)
                self.global_samples = ckpt.get("global_samples", self.global_steps * self.train_batch_size())
                self.skipped_steps = ckpt.get("skipped_steps", 0)
                if load_lr_scheduler_states and self.lr_scheduler is not None and ckpt.get("lr_scheduler") is not None:
                    self.lr_scheduler.load_state_dict(ckpt["lr_scheduler"])
            reserved = {"module", "buffer_names", "buffers", "optimizer", "param_shapes", "frozen_param_shapes",
                        "frozen_param_fragments", "shared_params", "lr_scheduler", "skipped_steps", "global_steps",
                        "global_samples", "dp_world_size", "mp_world_size", "ds_config", "ds_version", "universal_checkpoint_info"}
            client_state = {k: v for k, v in ckpt.items() if k not in reserved}
        return folder, client_state

    def _zero3_consolidated_16bit_state_dict(self, exclude_frozen_parameters: bool = False):
        from .zero.partition_parameters import GatheredParameters
        sd = {}
        for name, p in self.module.named_parameters():
            if exclude_frozen_parameters and not p.requires_grad:
                continue
            with GatheredParameters([p]):
                if self.global_rank == 0:
                    sd[name] = p.data.detach().cpu().clone()
        for name, b in self.module.named_buffers():
            if self.global_rank == 0:
                sd[name] = b.detach().cpu().clone()
        return sd if self.global_rank == 0 else None

    def save_16bit_model(self, save_dir, save_filename="pytorch_model.bin", exclude_frozen_parameters: bool = False):
        """Write a consolidated 16-bit ``state_dict`` from rank 0 (gathers ZeRO-3 partitions)."""
        if self.zero_optimization_stage() == 3:
            sd = self._zero3_consolidated_16bit_state_dict(exclude_frozen_parameters)
        else:
            sd = self.module_state_dict(exclude_frozen_parameters)
        if self.global_rank == 0:
            os.makedirs(save_dir, exist_ok=True)
            self.checkpoint_engine.save(sd, os.path.join(save_dir, save_filename))
        return True

    def save_fp16_model(self, save_dir, save_filename="pytorch_model.bin"):
        """Older name of :meth:`save_16bit_model`."""
        return self.save_16bit_model(save_dir, save_filename)


def _version() -> str:
    from ..version import __version__
    return __version__


def initialize(args=None, model: torch.nn.Module = None, optimizer=None, model_parameters=None, training_data=None,
               lr_scheduler=None, distributed_port: int = 29500, mpu=None, dist_init_required: Optional[bool] = None,
               collate_fn=None, config=None, mesh_param=None, config_params=None):
    """Returns ``(engine, optimizer, training_dataloader, lr_scheduler)`` — ``deepspeed.initialize``
    (``deepspeed/__init__.py:69-235``). ``config`` is a dict or a JSON path (``args.deepspeed_config`` also works)."""
    assert model is not None, "deepspeed.initialize requires a model"
    if config is None:
        config = config_params
    if config is None and args is not None:
        config = getattr(args, "deepspeed_config", None) or getattr(args, "deepscale_config", None)
    assert config is not None, "DeepSpeed requires --deepspeed_config to specify configuration file"
    if not dist.is_initialized() and int(os.environ.get("WORLD_SIZE", "1")) > 1 and dist_init_required is not False:
        dist.init_distributed(distributed_port=distributed_port)
    from .pipe.module import PipelineModule
    cfg = DeepSpeedConfig(config, mpu=mpu if not isinstance(model, PipelineModule) else model.mpu())
    if isinstance(model, PipelineModule):
        from .pipe.engine import PipelineEngine
        engine = PipelineEngine(args=args, model=model, optimizer=optimizer, model_parameters=model_parameters,
                                training_data=training_data, lr_scheduler=lr_scheduler, mpu=model.mpu(),
                                dist_init_required=dist_init_required, collate_fn=collate_fn, config=config,
                                config_class=cfg)
    elif cfg.hybrid_engine.get("enabled", False):
        from .hybrid_engine import DeepSpeedHybridEngine
        engine = DeepSpeedHybridEngine(args=args, model=model, optimizer=optimizer, model_parameters=model_parameters,
                                       training_data=training_data, lr_scheduler=lr_scheduler, mpu=mpu,
                                       dist_init_required=dist_init_required, collate_fn=collate_fn, config=config,
                                       config_class=cfg)
    else:
        engine = DeepSpeedEngine(args=args, model=model, optimizer=optimizer, model_parameters=model_parameters,
                                 training_data=training_data, lr_scheduler=lr_scheduler, mpu=mpu,
                                 dist_init_required=dist_init_required, collate_fn=collate_fn, config=config,
                                 config_class=cfg)
    return engine, engine.optimizer, engine.training_dataloader, engine.lr_scheduler


def split_half_float_double_sparse(tensors):
    """Gradients grouped for reduction: ([(dtype, [SparseTensor...])...], [(dtype, [dense tensor...])...])."""
    from .sparse_tensor import SparseTensor
    sparse, dense = [], []
    for dtype in (torch.float16, torch.bfloat16, torch.float32, torch.float64):
        sp = [t for t in tensors if isinstance(t, SparseTensor) and t.dtype == dtype]
        de = [t for t in tensors if not isinstance(t, SparseTensor) and t.dtype == dtype]
        if sp:
            sparse.append((dtype, sp))
        if de:
            dense.append((dtype, de))
    return sparse, dense
