"""Training configuration: JSON/dict -> :class:`DeepSpeedConfig`.

Same user-facing keys as the reference (``deepspeed/runtime/config.py``, ``constants.py``, ``zero/config.py``,
``zero/offload_config.py``, ``activation_checkpointing/config.py``, ``pipe`` section of ``config.py:706+``), but
modelled as pydantic sections on :class:`DeepSpeedConfigModel` instead of a table of string constants.
Batch-size resolution follows ``config.py:_configure_train_batch_size``:
``train_batch_size = micro_batch_per_gpu * gradient_accumulation_steps * data_parallel_world_size``;
any one (or two) of the three may be omitted.
"""
from __future__ import annotations

import json
import os
from enum import Enum
from typing import Any, Dict, List, Optional, Union

from pydantic import Field, model_validator

from ..config.config_utils import DeepSpeedConfigModel
from ..monitor.config import get_monitor_config


class DeepSpeedConfigError(Exception):
    pass


# ---------------------------------------------------------------------------------------------------
# sections
# ---------------------------------------------------------------------------------------------------
class FP16Config(DeepSpeedConfigModel):
    enabled: bool = False
    auto_cast: bool = False
    loss_scale: float = 0.0  # 0 = dynamic
    initial_scale_power: int = 16
    loss_scale_window: int = 1000
    hysteresis: int = 2
    consecutive_hysteresis: bool = False
    min_loss_scale: float = 1.0
    fp16_master_weights_and_grads: bool = False

    @property
    def dynamic(self) -> bool:
        return self.loss_scale == 0


class BF16Config(DeepSpeedConfigModel):
    enabled: bool = False
    immediate_grad_update: bool = False


class AMPConfig(DeepSpeedConfigModel):
    enabled: bool = False
    opt_level: str = "O1"


class OffloadDeviceEnum(str, Enum):
    none = "none"
    cpu = "cpu"
    nvme = "nvme"


class OffloadParamConfig(DeepSpeedConfigModel):
    device: OffloadDeviceEnum = OffloadDeviceEnum.none
    nvme_path: Optional[str] = None
    buffer_count: int = Field(5, ge=0)
    buffer_size: int = Field(int(1e8), ge=0)
    max_in_cpu: int = Field(int(1e9), ge=0)
    pin_memory: bool = False


class OffloadOptimizerConfig(DeepSpeedConfigModel):
    device: OffloadDeviceEnum = OffloadDeviceEnum.none
    nvme_path: Optional[str] = None
    buffer_count: int = Field(4, ge=0)
    pin_memory: bool = False
    pipeline_read: bool = False
    pipeline_write: bool = False
    fast_init: bool = False
    ratio: float = Field(1.0, ge=0.0, le=1.0)


class ZeroConfig(DeepSpeedConfigModel):
    """``zero_optimization`` section (reference ``runtime/zero/config.py:83-330``)."""
    stage: int = Field(0, ge=0, le=3)
    contiguous_gradients: bool = True
    reduce_scatter: bool = True
    reduce_bucket_size: int = Field(int(5e8), ge=0)
    use_multi_rank_bucket_allreduce: bool = True
    allgather_partitions: bool = True
    allgather_bucket_size: int = Field(int(5e8), ge=0)
    overlap_comm: Optional[bool] = None
    load_from_fp32_weights: bool = True
    elastic_checkpoint: bool = False
    offload_param: Optional[OffloadParamConfig] = None
    offload_optimizer: Optional[OffloadOptimizerConfig] = None
    sub_group_size: int = Field(int(1e9), ge=0)
    cpu_offload: Optional[bool] = Field(None, json_schema_extra={
        "deprecated": True, "new_param": "offload_optimizer",
        "new_param_fn": (lambda val: OffloadOptimizerConfig(device="cpu") if val else None)})
    prefetch_bucket_size: int = Field(int(5e7), ge=0, alias="stage3_prefetch_bucket_size")
    param_persistence_threshold: int = Field(int(1e5), ge=0, alias="stage3_param_persistence_threshold")
    model_persistence_threshold: int = Field(int(9e18), ge=0, alias="stage3_model_persistence_threshold")
    max_live_parameters: int = Field(int(1e9), ge=0, alias="stage3_max_live_parameters")
    max_reuse_distance: int = Field(int(1e9), ge=0, alias="stage3_max_reuse_distance")
    gather_16bit_weights_on_model_save: bool = Field(False, alias="stage3_gather_16bit_weights_on_model_save")
    module_granularity_threshold: int = Field(0, alias="stage3_module_granularity_threshold")
    use_all_reduce_for_fetch_params: bool = Field(False, alias="stage3_use_all_reduce_for_fetch_params")
    ignore_unused_parameters: bool = True
    legacy_stage1: bool = False
    round_robin_gradients: bool = False
    zero_hpz_partition_size: int = Field(1, ge=0)
    zero_quantized_weights: bool = False
    zero_quantized_nontrainable_weights: bool = False
    zero_quantized_gradients: bool = False
    zeropp_loco_param: Optional[Dict[str, Any]] = None
    mics_shard_size: int = -1
    mics_hierarchical_params_gather: bool = False
    memory_efficient_linear: bool = True
    pipeline_loading_checkpoint: bool = False
    override_module_apply: bool = True
    log_trace_cache_warnings: bool = False

    @model_validator(mode="after")
    def _resolve(self):
        if self.cpu_offload and self.offload_optimizer is None:
            object.__setattr__(self, "offload_optimizer", OffloadOptimizerConfig(device="cpu"))
        if self.overlap_comm is None:
            object.__setattr__(self, "overlap_comm", self.stage == 3)
        return self


class ActivationCheckpointingConfig(DeepSpeedConfigModel):
    partition_activations: bool = False
    contiguous_memory_optimization: bool = False
    cpu_checkpointing: bool = False
    number_checkpoints: Optional[int] = None
    synchronize_checkpoint_boundary: bool = False
    profile: bool = False


class PipelineConfig(DeepSpeedConfigModel):
    stages: Union[str, int] = "auto"
    partition: str = "best"
    seed_layers: bool = False
    activation_checkpoint_interval: int = 0
    pipe_partitioned: bool = True
    grad_partitioned: bool = True
    use_reentrant: bool = True


class CommsLoggerConfig(DeepSpeedConfigModel):
    enabled: bool = False
    prof_all: bool = True
    prof_ops: List[str] = []
    verbose: bool = False
    debug: bool = False


class FlopsProfilerConfig(DeepSpeedConfigModel):
    enabled: bool = False
    recompute_fwd_factor: float = 0.0
    profile_step: int = 1
    module_depth: int = -1
    top_modules: int = 1
    detailed: bool = True
    output_file: Optional[str] = None


class CheckpointConfig(DeepSpeedConfigModel):
    tag_validation: str = "Warn"
    load_universal: bool = False
    use_node_local_storage: bool = False
    parallel_write: Dict[str, Any] = {}

    @model_validator(mode="after")
    def _check(self):
        if self.tag_validation.upper() not in ("WARN", "IGNORE", "FAIL"):
            raise ValueError(f"checkpoint.tag_validation must be Warn/Ignore/Fail, got {self.tag_validation}")
        return self


class DataTypesConfig(DeepSpeedConfigModel):
    grad_accum_dtype: Optional[str] = None


class TensorParallelTrainConfig(DeepSpeedConfigModel):
    autotp_size: int = 0
    tp_grain_size: int = 1


class EigenvalueConfig(DeepSpeedConfigModel):
    enabled: bool = False
    verbose: bool = False
    max_iter: int = 100
    tol: float = 1e-2
    stability: float = 1e-6
    gas_boundary_resolution: int = 1
    layer_name: str = "bert.encoder.layer"
    layer_num: int = 0


class PLDConfig(DeepSpeedConfigModel):
    enabled: bool = False
    theta: float = 1.0
    gamma: float = 0.001


class CurriculumLegacyConfig(DeepSpeedConfigModel):
    enabled: bool = False
    curriculum_type: str = "seqlen"
    min_difficulty: int = 8
    max_difficulty: int = 1024
    schedule_type: str = "fixed_linear"
    schedule_config: Dict[str, Any] = {}


ADAM_OPTIMIZER, ADAMW_OPTIMIZER, LAMB_OPTIMIZER = "adam", "adamw", "lamb"
ONEBIT_ADAM_OPTIMIZER, ZERO_ONE_ADAM_OPTIMIZER, ONEBIT_LAMB_OPTIMIZER = "onebitadam", "zerooneadam", "onebitlamb"
ADAGRAD_OPTIMIZER, LION_OPTIMIZER, MUADAM_OPTIMIZER, SGD_OPTIMIZER = "adagrad", "lion", "muadam", "sgd"
DEEPSPEED_OPTIMIZERS = [ADAM_OPTIMIZER, ADAMW_OPTIMIZER, LAMB_OPTIMIZER, ONEBIT_ADAM_OPTIMIZER, ZERO_ONE_ADAM_OPTIMIZER,
                        ONEBIT_LAMB_OPTIMIZER, ADAGRAD_OPTIMIZER, LION_OPTIMIZER, SGD_OPTIMIZER]


class DeepSpeedConfig:
    """Parsed training configuration. ``config`` is a dict or a path to a JSON file; ``world_size`` is the
    data-parallel world size used to resolve the batch-size triple (``mpu`` may supply it instead)."""

    def __init__(self, config: Union[str, dict], mpu=None, mesh_device=None, world_size: Optional[int] = None):
        if isinstance(config, dict):
            self._param_dict = dict(config)
        elif isinstance(config, (str, os.PathLike)) and os.path.exists(config):
            with open(config, "r") as f:
                self._param_dict = json.load(f)
        else:
            raise DeepSpeedConfigError(f"Expected a dict or a path to an existing JSON config file, got: {config!r}")
        if world_size is None:
            try:
                from .. import comm as dist
                if mpu is not None:
                    world_size = mpu.get_data_parallel_world_size()
                elif dist.is_initialized():
                    world_size = dist.get_world_size()
                else:
                    world_size = int(os.environ.get("WORLD_SIZE", 1))
            except Exception:
                world_size = 1
            pd_ = self._param_dict
            sp = int(pd_.get("sequence_parallel_size", 1) or 1)
            tp = int((pd_.get("tensor_parallel") or {}).get("autotp_size", 0) or 0)
            if mpu is None:
                world_size = max(1, world_size // max(1, sp) // max(1, tp if tp > 0 else 1))
        self.world_size = world_size
        self._apply_elasticity()
        self._initialize_params(self._param_dict)
        self._configure_train_batch_size()
        self._do_sanity_check()

    def _apply_elasticity(self):
        """``"elasticity": {"enabled": true, ...}``: the batch-size triple is DERIVED from the elastic plan for the current
        world size (global batch = the plan's, micro batch = the largest admissible one, accumulation = what is left);
        batch sizes given by the user are an error unless ``ignore_non_elastic_batch_info`` (ref: runtime/config.py:739-790)."""
        pd = self._param_dict
        el = pd.get("elasticity") or {}
        if not el.get("enabled", False):
            return
        from ..elasticity import compute_elastic_config
        from ..elasticity.constants import IGNORE_NON_ELASTIC_BATCH_INFO
        ws = max(1, int(self.world_size or 1))
        final_bs, valid_gpus, micro = compute_elastic_config(ds_config=pd, world_size=ws, return_microbatch=True)
        if not el.get(IGNORE_NON_ELASTIC_BATCH_INFO, False):
            given = [k for k in ("train_batch_size", "train_micro_batch_size_per_gpu", "gradient_accumulation_steps") if k in pd]
            if given:
                raise DeepSpeedConfigError(f"elasticity is enabled: {given} must not be set in the config (they follow from the "
                                           f"elastic plan); remove them or set elasticity.{IGNORE_NON_ELASTIC_BATCH_INFO}")
        if micro is None:
            mbs = sorted(el.get("micro_batch_sizes", [1]), reverse=True)
            micro = next((m for m in mbs if final_bs % (m * ws) == 0), None)
            if micro is None:
                raise DeepSpeedConfigError(f"no elastic micro batch size of {mbs} divides batch {final_bs} over {ws} ranks")
        gas = final_bs // (micro * ws)
        if gas < 1 or gas * micro * ws != final_bs:
            raise DeepSpeedConfigError(f"elastic batch {final_bs} cannot be split into micro batches of {micro} over {ws} ranks")
        pd["train_batch_size"], pd["train_micro_batch_size_per_gpu"], pd["gradient_accumulation_steps"] = final_bs, micro, gas
        self.elastic_plan = {"train_batch_size": final_bs, "valid_gpus": valid_gpus, "micro_batch_size": micro,
                             "gradient_accumulation_steps": gas}

    # -----------------------------------------------------------------------------------------------
    def _initialize_params(self, pd: dict):
        g = pd.get
        self.train_batch_size = g("train_batch_size")
        self.train_micro_batch_size_per_gpu = g("train_micro_batch_size_per_gpu")
        self.gradient_accumulation_steps = g("gradient_accumulation_steps")
        self.steps_per_print = g("steps_per_print", 10)
        self.dump_state = g("dump_state", False)
        self.disable_allgather = g("disable_allgather", False)
        self.communication_data_type = g("communication_data_type")
        self.seq_parallel_communication_data_type = g("seq_parallel_communication_data_type", "fp32")
        self.prescale_gradients = g("prescale_gradients", False)
        self.gradient_predivide_factor = g("gradient_predivide_factor", 1.0)
        self.sparse_gradients_enabled = g("sparse_gradients", False)
        self.gradient_clipping = float(g("gradient_clipping", 0.0))
        self.graph_harvesting = g("graph_harvesting", False)
        self.wall_clock_breakdown = g("wall_clock_breakdown", False)
        self.memory_breakdown = g("memory_breakdown", False)
        self.vocabulary_size = g("vocabulary_size")
        self.dataloader_drop_last = g("dataloader_drop_last", False)
        self.zero_allow_untested_optimizer = g("zero_allow_untested_optimizer", False)
        self.zero_force_ds_cpu_optimizer = g("zero_force_ds_cpu_optimizer", True)
        self.sequence_parallel_size = int(g("sequence_parallel_size", 1) or 1)
        self.use_data_before_expert_parallel_ = g("use_data_before_expert_parallelism", False)

        self.fp16_config = FP16Config(**(g("fp16") or {}))
        self.bf16_config = BF16Config(**(g("bf16") or g("bfloat16") or {}))
        self.amp_config = AMPConfig(**{k: v for k, v in (g("amp") or {}).items() if k in ("enabled", "opt_level")})
        self.amp_params = {k: v for k, v in (g("amp") or {}).items() if k != "enabled"}
        if self.fp16_config.enabled and self.bf16_config.enabled:
            raise DeepSpeedConfigError("bf16 and fp16 modes cannot be simultaneously enabled")
        self.fp16_enabled, self.bfloat16_enabled, self.amp_enabled = (self.fp16_config.enabled, self.bf16_config.enabled,
                                                                     self.amp_config.enabled)
        self.loss_scale = self.fp16_config.loss_scale
        self.initial_dynamic_scale = 2 ** self.fp16_config.initial_scale_power
        self.dynamic_loss_scale_args = dict(init_scale=2 ** self.fp16_config.initial_scale_power,
                                            scale_window=self.fp16_config.loss_scale_window,
                                            delayed_shift=self.fp16_config.hysteresis,
                                            consecutive_hysteresis=self.fp16_config.consecutive_hysteresis,
                                            min_scale=self.fp16_config.min_loss_scale)

        self.zero_config = ZeroConfig(**(g("zero_optimization") or {}))
        self.zero_optimization_stage = self.zero_config.stage
        self.zero_enabled = self.zero_optimization_stage > 0

        opt = g("optimizer") or {}
        self.optimizer_name = opt.get("type").lower() if isinstance(opt.get("type"), str) and \
            opt.get("type").lower() in DEEPSPEED_OPTIMIZERS else opt.get("type")
        self.optimizer_params = dict(opt.get("params") or {})
        self.optimizer_legacy_fusion = opt.get("legacy_fusion", False)
        sch = g("scheduler") or {}
        self.scheduler_name = sch.get("type")
        self.scheduler_params = dict(sch.get("params") or {})

        self.activation_checkpointing_config = ActivationCheckpointingConfig(**(g("activation_checkpointing") or {}))
        self.pipeline = PipelineConfig(**(g("pipeline") or {})).model_dump()
        self.comms_config = CommsLoggerConfig(**(g("comms_logger") or {}))
        self.flops_profiler_config = FlopsProfilerConfig(**(g("flops_profiler") or {}))
        self.monitor_config = get_monitor_config(pd)
        self.checkpoint_config = CheckpointConfig(**(g("checkpoint") or {}))
        self.checkpoint_tag_validation_enabled = self.checkpoint_config.tag_validation.upper() != "IGNORE"
        self.checkpoint_tag_validation_fail = self.checkpoint_config.tag_validation.upper() == "FAIL"
        self.load_universal_checkpoint = self.checkpoint_config.load_universal
        self.use_node_local_storage = self.checkpoint_config.use_node_local_storage
        dt = DataTypesConfig(**(g("data_types") or {}))
        self.grad_accum_dtype = dt.grad_accum_dtype
        self.tensor_parallel_config = TensorParallelTrainConfig(**{k: v for k, v in (g("tensor_parallel") or {}).items()
                                                                   if k in ("autotp_size", "tp_grain_size")})
        self.eigenvalue_config = EigenvalueConfig(**(g("eigenvalue") or {}))
        self.eigenvalue_enabled = self.eigenvalue_config.enabled
        self.pld_config = PLDConfig(**(g("progressive_layer_drop") or {}))
        self.pld_enabled = self.pld_config.enabled
        self.pld_params = self.pld_config.model_dump() if self.pld_enabled else False
        self.curriculum_config_legacy = CurriculumLegacyConfig(**(g("curriculum_learning") or {}))
        self.curriculum_enabled_legacy = self.curriculum_config_legacy.enabled
        self.curriculum_params_legacy = dict(g("curriculum_learning") or {})
        self.data_efficiency_config = dict(g("data_efficiency") or {})
        self.data_efficiency_enabled = bool(self.data_efficiency_config.get("enabled", False))
        self.elasticity_enabled = bool((g("elasticity") or {}).get("enabled", False))
        self.elasticity_config = dict(g("elasticity") or {})
        self.autotuning_config = dict(g("autotuning") or {})
        self.compression_config = dict(g("compression_training") or {})
        self.aio_config = {"block_size": 1048576, "queue_depth": 8, "intra_op_parallelism": 1, "single_submit": False,
                           "overlap_events": True, "use_gds": False, **(g("aio") or {})}
        self.hybrid_engine = dict(g("hybrid_engine") or {"enabled": False})
        self.compile_config = dict(g("compile") or {})
        self.timers_config = dict(g("timers") or {})

    # -----------------------------------------------------------------------------------------------
    def _configure_train_batch_size(self):
        tb, mb, gas, ws = (self.train_batch_size, self.train_micro_batch_size_per_gpu,
                           self.gradient_accumulation_steps, self.world_size)
        if tb is not None and mb is not None and gas is not None:
            pass
        elif tb is not None and mb is not None:
            gas = tb // (mb * ws)
        elif tb is not None and gas is not None:
            mb = tb // (ws * gas)
        elif mb is not None and gas is not None:
            tb = mb * gas * ws
        elif tb is not None:
            gas = 1
            mb = tb // ws
        elif mb is not None:
            gas = 1
            tb = mb * ws
        else:
            raise DeepSpeedConfigError("Either train_batch_size or train_micro_batch_size_per_gpu needs to be provided")
        self.train_batch_size, self.train_micro_batch_size_per_gpu, self.gradient_accumulation_steps = tb, mb, gas
        for name, v in (("train_batch_size", tb), ("train_micro_batch_size_per_gpu", mb),
                        ("gradient_accumulation_steps", gas)):
            if not isinstance(v, int) or v <= 0:
                raise DeepSpeedConfigError(f"{name}: {v} has to be a positive integer (world size {ws})")
        if tb != mb * gas * ws:
            raise DeepSpeedConfigError(
                f"Check batch related parameters. train_batch_size is not equal to micro_batch_per_gpu * gradient_acc_step"
                f" * world_size {tb} != {mb} * {gas} * {ws}")

    def _do_sanity_check(self):
        if self.zero_enabled and self.zero_optimization_stage > 3:
            raise DeepSpeedConfigError("ZeRO stage must be in 0..3")
        if self.fp16_config.fp16_master_weights_and_grads and self.zero_optimization_stage != 2:
            raise DeepSpeedConfigError("fp16_master_weights_and_grads is only supported with ZeRO stage 2")
        if self.optimizer_name is not None and self.optimizer_params.get("max_grad_norm", 0) and self.fp16_enabled:
            self.optimizer_params["max_grad_norm"] = 0.0  # fp16 wrapper does the clipping (reference warns and resets)

    def print(self, name: str = "DeepSpeedConfig"):
        from ..utils.logging import logger
        logger.info(f"{name}:")
        for k in sorted(vars(self)):
            if k != "_param_dict":
                logger.info(f"  {k} {'.' * max(1, 30 - len(k))} {getattr(self, k)}")

    def print_user_config(self):
        from ..utils.logging import logger
        logger.info("  json = " + json.dumps(self._param_dict, sort_keys=True, indent=4, default=str))


# single-entry readers of a raw config dict under the reference's names (used by integrations that hold only the JSON)
from .config_getters import *  # noqa: E402,F401,F403
from .config_getters import DeepSpeedConfigWriter, DtypeEnum, HybridEngineConfig, ValidationMode, get_scalar_param  # noqa: E402,F401
