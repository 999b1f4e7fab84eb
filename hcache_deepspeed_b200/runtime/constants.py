"""Config key names and defaults as module constants, for code written against ``deepspeed/runtime/constants.py``
(``from ...constants import TRAIN_BATCH_SIZE``). The framework itself reads its pydantic sections (:mod:`.config`); this
table is generated from ``_KEYS`` below: each entry ``key: default`` exports ``KEY = "key"`` and ``KEY_DEFAULT``."""
from __future__ import annotations

_KEYS = {
    # batch geometry
    "train_batch_size": None, "train_micro_batch_size_per_gpu": None, "gradient_accumulation_steps": None,
    # sections
    "optimizer": None, "scheduler": None, "fp16": None, "bf16": None, "amp": None, "zero_optimization": None,
    "activation_checkpointing": None, "pipeline": None, "checkpoint": None, "data_types": None, "tensor_parallel": None,
    "eigenvalue": None, "progressive_layer_drop": None, "curriculum_learning": None, "data_efficiency": None,
    "elasticity": None, "autotuning": None, "compression_training": None, "aio": None, "hybrid_engine": None,
    "compile": None, "timers": None, "comms_logger": None, "flops_profiler": None,
    # scalars
    "steps_per_print": 10, "dump_state": False, "disable_allgather": False, "communication_data_type": None,
    "seq_parallel_communication_data_type": "fp32", "prescale_gradients": False, "gradient_predivide_factor": 1.0,
    "sparse_gradients": False, "gradient_clipping": 0.0, "graph_harvesting": False, "wall_clock_breakdown": False,
    "memory_breakdown": False, "vocabulary_size": None, "dataloader_drop_last": False,
    "zero_allow_untested_optimizer": False, "zero_force_ds_cpu_optimizer": True, "sequence_parallel_size": 1,
    "use_data_before_expert_parallelism": False,
}

for _k, _d in _KEYS.items():
    globals()[_k.upper()] = _k
    globals()[_k.upper() + "_DEFAULT"] = _d

# names that differ from their key
BFLOAT16, BFLOAT16_OLD = "bf16", "bfloat16"
TYPE, OPTIMIZER_PARAMS, SCHEDULER_PARAMS, LEGACY_FUSION = "type", "params", "params", "legacy_fusion"
MAX_GRAD_NORM = "max_grad_norm"
OPTIMIZER_TYPE_DEFAULT = SCHEDULER_TYPE_DEFAULT = None
FP16_ENABLED, FP16_LOSS_SCALE, FP16_INITIAL_SCALE_POWER = "enabled", "loss_scale", "initial_scale_power"
FP16_LOSS_SCALE_WINDOW, FP16_HYSTERESIS, FP16_MIN_LOSS_SCALE = "loss_scale_window", "hysteresis", "min_loss_scale"
FP16_CONSECUTIVE_HYSTERESIS, FP16_MASTER_WEIGHTS_AND_GRADS = "consecutive_hysteresis", "fp16_master_weights_and_grads"
BFLOAT16_ENABLED, BFLOAT16_IMMEDIATE_GRAD_UPDATE = "enabled", "immediate_grad_update"
AMP_ENABLED = "enabled"
GRAD_ACCUM_DTYPE = "grad_accum_dtype"
CHECKPOINT_TAG_VALIDATION, LOAD_UNIVERSAL_CHECKPOINT, USE_NODE_LOCAL_STORAGE_CHECKPOINT = (
    "tag_validation", "load_universal", "use_node_local_storage")
CHECKPOINT_PARALLEL_WRITE, CHECKPOINT_PARALLEL_WRITE_PIPELINE_STAGE = "parallel_write", "pipeline_stage"
CHECKPOINT_TAG_VALIDATION_MODES = ["WARN", "IGNORE", "FAIL"]
PLD_ENABLED, PLD_THETA, PLD_GAMMA = "enabled", "theta", "gamma"
EIGENVALUE_ENABLED = "enabled"
PIPE_REPLICATED = "ds_pipe_replicated"
DATA_PARALLEL_GROUP = "data_parallel_group"
GLOBAL_RANK = "global_rank"
ROUTE_TRAIN, ROUTE_EVAL, ROUTE_PREDICT, ROUTE_ENCODE = "train", "eval", "predict", "encode"
TRAIN_BATCH_SIZE_EXAMPLE = {"train_batch_size": 32}


def defaults() -> dict:
    """``key -> default`` of every scalar top-level key."""
    return {k: d for k, d in _KEYS.items() if d is not None}

from .config_getters import ValidationMode  # noqa: E402,F401
