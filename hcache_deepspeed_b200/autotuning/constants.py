"""Names used by the autotuner (``deepspeed/autotuning/constants.py``)."""
from .config import (METRIC_FLOPS, METRIC_LATENCY, METRIC_THROUGHPUT, TUNER_GRIDSEARCH, TUNER_MODELBASED,  # noqa: F401
                     TUNER_RANDOM)

AUTOTUNING = "autotuning"
AUTOTUNING_ENABLED, AUTOTUNING_FAST, AUTOTUNING_METRIC, AUTOTUNING_METRIC_PATH = "enabled", "fast", "metric", "metric_path"
AUTOTUNING_RESULTS_DIR, AUTOTUNING_EXPS_DIR, AUTOTUNING_OVERWRITE = "results_dir", "exps_dir", "overwrite"
AUTOTUNING_START_PROFILE_STEP, AUTOTUNING_END_PROFILE_STEP = "start_profile_step", "end_profile_step"
AUTOTUNING_TUNER_TYPE, AUTOTUNING_TUNER_EARLY_STOPPING, AUTOTUNING_TUNER_NUM_TRIALS = "tuner_type", "tuner_early_stopping", "tuner_num_trials"
AUTOTUNING_ARG_MAPPINGS, AUTOTUNING_MODEL_INFO, AUTOTUNING_MODEL_INFO_PATH = "arg_mappings", "model_info", "model_info_path"
AUTOTUNING_MAX_TRAIN_BATCH_SIZE, AUTOTUNING_MIN_TRAIN_BATCH_SIZE = "max_train_batch_size", "min_train_batch_size"
AUTOTUNING_MAX_TRAIN_MICRO_BATCH_SIZE_PER_GPU, AUTOTUNING_MIN_TRAIN_MICRO_BATCH_SIZE_PER_GPU = (
    "max_train_micro_batch_size_per_gpu", "min_train_micro_batch_size_per_gpu")
AUTOTUNING_NUM_TUNING_MICRO_BATCH_SIZES, AUTOTUNING_MP_SIZE = "num_tuning_micro_batch_sizes", "mp_size"
MODEL_INFO_PROFILE, MODEL_INFO_NUM_PARAMS, MODEL_INFO_HIDDEN_SIZE, MODEL_INFO_NUM_LAYERS = "profile", "num_params", "hidden_size", "num_layers"
DEFAULT_MIN_MEM_CONFIG = {"train_micro_batch_size_per_gpu": 1, "zero_optimization": {"stage": 3}, "memory_break_down": False}
