"""``"autotuning"`` config section (``deepspeed/autotuning/config.py`` + ``constants.py``)."""
from __future__ import annotations

from typing import Dict, Optional

from ..config.config_utils import DeepSpeedConfigModel

METRIC_THROUGHPUT, METRIC_LATENCY, METRIC_FLOPS = "throughput", "latency", "flops"
TUNER_GRIDSEARCH, TUNER_RANDOM, TUNER_MODELBASED = "gridsearch", "random", "model_based"


class DeepSpeedAutotuningConfig(DeepSpeedConfigModel):
    enabled: bool = False
    fast: bool = True
    results_dir: str = "autotuning_results"
    exps_dir: str = "autotuning_exps"
    overwrite: bool = True
    start_profile_step: int = 3
    end_profile_step: int = 5
    metric: str = METRIC_THROUGHPUT
    metric_path: Optional[str] = None
    tuner_type: str = TUNER_GRIDSEARCH
    tuner_early_stopping: int = 5
    tuner_num_trials: int = 50
    arg_mappings: Optional[Dict[str, str]] = None
    max_train_batch_size: Optional[int] = None
    min_train_batch_size: int = 1
    max_train_micro_batch_size_per_gpu: int = 1024
    min_train_micro_batch_size_per_gpu: int = 1
    num_tuning_micro_batch_sizes: int = 3
    mp_size: int = 1
    model_info: Optional[Dict] = None
    model_info_path: Optional[str] = None


def get_autotuning_config(param_dict: dict) -> DeepSpeedAutotuningConfig:
    return DeepSpeedAutotuningConfig(**(param_dict.get("autotuning") or {}))


def get_default_model_info_config():
    return {"profile": False, "num_params": None, "hidden_size": None, "num_layers": None}


def get_model_info_config(param_dict):
    """The ``autotuning.model_info`` block over its defaults (``None`` without the block)."""
    at = param_dict.get("autotuning") or {}
    if not at.get("model_info"):
        return None
    return {**get_default_model_info_config(), **at["model_info"]}
