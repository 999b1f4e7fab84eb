"""Autotuner: explores ZeRO stage x micro-batch size (x user-defined knobs) for the best throughput / latency.

For each ZeRO stage it first rules configs out with the model-state memory model (params / grads / optimizer state
per device), then measures the survivors — in process through a user ``run_fn(config) -> metric`` (a few profiling
steps), or by launching the training script with ``--deepspeed_config`` pointed at a generated file and reading the
metric file it writes. Reference: ``deepspeed/autotuning/autotuner.py:42-1100``, ``scheduler.py``, ``config.py``."""
from __future__ import annotations

import copy
import json
import os
import subprocess
import sys
from typing import Callable, Dict, List, Optional

from ..utils.logging import logger
from .staged import StagedSearch
from .tuner import GridSearchTuner, ModelBasedTuner, RandomTuner

DEFAULT_TUNING_SPACE = {
    0: {"zero_optimization": {"stage": 0}},
    1: {"zero_optimization": {"stage": 1, "reduce_bucket_size": [5e7, 5e8, 1e9], "allgather_bucket_size": [5e7, 5e8, 1e9]}},
    2: {"zero_optimization": {"stage": 2, "overlap_comm": [True, False], "reduce_scatter": [False, True],
                              "reduce_bucket_size": [5e7, 5e8, 1e9], "contiguous_gradients": [False, True]}},
    3: {"zero_optimization": {"stage": 3, "overlap_comm": [True, False], "reduce_bucket_size": [5e7, 5e8, 1e9],
                              "stage3_prefetch_bucket_size": [5e7, 5e8], "stage3_max_live_parameters": [1e9, 1e8]}},
}


def _expand(space: Dict) -> List[Dict]:
    """Cartesian product over every list-valued leaf."""
    out = [{}]
    for k, v in space.items():
        if isinstance(v, dict):
            subs = _expand(v)
            out = [{**o, k: s} for o in out for s in subs]
        elif isinstance(v, list):
            out = [{**o, k: x} for o in out for x in v]
        else:
            out = [{**o, k: v} for o in out]
    return out


def _merge(a: Dict, b: Dict) -> Dict:
    out = copy.deepcopy(a)
    for k, v in b.items():
        out[k] = _merge(out[k], v) if isinstance(v, dict) and isinstance(out.get(k), dict) else copy.deepcopy(v)
    return out


class Autotuner(StagedSearch):

    def __init__(self, user_config: Dict, num_params: int, num_gpus: int = 1, gpu_memory_bytes: Optional[int] = None,
                 run_fn: Optional[Callable[[Dict], Optional[float]]] = None, results_dir: str = "autotuning_results",
                 exps_dir: str = "autotuning_exps", user_script: Optional[str] = None, user_args: Optional[List[str]] = None):
        self.user_config = copy.deepcopy(user_config)
        at = dict(self.user_config.get("autotuning", {}))
        self._at = at
        self._metric = at.get("metric", "throughput")
        self.search = at.get("search", "staged")    # "staged": the reference's stage-by-stage walk; "flat": one tuner over all
        self.model_info = at.get("model_info") if isinstance(at.get("model_info"), dict) and "num_params" in at.get("model_info", {}) else None
        self.tuner_type = at.get("tuner_type", "gridsearch")
        self.early_stopping = at.get("tuner_early_stopping", 5)
        self.num_trials = at.get("tuner_num_trials", 50)
        self.fast = at.get("fast", True)
        self.max_mbs = at.get("max_train_micro_batch_size_per_gpu", 1024)
        self.min_mbs = at.get("min_train_micro_batch_size_per_gpu", 1)
        self.num_tuning_mbs = at.get("num_tuning_micro_batch_sizes", 3)
        self.mbs_list = at.get("micro_batch_sizes")
        self.num_params, self.num_gpus = num_params, num_gpus
        self.gpu_memory = gpu_memory_bytes if gpu_memory_bytes is not None else 180 * (1 << 30)  # B200
        self.run_fn = run_fn
        self.results_dir, self.exps_dir = at.get("results_dir", results_dir), at.get("exps_dir", exps_dir)
        self.user_script, self.user_args = user_script, user_args or []
        self.records: Dict[str, list] = {}
        self.best_config, self.best_metric = None, None

    # ---- memory model (reference: get_instantiation_memory_required_per_gpu :278) ------------------------------
    def model_state_bytes_per_gpu(self, stage: int, fp16: bool = True) -> int:
        n = self.num_params
        params = (2 if fp16 else 4) * n
        grads = (2 if fp16 else 4) * n
        optim = (16 if fp16 else 8) * n  # fp32 master + 2 Adam moments (+ fp32 grads in mixed precision)
        g = max(1, self.num_gpus)
        if stage >= 1:
            optim //= g
        if stage >= 2:
            grads //= g
        if stage >= 3:
            params //= g
        return params + grads + optim

    def feasible_stages(self) -> List[int]:
        fp16 = bool(self.user_config.get("fp16", {}).get("enabled") or self.user_config.get("bf16", {}).get("enabled"))
        want = self.user_config.get("zero_optimization", {}).get("stage")
        stages = [want] if isinstance(want, int) else [0, 1, 2, 3]
        return [s for s in stages if self.model_state_bytes_per_gpu(s, fp16) < self.gpu_memory]

    def _mbs_candidates(self) -> List[int]:
        if self.mbs_list:
            return list(self.mbs_list)
        out, m = [], max(1, self.min_mbs)
        while m <= self.max_mbs and len(out) < self.num_tuning_mbs:
            out.append(m)
            m *= 2
        return out

    def experiments(self) -> List[Dict]:
        exps = []
        for stage in self.feasible_stages():
            space = {"zero_optimization": {"stage": stage}} if self.fast else DEFAULT_TUNING_SPACE[stage]
            for zc in _expand(space):
                for mbs in self._mbs_candidates():
                    cfg = _merge(self.user_config, zc)
                    cfg.pop("autotuning", None)
                    cfg.pop("train_batch_size", None)
                    cfg["train_micro_batch_size_per_gpu"] = mbs
                    cfg.setdefault("gradient_accumulation_steps", 1)
                    exps.append(cfg)
        return exps

    # ---- running ------------------------------------------------------------------------------------------------
    def _run_script(self, cfg: Dict) -> Optional[float]:
        os.makedirs(self.exps_dir, exist_ok=True)
        os.makedirs(self.results_dir, exist_ok=True)
        idx = len(os.listdir(self.exps_dir))
        cfg_path = os.path.join(self.exps_dir, f"exp_{idx}.json")
        metric_path = os.path.join(self.results_dir, f"exp_{idx}_metrics.json")
        cfg = _merge(cfg, {"autotuning": {"enabled": True, "metric_path": metric_path, "start_profile_step": 3, "end_profile_step": 5}})
        with open(cfg_path, "w") as f:
            json.dump(cfg, f)
        cmd = [sys.executable, self.user_script, *self.user_args, "--deepspeed_config", cfg_path]
        try:
            subprocess.run(cmd, check=True, timeout=3600, stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
        except (subprocess.CalledProcessError, subprocess.TimeoutExpired):
            return None
        if not os.path.exists(metric_path):
            return None
        return json.load(open(metric_path)).get(self._metric)

    def _launch_with_config(self, cfg: Dict, file_name: str) -> int:
        os.makedirs(self.results_dir, exist_ok=True)
        path = os.path.join(self.results_dir, file_name)
        with open(path, "w") as f:
            json.dump(cfg, f, indent=2)
        cmd = [sys.executable, self.user_script, *self.user_args, "--deepspeed_config", path]
        return subprocess.run(cmd, stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL).returncode

    def tune(self):
        """Run the search (``autotuning.search``: "staged" - default - or "flat") and return the best config."""
        if self.search == "staged":
            if not self.feasible_stages():
                logger.warning("autotuning: the model states do not fit on the device under any ZeRO stage")
                return None
            best = self.tune_staged()
            logger.info(f"autotuning: best {self._metric} = {self.best_metric}")
            return best
        return self.tune_flat()

    def tune_flat(self):
        exps = self.experiments()
        if not exps:
            logger.warning("autotuning: the model states do not fit on the device under any ZeRO stage")
            return None
        run = self.run_fn if self.run_fn is not None else self._run_script
        cls = {"gridsearch": GridSearchTuner, "random": RandomTuner, "model_based": ModelBasedTuner}[self.tuner_type]
        tuner = cls(exps, run, self._metric)
        n = tuner.tune(sample_size=1, n_trials=self.num_trials, early_stopping=self.early_stopping)
        self.records["all"] = tuner.records
        self.best_config, self.best_metric = tuner.best_exp, tuner.best_metric_val
        logger.info(f"autotuning: {n} experiments, best {self._metric} = {self.best_metric}")
        return self.best_config

    def print_tuning_results(self):
        for exp, val in self.records.get("all", []):
            z = exp.get("zero_optimization", {}).get("stage", 0)
            print(f"z{z} mbs={exp['train_micro_batch_size_per_gpu']:<5d} {self._metric}={val}")

    def write_optimal_config(self, path: Optional[str] = None):
        if self.best_config is None:
            return None
        path = path or os.path.join(self.results_dir, "ds_config_optimal.json")
        os.makedirs(os.path.dirname(path) or ".", exist_ok=True)
        with open(path, "w") as f:
            json.dump(self.best_config, f, indent=2)
        return path
