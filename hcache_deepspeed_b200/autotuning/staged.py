"""The stage-by-stage search of the autotuner and the accessors around it, mixed into :class:`Autotuner`.

Order of work (the reference's algorithm, ``deepspeed/autotuning/autotuner.py:404-1100``):

1. ZeRO stages are visited in increasing order; a stage whose model states cannot fit (memory model) is skipped.
2. Within a stage the micro-batch size is tuned first: find the smallest and the largest size that run (an experiment
   that fails or reports nothing counts as out of memory; the upper end is found by doubling, then bisecting), measure
   a short list of sizes between them and stop at the plateau - the first size that no longer improves the metric.
3. With the best micro-batch size fixed, the stage's other knobs (bucket sizes, overlap, ...) are searched with the
   configured tuner (grid / random / model based), unless ``fast`` is on.
4. In ``fast`` mode the walk over stages ends as soon as a stage does not beat the best one before it.

Every measurement lands in ``records[space name]``; ``get_best_space_record`` / ``get_best_space_records`` read them.
"""
from __future__ import annotations

import copy
import re
from typing import Dict, List, Optional, Tuple

from ..utils.logging import logger

GLOBAL_SPACE = "global"
METRIC_PLATEAU_RATIO = 0.05      # "no longer improves": less than this relative gain over the best so far


class StagedSearch:

    # ---- accessors of the "autotuning" block (ref: autotuner.py:126-170) -------------------------------------------------
    def metric(self) -> str:
        return self._metric

    def fast_enabled(self) -> bool:
        return bool(self.fast)

    def max_train_batch_size(self) -> Optional[int]:
        return self._at.get("max_train_batch_size")

    def mp_size(self) -> int:
        return int(self._at.get("mp_size", 1))

    def max_train_micro_batch_size_per_gpu(self) -> int:
        cap = self.max_train_batch_size()
        if cap:                              # the global cap bounds the micro batch: batch = mbs * gas * dp
            dp = max(1, self.num_gpus // self.mp_size())
            return max(1, min(self.max_mbs, cap // (self.get_gas_from_user_config() * dp)))
        return self.max_mbs

    def min_train_micro_batch_size_per_gpu(self) -> int:
        return self.min_mbs

    def num_tuning_micro_batch_sizes(self) -> int:
        return self.num_tuning_mbs

    def fp16_enabled(self) -> bool:
        return bool((self.user_config.get("fp16") or {}).get("enabled") or (self.user_config.get("bf16") or {}).get("enabled"))

    def get_gpu_memory_info(self) -> int:
        return self.gpu_memory

    def get_model_num_params(self) -> int:
        return self.num_params

    def get_activation_memory_per_gpu(self) -> int:
        """Bytes of activations per micro-batch sample as reported by the model-info run (0 when unknown)."""
        return int((self.model_info or {}).get("activation_mem_per_gpu", 0))

    def get_instantiation_memory_required_per_gpu(self, zero_stage: int) -> int:
        """Model states per device under ``zero_stage`` (tensor parallelism divides everything first)."""
        return self.model_state_bytes_per_gpu(zero_stage, self.fp16_enabled()) // self.mp_size()

    def get_gas_from_user_config(self) -> int:
        gas = self.user_config.get("gradient_accumulation_steps", 1)
        if isinstance(gas, str):
            gas = self.get_val_from_user_args(gas) if gas.startswith("$") or gas == "auto" else gas
        return max(1, int(gas or 1)) if not isinstance(gas, str) or gas.isdigit() else 1

    def get_val_from_user_args(self, ds_name: str):
        """Config values may point at a script argument (``"$--batch"`` style, or ``"auto"`` with an ``arg_mappings``
        entry): return that argument's value from the user command line, or ``None``."""
        name = ds_name[1:] if ds_name.startswith("$") else (self._at.get("arg_mappings") or {}).get(ds_name, ds_name)
        args = list(self.user_args)
        for i, a in enumerate(args):
            if a == name and i + 1 < len(args):
                return _number(args[i + 1])
            if a.startswith(name + "="):
                return _number(a.split("=", 1)[1])
        return None

    # ---- records -------------------------------------------------------------------------------------------------------
    def update_records(self, space_name: str, exp: Dict, metric_val: Optional[float], num_exps: int = 1) -> None:
        self.records.setdefault(space_name, []).append((exp, metric_val, num_exps))

    def get_best_space_record(self, space_name: str) -> Optional[Tuple[Dict, float, int]]:
        """(best experiment, its metric, number of experiments run) of one space; ``None`` if nothing ran there."""
        rows = [r for r in self.records.get(space_name, []) if r[1] is not None]
        if not rows:
            return None
        best = (min if self._metric == "latency" else max)(rows, key=lambda r: r[1])
        return best[0], best[1], sum(r[2] for r in self.records[space_name])

    def get_best_space_records(self) -> Dict[str, Tuple[Dict, float, int]]:
        out = {name: self.get_best_space_record(name) for name in self.records if name != "all"}
        out = {k: v for k, v in out.items() if v is not None}
        if out:
            pick = (min if self._metric == "latency" else max)(out.values(), key=lambda r: r[1])
            out[GLOBAL_SPACE] = (pick[0], pick[1], sum(v[2] for v in out.values()))
        return out

    def _better(self, a: Optional[float], b: Optional[float]) -> bool:
        if a is None:
            return False
        if b is None:
            return True
        return a < b if self._metric == "latency" else a > b

    # ---- running one config ------------------------------------------------------------------------------------------------
    def run_ds_config(self, ds_config: Dict, exp_name: str) -> Optional[float]:
        """Measure one config (in process through ``run_fn``, else by launching the script); ``None`` = did not run."""
        cfg = copy.deepcopy(ds_config)
        cfg.pop("autotuning", None)
        run = self.run_fn if self.run_fn is not None else self._run_script
        val = run(cfg)
        logger.info(f"autotuning: {exp_name}: {self._metric} = {val}")
        return val

    def _stage_config(self, stage: int, mbs: int, knobs: Optional[Dict] = None) -> Dict:
        from .autotuner import _merge
        cfg = _merge(self.user_config, knobs or {"zero_optimization": {"stage": stage}})
        cfg.setdefault("zero_optimization", {})["stage"] = stage
        cfg.pop("train_batch_size", None)
        cfg["train_micro_batch_size_per_gpu"] = int(mbs)
        cfg["gradient_accumulation_steps"] = self.get_gas_from_user_config()
        return cfg

    # ---- micro-batch size ----------------------------------------------------------------------------------------------------
    def get_min_max_micro_batch_size(self, stage: int, min_mbs: int, calculated_max: int) -> Tuple[int, int]:
        """Smallest and largest micro-batch size that run under ``stage`` within ``[min_mbs, calculated_max]``; (-1, -1)
        when not even the smallest runs. The upper end is searched by doubling and then bisecting the failing interval."""
        name = f"z{stage}"
        tried: Dict[int, Optional[float]] = {}

        def runs(m: int) -> bool:
            if m not in tried:
                tried[m] = self.run_ds_config(self._stage_config(stage, m), f"{name}_mbs{m}")
                self.update_records(name, self._stage_config(stage, m), tried[m])
            return tried[m] is not None

        if not runs(min_mbs):
            return -1, -1
        lo, hi = min_mbs, None
        m = min_mbs
        while m * 2 <= calculated_max:
            m *= 2
            if runs(m):
                lo = m
            else:
                hi = m
                break
        if hi is None and lo < calculated_max and runs(calculated_max):
            lo = calculated_max
        elif hi is None and lo < calculated_max:
            hi = calculated_max
        while hi is not None and hi - lo > max(1, lo // 8):       # bisect to within ~12 %
            mid = (lo + hi) // 2
            if runs(mid):
                lo = mid
            else:
                hi = mid
        return min_mbs, lo

    def get_tuning_micro_batch_size_list(self, min_mbs: int, max_mbs: int, num: int) -> List[int]:
        """Up to ``num`` sizes to measure in ``[min_mbs, max_mbs]``: evenly spaced, always including both ends."""
        if min_mbs <= 0 or max_mbs < min_mbs:
            return []
        if num <= 1 or max_mbs == min_mbs:
            return [max_mbs]
        step = (max_mbs - min_mbs) / (num - 1)
        out = sorted({int(round(min_mbs + i * step)) for i in range(num)})
        return [m for m in out if min_mbs <= m <= max_mbs]

    def run_tuning_micro_batch_sizes(self, mbs_list: List[int], gas: int, stage: int, space_name: str) -> Tuple[Optional[int], Optional[float]]:
        """Measure the sizes in increasing order and stop at the plateau; returns (best size, its metric)."""
        best_m, best_v = None, None
        already = {r[0]["train_micro_batch_size_per_gpu"]: r[1] for r in self.records.get(space_name, [])}
        for m in sorted(mbs_list):
            if m in already:
                val = already[m]
            else:
                cfg = self._stage_config(stage, m)
                val = self.run_ds_config(cfg, f"{space_name}_gas{gas}_mbs{m}")
                self.update_records(space_name, cfg, val)
            if val is None:
                break                                              # larger sizes will not run either
            if self._better(val, best_v):
                gain = abs(val - best_v) / max(abs(best_v), 1e-12) if best_v is not None else 1.0
                best_m, best_v = m, val
                if gain < METRIC_PLATEAU_RATIO:
                    break
            else:
                break
        return best_m, best_v

    def get_plateau_mbs(self, space_name: str) -> Optional[int]:
        """The micro-batch size after which the metric of ``space_name`` stopped improving."""
        rows = sorted(((r[0]["train_micro_batch_size_per_gpu"], r[1]) for r in self.records.get(space_name, []) if r[1] is not None))
        best_m, best_v = None, None
        for m, v in rows:
            if self._better(v, best_v):
                gain = abs(v - best_v) / max(abs(best_v), 1e-12) if best_v is not None else 1.0
                best_m, best_v = m, v
                if gain < METRIC_PLATEAU_RATIO:
                    break
            else:
                break
        return best_m

    # ---- one stage ---------------------------------------------------------------------------------------------------------
    def tune_space(self, stage: int, prev_max_mbs: int = 0, prev_best_mbs: int = 0, prev_best_metric_val: Optional[float] = None):
        """Tune ZeRO ``stage``; returns (largest runnable mbs, best mbs, best metric) or (0, 0, None) if nothing ran."""
        from .autotuner import DEFAULT_TUNING_SPACE, _expand
        from .tuner import GridSearchTuner, ModelBasedTuner, RandomTuner
        name = f"z{stage}"
        if self.mbs_list:
            lo, hi, sizes = min(self.mbs_list), max(self.mbs_list), sorted(self.mbs_list)
        else:
            start = max(self.min_train_micro_batch_size_per_gpu(), 1)
            lo, hi = self.get_min_max_micro_batch_size(stage, start, self.max_train_micro_batch_size_per_gpu())
            if lo < 0:
                logger.info(f"autotuning: {name}: the smallest micro-batch size does not run; stage skipped")
                return 0, 0, None
            # a later stage frees memory: start its list where the previous stage's best was
            sizes = self.get_tuning_micro_batch_size_list(max(lo, min(prev_best_mbs, hi)) if prev_best_mbs else lo, hi,
                                                          self.num_tuning_micro_batch_sizes())
        best_mbs, best_val = self.run_tuning_micro_batch_sizes(sizes, self.get_gas_from_user_config(), stage, name)
        if best_mbs is None:
            return 0, 0, None
        if not self.fast_enabled():
            space = copy.deepcopy(DEFAULT_TUNING_SPACE[stage])
            exps = [self._stage_config(stage, best_mbs, knobs) for knobs in _expand(space)]
            cls = {"gridsearch": GridSearchTuner, "random": RandomTuner, "model_based": ModelBasedTuner}[self.tuner_type]
            tuner = cls(exps, lambda cfg: self.run_ds_config(cfg, f"{name}_knobs"), self._metric)
            n = tuner.tune(sample_size=1, n_trials=self.num_trials, early_stopping=self.early_stopping)
            for exp, val in tuner.records:
                self.update_records(name, exp, val)
            if self._better(tuner.best_metric_val, best_val):
                best_val = tuner.best_metric_val
            logger.info(f"autotuning: {name}: {n} knob experiments at mbs={best_mbs}")
        return hi, best_mbs, best_val

    def tune_staged(self):
        prev_max, prev_best_mbs, prev_best_val = 0, 0, None
        for stage in self.feasible_stages():
            hi, mbs, val = self.tune_space(stage, prev_max, prev_best_mbs, prev_best_val)
            if val is None:
                continue
            if self.fast_enabled() and prev_best_val is not None and not self._better(val, prev_best_val):
                logger.info(f"autotuning: z{stage} does not improve on the previous stage; stopping (fast mode)")
                break
            if self._better(val, prev_best_val):
                prev_max, prev_best_mbs, prev_best_val = hi, mbs, val
        best = self.get_best_space_records().get(GLOBAL_SPACE)
        if best is None:
            logger.warning("autotuning: no experiment produced a metric")
            return None
        self.best_config, self.best_metric = best[0], best[1]
        self.records["all"] = [(e, v) for name, rows in self.records.items() if name != "all" for e, v, _ in rows]
        return self.best_config

    # ---- before / after ------------------------------------------------------------------------------------------------------
    def model_info_profile_run(self) -> Optional[Dict]:
        """One short run that only reports the model size (``autotuning.model_info.profile``); fills ``num_params``."""
        from .autotuner import _merge
        import json
        import os
        if self.run_fn is not None or self.user_script is None:
            return self.model_info
        os.makedirs(self.results_dir, exist_ok=True)
        info_path = os.path.join(self.results_dir, "model_info.json")
        cfg = _merge(self.user_config, {"autotuning": {"enabled": True, "model_info_path": info_path, "model_info": {"profile": True}}})
        self._launch_with_config(cfg, "ds_config_model_info.json")
        if os.path.exists(info_path):
            self.model_info = json.load(open(info_path))
            self.num_params = int(self.model_info.get("num_params", self.num_params))
        return self.model_info

    def run_after_tuning(self) -> Optional[int]:
        """``--autotuning run``: launch the script once more with the best config found."""
        if self.best_config is None or self.user_script is None:
            return None
        return self._launch_with_config(self.best_config, "ds_config_optimal.json")


def _number(text: str):
    if re.fullmatch(r"-?\d+", text):
        return int(text)
    try:
        return float(text)
    except ValueError:
        return text
