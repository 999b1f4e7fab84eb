"""Curriculum-aware distributed batch sampler: yields, per data-parallel rank, the indices of each micro-batch; with
curriculum learning enabled only samples whose difficulty metric is <= the current difficulty are eligible.
Reference: ``deepspeed/runtime/data_pipeline/data_sampling/data_sampler.py:36-340`` (DeepSpeedDataSampler)."""
from __future__ import annotations

from typing import Dict, Iterator, List, Optional

import numpy as np

from .curriculum_scheduler import CurriculumScheduler


class DeepSpeedDataSampler:

    def __init__(self, data_efficiency_config: Dict, one_epoch_total_samples: int, micro_batch_size: int, data_parallel_rank: int,
                 data_parallel_size: int, data_parallel_group=None, gradient_accumulation_steps: int = 1, global_rank: int = 0,
                 drop_last: bool = True, difficulties: Optional[Dict[str, np.ndarray]] = None):
        self.cfg = data_efficiency_config or {}
        self.total_samples = one_epoch_total_samples
        self.micro_batch_size, self.dp_rank, self.dp_size = micro_batch_size, data_parallel_rank, data_parallel_size
        self.gas = gradient_accumulation_steps
        self.global_batch_size = micro_batch_size * data_parallel_size * gradient_accumulation_steps
        self.drop_last = drop_last
        self.seed = int((self.cfg.get("data_sampling") or {}).get("seed", self.cfg.get("seed", 1234)))
        self.np_rng = np.random.default_rng(self.seed)
        self.consumed_samples = 0
        self.curriculum_step = 0
        self.difficulties = difficulties or {}
        cl = ((self.cfg.get("data_sampling") or {}).get("curriculum_learning") or {})
        self.curriculum_enabled = bool(cl.get("enabled", False))
        self.curriculum_schedulers: Dict[str, CurriculumScheduler] = {}
        if self.curriculum_enabled:
            for metric, mcfg in (cl.get("curriculum_metrics") or {}).items():
                self.curriculum_schedulers[metric] = CurriculumScheduler(mcfg)
                if metric not in self.difficulties and mcfg.get("index_to_sample_path") and mcfg.get("index_to_metric_path"):
                    self.difficulties[metric] = self._load_difficulties(mcfg["index_to_sample_path"],
                                                                        mcfg["index_to_metric_path"])
        self.batch: List[int] = []

    def _load_difficulties(self, index_to_sample_path: str, index_to_metric_path: str) -> np.ndarray:
        """Per-sample difficulty from the data analyzer's ``*_index_to_sample`` / ``*_index_to_metric`` files."""
        from .data_sampling.indexed_dataset import MMapIndexedDataset
        i2s, i2m = MMapIndexedDataset(index_to_sample_path), MMapIndexedDataset(index_to_metric_path)
        out = np.zeros(self.total_samples, dtype=np.int64)
        for j in range(len(i2s)):
            out[np.asarray(i2s[j], dtype=np.int64)] = int(i2m[j][0])
        return out

    def __len__(self):
        return self.total_samples

    def set_custom_curriculum_learning_schedule(self, schedule_func_dict):
        for metric, fn in schedule_func_dict.items():
            if metric in self.curriculum_schedulers:
                self.curriculum_schedulers[metric].set_custom_get_difficulty(fn)

    def _eligible(self) -> np.ndarray:
        idx = np.arange(self.total_samples)
        if not self.curriculum_enabled:
            return idx
        keep = np.ones(self.total_samples, dtype=bool)
        for metric, sched in self.curriculum_schedulers.items():
            d = sched.update_difficulty(self.curriculum_step)
            if metric in self.difficulties:
                keep &= self.difficulties[metric] <= d
        out = idx[keep]
        return out if out.size >= self.global_batch_size else idx[np.argsort(sum(self.difficulties.values()))[:self.global_batch_size]]

    def get_next_global_batch(self) -> np.ndarray:
        self.curriculum_step += 1
        pool = self._eligible()
        return self.np_rng.choice(pool, size=self.global_batch_size, replace=pool.size < self.global_batch_size)

    def __iter__(self) -> Iterator[List[int]]:
        while self.consumed_samples <= self.total_samples - (self.global_batch_size if self.drop_last else 1):
            g = self.get_next_global_batch()
            per_step = self.micro_batch_size * self.dp_size
            for a in range(self.gas):
                chunk = g[a * per_step:(a + 1) * per_step]
                lo = self.dp_rank * self.micro_batch_size
                yield chunk[lo:lo + self.micro_batch_size].tolist()
            self.consumed_samples += self.global_batch_size

    def state_dict(self):
        return {"consumed_samples": self.consumed_samples, "curriculum_step": self.curriculum_step,
                "np_rng_state": self.np_rng.bit_generator.state,
                "current_difficulties": {m: s.get_current_difficulty() for m, s in self.curriculum_schedulers.items()}}

    def load_state_dict(self, sd):
        self.consumed_samples, self.curriculum_step = sd["consumed_samples"], sd["curriculum_step"]
        self.np_rng.bit_generator.state = sd["np_rng_state"]
        for m, d in sd.get("current_difficulties", {}).items():
            if m in self.curriculum_schedulers:
                self.curriculum_schedulers[m].set_current_difficulty(d)
