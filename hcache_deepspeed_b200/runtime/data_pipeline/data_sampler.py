"""Curriculum-aware distributed batch sampler.

Yields, per data-parallel rank, the indices of each micro-batch. With curriculum learning enabled the data is consumed
cluster by cluster (the scheme of ``deepspeed/runtime/data_pipeline/data_sampling/data_sampler.py:36-340``):

* every time a metric's difficulty moves, the samples that became eligible since the last move (for every metric at
  once: the intersection of the per-metric eligible sets, minus what older clusters already hold) form a NEW cluster;
* a global batch draws from all clusters in proportion to their sizes;
* inside a cluster samples are handed out in a shuffled order WITHOUT replacement; an exhausted cluster is reshuffled.

So easy samples keep being seen after harder ones unlock, each sample of a cluster is used once per pass over it, and
the stream is reproducible from ``state_dict()``. A metric's difficulty is either a value (``difficulty_type: "value"``:
eligible = metric <= difficulty) or a percentile of the data ordered by the metric (``"percentile"``). Clusters live in
memory (one int64 array each); the reference keeps them as mmap files under ``data_cluster_path`` because its indexes do
- set ``data_cluster_path`` and they are also written there, one ``.npy`` per cluster, for inspection.
"""
from __future__ import annotations

import os
from typing import Dict, Iterator, List, Optional

import numpy as np

from .curriculum_scheduler import CurriculumScheduler

VALUE_BASED, PERCENTILE_BASED = "value", "percentile"
SINGLE_CLUSTER, SCHEDULE_BASED = "single_cluster", "schedule_based"


class DeepSpeedDataSampler:

    def __init__(self, data_efficiency_config: Dict, one_epoch_total_samples: int, micro_batch_size: int, data_parallel_rank: int,
                 data_parallel_size: int, data_parallel_group=None, gradient_accumulation_steps: int = 1, global_rank: int = 0,
                 drop_last: bool = True, difficulties: Optional[Dict[str, np.ndarray]] = None):
        self.cfg = data_efficiency_config or {}
        self.total_samples = self.one_epoch_total_samples = one_epoch_total_samples
        self.micro_batch_size, self.dp_rank, self.dp_size = micro_batch_size, data_parallel_rank, data_parallel_size
        self.data_parallel_group, self.global_rank = data_parallel_group, global_rank
        self.gas = gradient_accumulation_steps
        self.micro_batch_times_data_parallel_size = micro_batch_size * data_parallel_size
        self.global_batch_size = self.micro_batch_times_data_parallel_size * gradient_accumulation_steps
        self.drop_last = drop_last
        sampling = self.cfg.get("data_sampling") or {}
        self.seed = int(sampling.get("seed", self.cfg.get("seed", 1234)))
        self.np_rng = np.random.default_rng(self.seed)
        self.consumed_samples = 0
        self.curriculum_step = 0
        self.difficulties = dict(difficulties or {})
        cl = sampling.get("curriculum_learning") or {}
        self.curriculum_enabled = bool(cl.get("enabled", False))
        self.cluster_path = cl.get("data_cluster_path")
        self.curriculum_schedulers: Dict[str, CurriculumScheduler] = {}
        self.difficulty_type: Dict[str, str] = {}
        self.clustering_type: Dict[str, str] = {}
        self.max_difficulty: Dict[str, int] = {}
        self.current_difficulties: Dict[str, int] = {}
        self.data_clusters: List[np.ndarray] = []
        self.data_cluster_current_position: List[int] = []
        self._clustered = np.zeros(self.total_samples, dtype=bool)     # sample already belongs to a cluster
        self._order: Dict[str, np.ndarray] = {}                          # metric -> sample ids sorted by the metric
        if self.curriculum_enabled:
            for metric, mcfg in (cl.get("curriculum_metrics") or {}).items():
                self.curriculum_schedulers[metric] = CurriculumScheduler(mcfg)
                self.difficulty_type[metric] = mcfg.get("difficulty_type", VALUE_BASED)
                self.clustering_type[metric] = mcfg.get("clustering_type", SCHEDULE_BASED)
                self.max_difficulty[metric] = mcfg.get("max_difficulty", 100)
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

    def get_start_end_idx(self, batch_len: Optional[int] = None):
        """This rank's [start, end) inside a mini-batch of ``batch_len`` samples (default: micro batch x dp size)."""
        n = batch_len or self.micro_batch_times_data_parallel_size
        at = lambda r: round(r * n / self.dp_size)     # noqa: E731
        return at(self.dp_rank), at(self.dp_rank + 1)

    # ---- eligible sets -------------------------------------------------------------------------------------------------------
    def get_sample_based_on_metric_value(self, metric: str, value_start, value_end) -> np.ndarray:
        """Samples with ``value_start < metric <= value_end``."""
        d = self.difficulties[metric]
        return np.nonzero((d > value_start) & (d <= value_end))[0].astype(np.int64)

    def get_sample_based_on_metric_percentile(self, metric: str, percentile_start, percentile_end) -> np.ndarray:
        """Samples between two percentiles (of ``max_difficulty`` steps) of the data ordered by the metric."""
        if metric not in self._order:
            self._order[metric] = np.argsort(self.difficulties[metric], kind="stable").astype(np.int64)
        steps = max(1, int(self.max_difficulty[metric]))
        per = self.total_samples // steps
        lo = per * int(percentile_start)
        hi = self.total_samples if percentile_end >= steps else per * int(percentile_end)
        return self._order[metric][lo:hi]

    def _eligible_of(self, metric: str, previous) -> np.ndarray:
        if self.clustering_type[metric] == SINGLE_CLUSTER or metric not in self.difficulties:
            return np.arange(self.total_samples, dtype=np.int64)
        cur = self.current_difficulties[metric]
        if self.difficulty_type[metric] == PERCENTILE_BASED:
            return self.get_sample_based_on_metric_percentile(metric, 0, cur)
        return self.get_sample_based_on_metric_value(metric, float("-inf"), cur)

    def get_new_cluster(self, previous_difficulties: Dict[str, int]) -> Optional[np.ndarray]:
        """The samples eligible under every metric's current difficulty that no older cluster holds; appended as a new
        (shuffled) cluster. ``None`` when the move unlocked nothing."""
        keep = np.ones(self.total_samples, dtype=bool)
        for metric in self.curriculum_schedulers:
            mask = np.zeros(self.total_samples, dtype=bool)
            mask[self._eligible_of(metric, previous_difficulties.get(metric))] = True
            keep &= mask
        keep &= ~self._clustered
        new = np.nonzero(keep)[0].astype(np.int64)
        if new.size == 0:
            return None
        self._clustered[new] = True
        self.np_rng.shuffle(new)
        self.data_clusters.append(new)
        self.data_cluster_current_position.append(0)
        if self.cluster_path and self.global_rank == 0:
            os.makedirs(self.cluster_path, exist_ok=True)
            name = "cluster" + "".join(f"_{m}{self.current_difficulties[m]}" for m in self.curriculum_schedulers)
            np.save(os.path.join(self.cluster_path, name + ".npy"), new)
        return new

    # ---- drawing ---------------------------------------------------------------------------------------------------------------
    @property
    def data_cluster_sizes(self) -> List[int]:
        return [len(c) for c in self.data_clusters]

    def sample_from_clusters(self) -> np.ndarray:
        """How many samples of the next global batch come from each cluster (multinomial over the cluster sizes)."""
        sizes = np.asarray(self.data_cluster_sizes, dtype=np.float64)
        picks = self.np_rng.choice(len(sizes), self.global_batch_size, replace=True, p=sizes / sizes.sum())
        return np.bincount(picks, minlength=len(sizes))

    def reshuffle_clusters(self, cidx: int) -> None:
        self.np_rng.shuffle(self.data_clusters[cidx])

    def get_sample_from_cluster(self, cidx: int, num_samples: int) -> List[int]:
        """The next ``num_samples`` of cluster ``cidx`` in its shuffled order; wraps around with a reshuffle."""
        out: List[int] = []
        cluster = self.data_clusters[cidx]
        while len(out) < num_samples:
            pos = self.data_cluster_current_position[cidx]
            take = cluster[pos:pos + num_samples - len(out)]
            out += take.tolist()
            self.data_cluster_current_position[cidx] = pos + len(take)
            if len(out) < num_samples:
                self.reshuffle_clusters(cidx)
                self.data_cluster_current_position[cidx] = 0
        return out

    def _eligible(self) -> np.ndarray:
        """Union of the clusters (every sample eligible so far)."""
        return np.nonzero(self._clustered)[0] if self.curriculum_enabled else np.arange(self.total_samples)

    def get_next_global_batch(self) -> np.ndarray:
        if not self.curriculum_enabled:
            pool = np.arange(self.total_samples)
            return self.np_rng.choice(pool, size=self.global_batch_size, replace=pool.size < self.global_batch_size)
        self.curriculum_step += 1
        moved, previous = False, dict(self.current_difficulties)
        for metric, sched in self.curriculum_schedulers.items():
            nxt = sched.update_difficulty(self.curriculum_step)
            if self.current_difficulties.get(metric) != nxt:
                moved = True
            self.current_difficulties[metric] = nxt
        if moved:
            self.get_new_cluster(previous)
        if not self.data_clusters:
            # nothing is eligible yet (difficulty below every sample): fall back to the easiest global batch
            order = np.argsort(sum(self.difficulties.values())) if self.difficulties else np.arange(self.total_samples)
            easiest = order[:self.global_batch_size].astype(np.int64)
            self._clustered[easiest] = True
            self.data_clusters.append(easiest.copy())
            self.data_cluster_current_position.append(0)
        batch: List[int] = []
        for cidx, n in enumerate(self.sample_from_clusters()):
            if n:
                batch += self.get_sample_from_cluster(cidx, int(n))
        out = np.asarray(batch, dtype=np.int64)
        self.np_rng.shuffle(out)
        return out

    def __iter__(self) -> Iterator[List[int]]:
        while self.consumed_samples <= self.total_samples - (self.global_batch_size if self.drop_last else 1):
            g = self.get_next_global_batch()
            per_step = self.micro_batch_times_data_parallel_size
            for a in range(self.gas):
                chunk = g[a * per_step:(a + 1) * per_step]
                lo, hi = self.get_start_end_idx(len(chunk))
                yield chunk[lo:hi].tolist()
            self.consumed_samples += self.global_batch_size

    # ---- checkpoint ----------------------------------------------------------------------------------------------------------------
    def state_dict(self):
        return {"consumed_samples": self.consumed_samples, "curriculum_step": self.curriculum_step,
                "np_rng_state": self.np_rng.bit_generator.state,
                "current_difficulties": {m: s.get_current_difficulty() for m, s in self.curriculum_schedulers.items()},
                "data_clusters": [c.copy() for c in self.data_clusters],
                "data_cluster_current_position": list(self.data_cluster_current_position)}

    def load_state_dict(self, sd):
        self.consumed_samples, self.curriculum_step = sd["consumed_samples"], sd["curriculum_step"]
        self.np_rng.bit_generator.state = sd["np_rng_state"]
        for m, d in sd.get("current_difficulties", {}).items():
            if m in self.curriculum_schedulers:
                self.curriculum_schedulers[m].set_current_difficulty(d)
                self.current_difficulties[m] = d
        if "data_clusters" in sd:
            self.data_clusters = [np.asarray(c, dtype=np.int64).copy() for c in sd["data_clusters"]]
            self.data_cluster_current_position = list(sd["data_cluster_current_position"])
            self._clustered[:] = False
            for c in self.data_clusters:
                self._clustered[c] = True
