"""Offline data analysis for curriculum learning: evaluate difficulty metrics over a dataset (map, sharded over
workers) and build the indexes the sampler consumes (reduce).

Per ``single_value_per_sample`` metric ``m`` the reduce step writes, under ``<save_path>/<m>/`` (mmap indexed datasets):
``m_sample_to_metric`` (item i = metric of sample i), ``m_index_to_metric`` (item j = j-th smallest distinct value),
``m_index_to_sample`` (item j = samples with that value) and ``m_index_to_sample_percentile_merged`` (~100 buckets of
samples in ascending difficulty). Per ``accumulate_value_over_samples`` metric: ``m_metric_value`` (the summed vector).

Reference: ``data_sampling/data_analyzer.py:22-452`` (multi-process map / per-thread files / merge) and ``:455-730``
(``DistributedDataAnalyzer``). File names and contents match; the reduce here is one vectorised sort instead of a
per-value file merge.
"""
from __future__ import annotations

import os
from typing import Callable, List, Optional, Sequence

import numpy as np
import torch

from ....utils.logging import logger
from .indexed_dataset import MMapIndexedDataset
from .utils import close_mmap_dataset_builder, create_mmap_dataset_builder, find_fit_int_dtype, split_index

SINGLE, ACCUMULATE = "single_value_per_sample", "accumulate_value_over_samples"


def _write_items(fname: str, items: Sequence[np.ndarray], dtype) -> None:
    os.makedirs(os.path.dirname(fname), exist_ok=True)
    b = create_mmap_dataset_builder(fname, dtype)
    for it in items:
        b.add_item_numpy(np.asarray(it))
    close_mmap_dataset_builder(b, fname)


def write_metric_indexes(values: np.ndarray, metric_name: str, save_path: str) -> dict:
    """Reduce step for one single-value metric given the value of every sample (in dataset order)."""
    values = np.asarray(values).reshape(-1).astype(np.int64)
    n = values.size
    mdir = os.path.join(save_path, metric_name)
    vdtype = find_fit_int_dtype(int(values.min()), int(values.max()))
    sdtype = find_fit_int_dtype(0, max(0, n - 1))
    _write_items(f"{mdir}/{metric_name}_sample_to_metric", values.reshape(-1, 1), vdtype)
    order = np.argsort(values, kind="stable")
    uniq, starts = np.unique(values[order], return_index=True)
    groups = np.split(order, starts[1:])
    _write_items(f"{mdir}/{metric_name}_index_to_metric", uniq.reshape(-1, 1), vdtype)
    _write_items(f"{mdir}/{metric_name}_index_to_sample", groups, sdtype)
    step = max(1, len(groups) // 100)
    merged = [np.concatenate(groups[i:i + step]) for i in range(0, len(groups), step)]
    _write_items(f"{mdir}/{metric_name}_index_to_sample_percentile_merged", merged, sdtype)
    return {int(v): int(len(g)) for v, g in zip(uniq, groups)}


def metric_value_percentiles(num_sample_per_value: dict, total: int) -> dict:
    """{percentile: metric value} for every 5th percentile (the table the reference logs)."""
    out, seen, keys = {}, 0, sorted(num_sample_per_value)
    targets = list(range(5, 101, 5))
    for v in keys:
        seen += num_sample_per_value[v]
        while targets and seen * 100 >= targets[0] * total:
            out[targets.pop(0)] = v
    return out


class DataAnalyzer:

    def __init__(self, dataset, num_workers: int = 1, worker_id: int = 0, num_threads: int = 1, num_threads_reduce: int = 1,
                 specific_threads: Sequence[int] = (), batch_size: int = 1, metric_names: Sequence[str] = (),
                 metric_functions: Sequence[Callable] = (), metric_types: Sequence[str] = (), metric_dtypes: Sequence = (),
                 save_path: str = "./", collate_fn=None, custom_map_init=None, custom_map_update=None,
                 custom_map_finalize=None, custom_reduce=None, sample_indices=None):
        self.dataset, self.num_workers, self.worker_id = dataset, num_workers, worker_id
        self.batch_size, self.collate_fn, self.save_path = batch_size, collate_fn, save_path
        self.metric_names, self.metric_functions = list(metric_names), list(metric_functions)
        self.metric_types, self.metric_dtypes = list(metric_types), list(metric_dtypes)
        self.custom_reduce, self.sample_indices = custom_reduce, sample_indices
        assert all(t in (SINGLE, ACCUMULATE) for t in self.metric_types), f"metric types must be {SINGLE} or {ACCUMULATE}"

    def _worker_file(self, metric: str, worker: int) -> str:
        return f"{self.save_path}/{metric}/worker{worker}_thread0/{metric}_" + \
            ("sample_to_metric" if self.metric_types[self.metric_names.index(metric)] == SINGLE else "metric_value")

    def _evaluate(self, start: int, end: int) -> List[np.ndarray]:
        """Metric values for samples [start, end): one int64 array per metric."""
        loader = torch.utils.data.DataLoader(torch.utils.data.Subset(self.dataset, range(start, end)),
                                             batch_size=self.batch_size, shuffle=False, collate_fn=self.collate_fn)
        per_metric: List[list] = [[] for _ in self.metric_names]
        acc: List[Optional[np.ndarray]] = [None] * len(self.metric_names)
        for batch in loader:
            for i, (fn, mtype) in enumerate(zip(self.metric_functions, self.metric_types)):
                v = fn(batch)
                v = v.detach().cpu().numpy() if torch.is_tensor(v) else np.asarray(v)
                if mtype == SINGLE:
                    per_metric[i].append(v.reshape(-1).astype(np.int64))
                else:
                    acc[i] = v.astype(np.int64) if acc[i] is None else acc[i] + v
        return [np.concatenate(per_metric[i]) if t == SINGLE and per_metric[i] else
                (np.empty(0, np.int64) if t == SINGLE else acc[i]) for i, t in enumerate(self.metric_types)]

    def run_map(self) -> None:
        start, end = split_index(0, len(self.dataset), self.num_workers)[self.worker_id]
        logger.info(f"data analyzer worker {self.worker_id}: samples [{start}, {end})")
        for name, mtype, vals in zip(self.metric_names, self.metric_types, self._evaluate(start, end)):
            _write_items(self._worker_file(name, self.worker_id), [vals] if vals is not None else [], np.int64)

    def run_reduce(self) -> None:
        if self.custom_reduce is not None:
            return self.custom_reduce(self.dataset, self.metric_names, self.metric_types, self.save_path, self.num_workers, 1, 1)
        for name, mtype in zip(self.metric_names, self.metric_types):
            parts = [MMapIndexedDataset(self._worker_file(name, w)) for w in range(self.num_workers)]
            parts = [np.copy(p[0]) for p in parts if len(p)]
            if mtype == SINGLE:
                values = np.concatenate(parts)
                assert values.size == len(self.dataset), f"{name}: {values.size} values for {len(self.dataset)} samples"
                counts = write_metric_indexes(values, name, self.save_path)
                logger.info(f"{name} percentiles: {metric_value_percentiles(counts, values.size)}")
            else:
                total = np.sum(parts, axis=0)
                _write_items(f"{self.save_path}/{name}/{name}_metric_value", [total],
                             find_fit_int_dtype(int(total.min()), int(total.max())))

    def run_map_reduce(self, comm_group=None) -> None:
        self.run_map()
        if torch.distributed.is_available() and torch.distributed.is_initialized():
            torch.distributed.barrier(group=comm_group)
        if self.worker_id == 0:
            self.run_reduce()
        if torch.distributed.is_available() and torch.distributed.is_initialized():
            torch.distributed.barrier(group=comm_group)


class DistributedDataAnalyzer(DataAnalyzer):
    """Every rank of ``comm_group`` maps its contiguous shard; values are gathered to rank 0, which writes the indexes
    (no intermediate per-worker files)."""

    def __init__(self, dataset, num_workers: int = 1, worker_id: int = 0, batch_size: int = 1, metric_names=(),
                 metric_functions=(), metric_types=(), save_path: str = "./", collate_fn=None, device="cpu", comm_group=None,
                 **kwargs):
        import torch.distributed as dist
        self.comm_group = comm_group
        if dist.is_available() and dist.is_initialized():
            num_workers, worker_id = dist.get_world_size(comm_group), dist.get_rank(comm_group)
        super().__init__(dataset, num_workers=num_workers, worker_id=worker_id, batch_size=batch_size, metric_names=metric_names,
                         metric_functions=metric_functions, metric_types=metric_types, save_path=save_path, collate_fn=collate_fn)

    def run_map_reduce(self, comm_group=None) -> None:
        import torch.distributed as dist
        start, end = split_index(0, len(self.dataset), self.num_workers)[self.worker_id]
        local = self._evaluate(start, end)
        live = dist.is_available() and dist.is_initialized() and self.num_workers > 1
        for name, mtype, vals in zip(self.metric_names, self.metric_types, local):
            if live:
                gathered = [None] * self.num_workers
                dist.all_gather_object(gathered, vals, group=self.comm_group)
            else:
                gathered = [vals]
            if self.worker_id != 0:
                continue
            if mtype == SINGLE:
                write_metric_indexes(np.concatenate(gathered), name, self.save_path)
            else:
                total = np.sum([g for g in gathered if g is not None], axis=0)
                _write_items(f"{self.save_path}/{name}/{name}_metric_value", [total],
                             find_fit_int_dtype(int(total.min()), int(total.max())))
        if live:
            dist.barrier(group=self.comm_group)


def load_sample_to_metric(save_path: str, metric_name: str) -> np.ndarray:
    """Difficulty of every sample, as consumed by ``DeepSpeedDataSampler(difficulties={metric: ...})``."""
    ds = MMapIndexedDataset(f"{save_path}/{metric_name}/{metric_name}_sample_to_metric")
    return np.concatenate([ds[i] for i in range(len(ds))]).astype(np.int64) if len(ds) else np.empty(0, np.int64)
