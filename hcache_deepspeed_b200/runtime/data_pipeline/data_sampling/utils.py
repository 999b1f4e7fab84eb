"""Helpers shared by the data analyzer and sampler (``data_sampling/utils.py:13-52``)."""
from __future__ import annotations

from typing import List, Tuple

import numpy as np

from .indexed_dataset import MMapIndexedDatasetBuilder


def find_fit_int_dtype(min_value: int, max_value: int):
    """Narrowest integer dtype that holds [min_value, max_value] (unsigned when possible)."""
    cands = (np.uint8, np.uint16, np.uint32, np.uint64) if min_value >= 0 else (np.int8, np.int16, np.int32, np.int64)
    for dt in cands:
        info = np.iinfo(dt)
        if info.min <= min_value and max_value <= info.max:
            return dt
    return cands[-1]


def split_index(start_idx: int, end_idx: int, num_partitions: int) -> List[Tuple[int, int]]:
    b = np.linspace(start_idx, end_idx, dtype=int, num=num_partitions + 1)
    return [(int(b[i]), int(b[i + 1])) for i in range(num_partitions)]


def split_dataset(dataset, num_workers: int, worker_id: int, num_threads: int):
    worker_splits = split_index(0, len(dataset), num_workers)
    return worker_splits, split_index(*worker_splits[worker_id], num_threads)


def create_mmap_dataset_builder(fname: str, dtype) -> MMapIndexedDatasetBuilder:
    return MMapIndexedDatasetBuilder(f"{fname}.bin", dtype=dtype)


def close_mmap_dataset_builder(builder: MMapIndexedDatasetBuilder, fname: str) -> None:
    builder.end_document()
    builder.finalize(f"{fname}.idx")
