"""Data sampling: curriculum-aware sampler, mmap indexed datasets, offline data analyzer, token-budget batching."""
from ..data_sampler import DeepSpeedDataSampler  # noqa: F401
from .data_analyzer import DataAnalyzer, DistributedDataAnalyzer, load_sample_to_metric  # noqa: F401
from .indexed_dataset import MMapIndexedDataset, MMapIndexedDatasetBuilder  # noqa: F401
from .variable_batch_size_and_lr import (VariableBatchSizeLR, batch_by_seqlens, dataloader_for_variable_batch_size,  # noqa: F401
                                         get_dataloader_and_lr_scheduler_for_variable_batch_size,
                                         lr_scheduler_for_variable_batch_size, scale_lr)
