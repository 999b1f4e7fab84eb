"""Token-budget batching with batch-size-aware learning-rate scaling: pack samples into micro-batches whose summed
length stays under ``max_tokens`` (so batches hold a varying NUMBER of sequences), hand each data-parallel replica its
share, and scale the LR of every step by that step's effective batch size.

Reference: ``data_sampling/variable_batch_size_and_lr.py:23-146`` (``batch_by_seqlens``), ``:149-162`` (``scale_lr``),
``:165-223`` (dataloader), ``:226-348`` (scheduler), ``:351-492`` (engine glue). The packing here is one greedy linear
pass (the constraints are monotone in the batch length once ``min_batch_size`` is met) and it may use the final sample
of the dataset, which the reference's ``range(..., len(metrics), ...)`` scan never reaches.
"""
from __future__ import annotations

import math
import random
from typing import Callable, List, Optional, Sequence, Tuple

import numpy as np
import torch
from torch.optim import Optimizer
from torch.optim.lr_scheduler import LRScheduler
from torch.utils.data import DataLoader, DistributedSampler

from ....utils.logging import logger


def batch_by_seqlens(seqlens: Sequence[int], max_tokens: int, sequence_ids_per_mb: Optional[Sequence[int]] = None,
                     min_batch_size: int = 1, max_batch_size: Optional[int] = None, sequence_picking_order: str = "dataloader",
                     effective_batch_size: int = 1, required_microbatches_of_same_size: bool = False, verbose: bool = False,
                     seed: Optional[int] = None) -> Tuple[List[Tuple[int, List[int]]], List[int], List[int]]:
    """Returns (``[(batch_id, sample ids of one micro-batch), ...]``, sequences per batch, longest sequence per batch);
    ``effective_batch_size`` micro-batches (one per replica x accumulation step) make one batch."""
    assert sequence_picking_order in ("random", "seqlen", "dataloader")
    ids = list(range(len(seqlens))) if sequence_ids_per_mb is None else list(sequence_ids_per_mb)
    items = [(int(seqlens[i]), i) for i in ids]
    if sequence_picking_order == "random":
        random.Random(seed).shuffle(items)
    elif sequence_picking_order == "seqlen":
        items.sort()
    too_long = [i for v, i in items if v > max_tokens]
    if too_long:
        logger.warning(f"Data indices {too_long} ignored as metrics exceed {max_tokens}.")
        items = [(v, i) for v, i in items if v <= max_tokens]

    mult = effective_batch_size if required_microbatches_of_same_size else 1
    microbatches: List[List[Tuple[int, int]]] = []
    pos = 0
    while pos < len(items):
        # grow the group `mult` samples at a time; micro-batch b takes samples b, b + mult, ... of the group
        tokens, best = [0] * mult, 0
        for end in range(pos + mult, len(items) + 1, mult):
            for b in range(mult):
                tokens[b] += items[end - mult + b][0]
            per_mb = (end - pos) // mult
            if max(tokens) > max_tokens or (max_batch_size and per_mb > max_batch_size):
                break
            if per_mb >= (min_batch_size or 1):
                best = end
        if best == 0:
            break
        group = items[pos:best]
        microbatches += [group[b::mult] for b in range(mult)]
        pos = best
    assert len(microbatches) >= effective_batch_size, "not enough datapoints to create a single sample per dataloader"
    microbatches = microbatches[:len(microbatches) - len(microbatches) % effective_batch_size]

    microbatch_ids, batch_sizes, batch_max_seqlens = [], [], []
    for first in range(0, len(microbatches), effective_batch_size):
        mbs = microbatches[first:first + effective_batch_size]
        batch_id = first // effective_batch_size
        batch_sizes.append(sum(len(mb) for mb in mbs))
        batch_max_seqlens.append(max(v for mb in mbs for v, _ in mb))
        microbatch_ids += [(batch_id, [i for _, i in mb]) for mb in mbs]
        if verbose:
            logger.info(f"batch {batch_id}: {len(mbs)} micro-batches, {batch_sizes[-1]} sequences, "
                        f"tokens per micro-batch {[sum(v for v, _ in mb) for mb in mbs]}")
    return microbatch_ids, batch_sizes, batch_max_seqlens


def scale_lr(base_batch_size: int, batch_size: int, base_lr: float = 1.0, method: Optional[str] = "linear") -> float:
    """Linear (Goyal et al.) or square-root (Krizhevsky) scaling of ``base_lr`` from ``base_batch_size`` to ``batch_size``."""
    if method is None or str(method).upper() == "NONE":
        return base_lr
    if method == "linear":
        return base_lr * batch_size / base_batch_size
    if method == "sqrt":
        return base_lr * math.sqrt(batch_size / base_batch_size)
    raise ValueError(f"Unknown scaling method: {method}")


def dataloader_for_variable_batch_size(dataset, microbatch_ids, batch_max_seqlens, dataloader_rank: int = 0,
                                       dataloader_batch_size: int = 1, dataloader_num_replicas: int = 1,
                                       dataloader_collate_fn: Optional[Callable] = None, dataloader_num_workers: int = 0,
                                       dataloader_pin_memory: bool = False, required_microbatches_of_same_seqlen: bool = False,
                                       sample_padding_fn: Optional[Callable] = None):
    """A DataLoader over MICRO-BATCHES (each element = the sample ids of one), interleaved over replicas."""
    sampler = DistributedSampler(dataset=microbatch_ids, num_replicas=dataloader_num_replicas, rank=dataloader_rank,
                                 shuffle=False, drop_last=False)

    def collate(list_microbatch_ids):
        batch = []
        for batch_id, sample_ids in list_microbatch_ids:
            data = [dataset[i] for i in sample_ids]
            if required_microbatches_of_same_seqlen:
                assert sample_padding_fn is not None, "sample_padding_fn is required to pad micro-batches to one length"
                data = [sample_padding_fn(s, batch_max_seqlens[batch_id]) for s in data]
            batch += data
        return dataloader_collate_fn(batch) if dataloader_collate_fn else batch

    loader = DataLoader(dataset=microbatch_ids, batch_size=dataloader_batch_size, sampler=sampler,
                        num_workers=dataloader_num_workers, collate_fn=collate, pin_memory=dataloader_pin_memory)
    io_kwargs = dict(dataset=microbatch_ids, batch_size=dataloader_batch_size, pin_memory=dataloader_pin_memory,
                     data_sampler=sampler, collate_fn=collate, num_local_io_workers=dataloader_num_workers)
    return loader, io_kwargs


class _ConstantLR(LRScheduler):
    def get_lr(self):
        return self.base_lrs


class VariableBatchSizeLR(LRScheduler):
    """Wraps a scheduler: after the base scheduler sets the step's LR, multiply it by the scale for that step's batch."""

    def __init__(self, lr_scheduler, base_batch_size: int, batch_sizes: Sequence[int], dataloader=None,
                 lr_scaling_method: str = "linear", last_epoch: int = -1, verbose: bool = False):
        self.base_lr_scheduler, self.base_batch_size = lr_scheduler, base_batch_size
        self.batch_sizes, self.lr_scaling_method, self.dataloader = list(batch_sizes), lr_scaling_method, dataloader
        self.base_lrs = list(lr_scheduler.get_lr())
        self.last_epoch, self.verbose = last_epoch, verbose
        self._unscaled = list(self.base_lrs)
        self.step(0)

    @property
    def optimizer(self):
        return self.base_lr_scheduler.optimizer

    def state_dict(self):
        return {"base_lr_scheduler": self.base_lr_scheduler.state_dict(), "base_batch_size": self.base_batch_size,
                "lr_scaling_method": self.lr_scaling_method, "batch_sizes": self.batch_sizes, "base_lrs": self.base_lrs,
                "last_epoch": self.last_epoch, "verbose": self.verbose, "unscaled": self._unscaled}

    def load_state_dict(self, sd):
        self.base_lr_scheduler.load_state_dict(sd["base_lr_scheduler"])
        self.base_batch_size, self.lr_scaling_method = sd["base_batch_size"], sd["lr_scaling_method"]
        self.batch_sizes, self.base_lrs, self.last_epoch = sd["batch_sizes"], sd["base_lrs"], sd["last_epoch"]
        self.verbose, self._unscaled = sd["verbose"], sd.get("unscaled", sd["base_lrs"])

    def get_last_lr(self):
        return list(self._unscaled)

    def get_lr(self):
        return [g["lr"] for g in self.optimizer.param_groups]

    def step(self, epoch: Optional[int] = None):
        # undo the previous step's scaling so the base scheduler advances from its own value
        for g, lr in zip(self.optimizer.param_groups, self.base_lrs if epoch == 0 else self._unscaled):
            g["lr"] = lr
        if epoch != 0:
            self.base_lr_scheduler.step()
        self._unscaled = [g["lr"] for g in self.optimizer.param_groups]
        self.last_epoch = self.last_epoch + 1 if epoch is None else epoch
        bs = self.batch_sizes[self.last_epoch % len(self.batch_sizes)]
        for g in self.optimizer.param_groups:
            g["lr"] = scale_lr(self.base_batch_size, bs, g["lr"], self.lr_scaling_method)
        if self.verbose:
            logger.info(f"next batch {self.last_epoch}: batch size {bs} (base {self.base_batch_size}), lr {self.get_lr()}")


def lr_scheduler_for_variable_batch_size(base_batch_size: int, batch_sizes: Sequence[int], dataloader,
                                         lr_scheduler_or_optimizer, lr_scaling_method: str = "linear", verbose: bool = False):
    if isinstance(lr_scheduler_or_optimizer, Optimizer):
        base = _ConstantLR(lr_scheduler_or_optimizer)
    else:
        assert hasattr(lr_scheduler_or_optimizer, "optimizer"), "expected an optimizer or an LR scheduler"
        base = lr_scheduler_or_optimizer
    return VariableBatchSizeLR(base, base_batch_size, batch_sizes, dataloader, lr_scaling_method, verbose=verbose)


def get_dataloader_and_lr_scheduler_for_variable_batch_size(dataset, dataset_seqlens, max_tokens: int, base_batch_size: int,
                                                            lr_scheduler_or_optimizer=None, effective_batch_size: int = 1,
                                                            dataloader_rank: int = 0, dataloader_num_replicas: int = 1,
                                                            dataloader_batch_size: int = 1, lr_scaling_method: str = "linear",
                                                            min_batch_size: int = 1, max_batch_size: Optional[int] = None,
                                                            sequence_picking_order: str = "dataloader",
                                                            required_microbatches_of_same_size: bool = False,
                                                            required_microbatches_of_same_seqlen: bool = False,
                                                            sample_padding_fn=None, dataloader_collate_fn=None,
                                                            dataloader_num_workers: int = 0, dataloader_pin_memory: bool = False,
                                                            seed: Optional[int] = None, verbose: bool = False):
    mb_ids, batch_sizes, max_seqlens = batch_by_seqlens(
        dataset_seqlens, max_tokens, min_batch_size=min_batch_size, max_batch_size=max_batch_size,
        sequence_picking_order=sequence_picking_order, effective_batch_size=effective_batch_size,
        required_microbatches_of_same_size=required_microbatches_of_same_size, verbose=verbose, seed=seed)
    loader, io_kwargs = dataloader_for_variable_batch_size(
        dataset, mb_ids, max_seqlens, dataloader_rank, dataloader_batch_size, dataloader_num_replicas, dataloader_collate_fn,
        dataloader_num_workers, dataloader_pin_memory, required_microbatches_of_same_seqlen, sample_padding_fn)
    sched = None if lr_scheduler_or_optimizer is None else lr_scheduler_for_variable_batch_size(
        base_batch_size, batch_sizes, loader, lr_scheduler_or_optimizer, lr_scaling_method, verbose)
    return loader, sched, io_kwargs


def get_dataloader_and_lr_scheduler_for_variable_batch_size_deepspeed(dataset, engine, dataset_seqlens=None,
                                                                      dataset_filter_ids=None, dataloader_collate_fn=None,
                                                                      sample_padding_fn=None, batch_seqlens_fn=None, **kw):
    """Engine glue: sizes come from the engine (dp world size x gradient accumulation = micro-batches per step) and
    the ``data_efficiency.data_sampling.dynamic_batching`` config section."""
    cfg = ((getattr(engine.config, "data_efficiency_config", None) or {}).get("data_sampling") or {}).get("dynamic_batching") or {}
    cfg = {**cfg, **kw}
    dp, gas = engine.dp_world_size, engine.gradient_accumulation_steps()
    loader, sched, io_kwargs = get_dataloader_and_lr_scheduler_for_variable_batch_size(
        dataset, dataset_seqlens, max_tokens=cfg["max_tokens"], base_batch_size=engine.train_batch_size(),
        lr_scheduler_or_optimizer=engine.lr_scheduler or engine.optimizer, effective_batch_size=dp * gas,
        dataloader_rank=getattr(engine, "dp_rank", engine.global_rank % dp), dataloader_num_replicas=dp,
        lr_scaling_method=cfg.get("lr_scaling_method", "linear"), min_batch_size=cfg.get("min_batch_size", 1),
        max_batch_size=cfg.get("max_batch_size"), sequence_picking_order=cfg.get("sequence_picking_order", "dataloader"),
        required_microbatches_of_same_size=engine.pipeline_parallelism if hasattr(engine, "pipeline_parallelism") else False,
        required_microbatches_of_same_seqlen=engine.pipeline_parallelism if hasattr(engine, "pipeline_parallelism") else False,
        sample_padding_fn=sample_padding_fn, dataloader_collate_fn=dataloader_collate_fn, verbose=cfg.get("verbose", False))
    return loader, sched, io_kwargs
