"""DeepSpeedDataLoader (distributed sampler + micro-batch sizing) and RepeatingLoader.
Reference: ``deepspeed/runtime/dataloader.py:17-162``."""
from __future__ import annotations

import torch
from torch.utils.data import DataLoader, DistributedSampler, RandomSampler


class RepeatingLoader:
    """Wraps an iterator so it restarts when exhausted (pipeline engines pull a fixed number of micro-batches)."""

    def __init__(self, loader):
        self.loader = loader
        self.data_iter = iter(loader)

    def __iter__(self):
        return self

    def __next__(self):
        try:
            return next(self.data_iter)
        except StopIteration:
            self.data_iter = iter(self.loader)
            return next(self.data_iter)


class DeepSpeedDataLoader:

    def __init__(self, dataset, batch_size, pin_memory, local_rank, tput_timer=None, collate_fn=None, num_local_io_workers=None,
                 data_sampler=None, data_parallel_world_size=None, data_parallel_rank=None, dataloader_drop_last=False,
                 deepspeed_dataloader_config=None):
        self.tput_timer = tput_timer
        self.batch_size = batch_size
        self.curriculum_learning_enabled = bool((deepspeed_dataloader_config or {}).get("curriculum_learning_enabled"))
        if data_sampler is None:
            if local_rank >= 0 and (data_parallel_world_size or 1) > 1:
                data_sampler = DistributedSampler(dataset, num_replicas=data_parallel_world_size, rank=data_parallel_rank)
            elif local_rank >= 0:
                data_sampler = RandomSampler(dataset)
            else:
                data_sampler = RandomSampler(dataset)
                batch_size *= max(1, torch.cuda.device_count())
        self.num_local_io_workers = num_local_io_workers if num_local_io_workers is not None else 0
        self.data_sampler = data_sampler
        self.dataset = dataset
        self.collate_fn = collate_fn
        self.pin_memory = pin_memory
        self.dataloader_drop_last = dataloader_drop_last
        if self.curriculum_learning_enabled:
            self.len = len(self.data_sampler) // max(1, batch_size * (data_parallel_world_size or 1))
        else:
            self.len = len(self.data_sampler) // batch_size if dataloader_drop_last else \
                (len(self.data_sampler) + batch_size - 1) // batch_size
        self.batch_size = batch_size
        self.data = None

    def __iter__(self):
        self._create_dataloader()
        return self

    def __len__(self):
        return self.len

    def __next__(self):
        if self.tput_timer:
            self.tput_timer.start()
        return next(self.data)

    def _create_dataloader(self):
        if self.curriculum_learning_enabled:
            # the data-efficiency sampler yields whole micro-batches of indices (ordered by the curriculum)
            kw = dict(pin_memory=self.pin_memory and torch.cuda.is_available(), batch_sampler=self.data_sampler,
                      num_workers=self.num_local_io_workers)
        else:
            kw = dict(batch_size=self.batch_size, pin_memory=self.pin_memory and torch.cuda.is_available(),
                      sampler=self.data_sampler, num_workers=self.num_local_io_workers, drop_last=self.dataloader_drop_last)
        if self.collate_fn is not None:
            kw["collate_fn"] = self.collate_fn
        self.dataloader = DataLoader(self.dataset, **kw)
        self.data = iter(self.dataloader)
        return self.dataloader
