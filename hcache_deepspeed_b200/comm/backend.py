"""Communication back-end object model (``deepspeed/comm/backend.py:20-44``): the module-level functions of
:mod:`.comm` are the API; a ``Backend`` object is what ``cdb`` points at and what alternative transports subclass."""
from __future__ import annotations


class Backend:

    def __init__(self, name: str = "backend", rank: int = 0, size: int = 1):
        self.name = name
        self.world_group = None
        self.world_size = size
        self.world_rank = rank
        self.process_groups = []
        self.initialized = False

    def is_initialized(self) -> bool:
        return self.initialized

    def new_group(self):
        raise NotImplementedError

    def init_process_group(self):
        self.initialized = True
