"""Cartesian process topologies for pipeline / data / model parallelism and the grid of process groups built on them.
Reference: ``deepspeed/runtime/pipe/topology.py`` (ProcessTopology :12, PipeDataParallelTopology :235,
PipeModelDataParallelTopology :247, PipelineParallelGrid :254)."""
from __future__ import annotations

from collections import namedtuple
from itertools import product
from typing import Dict, List, Optional

from ... import comm as dist


class ProcessTopology:
    """Maps n-dimensional coordinates (named axes, row-major: last axis fastest) to linear ranks."""

    def __init__(self, axes: List[str], dims: List[int]):
        self.axes = list(axes)
        self.dims = list(dims)
        self.ProcessCoord = namedtuple("ProcessCoord", axes)
        self.mapping: Dict[tuple, int] = {}
        for rank, coord in enumerate(product(*[range(d) for d in dims])):
            self.mapping[self.ProcessCoord(*coord)] = rank

    def get_rank(self, **coord_kwargs) -> int:
        if len(coord_kwargs) != len(self.axes):
            raise ValueError("get_rank() does not support slices. Use filter_match())")
        key = self.ProcessCoord(**coord_kwargs)
        assert key in self.mapping, f"key {coord_kwargs} invalid"
        return self.mapping[key]

    def get_axis_names(self) -> List[str]:
        return self.axes

    def get_rank_repr(self, rank: int, omit_axes=("data", "pipe"), inner_sep="_", outer_sep="-") -> str:
        omit = frozenset(omit_axes)
        coord = self.get_coord(rank)
        return outer_sep.join(f"{ax}{inner_sep}{getattr(coord, ax):02d}" for ax in self.axes if ax not in omit)

    def get_dim(self, axis: str) -> int:
        return self.dims[self.axes.index(axis)] if axis in self.axes else 0

    def get_coord(self, rank: int):
        for coord, r in self.mapping.items():
            if r == rank:
                return coord
        raise ValueError(f"rank {rank} not found in topology.")

    def get_axis_comm_lists(self, axis: str) -> List[List[int]]:
        """Rank lists that vary only along ``axis`` (one communicator per combination of the other axes)."""
        if axis not in self.axes:
            return []
        others = [a for a in self.axes if a != axis]
        lists = []
        for combo in product(*[range(self.get_dim(a)) for a in others]):
            fixed = dict(zip(others, combo))
            lists.append([self.get_rank(**{axis: i}, **fixed) for i in range(self.get_dim(axis))])
        return lists

    def filter_match(self, **filter_kwargs) -> List[int]:
        return [r for c, r in self.mapping.items() if all(getattr(c, k) == v for k, v in filter_kwargs.items())]

    def get_axis_list(self, axis: str, idx: int) -> List[int]:
        return sorted(r for c, r in self.mapping.items() if getattr(c, axis) == idx)

    def world_size(self) -> int:
        return len(self.mapping)

    def __str__(self):
        return str(self.mapping)


class PipeDataParallelTopology(ProcessTopology):
    """pipe x data, data fastest: adjacent ranks do the (bandwidth-heavy) gradient all-reduce."""

    def __init__(self, num_pp: int, num_dp: int):
        super().__init__(axes=["pipe", "data"], dims=[num_pp, num_dp])


class PipeModelDataParallelTopology(ProcessTopology):
    def __init__(self, num_pp: int, num_mp: int, num_dp: int):
        super().__init__(axes=["pipe", "data", "model"], dims=[num_pp, num_dp, num_mp])


class PipelineParallelGrid:
    """Process groups of a pipe x data (x model) topology; doubles as the ``mpu`` object engines expect."""

    def __init__(self, topology: Optional[ProcessTopology] = None, process_group=None):
        self.global_rank = dist.get_rank()
        self.world_size = dist.get_world_size()
        if topology is None:
            num_pp = 1
            for f in range(2, int(self.world_size ** 0.5) + 1):  # largest factor <= sqrt(world) as pipeline depth
                if self.world_size % f == 0:
                    num_pp = f
            topology = PipeDataParallelTopology(num_pp=num_pp, num_dp=self.world_size // num_pp)
        self._topo = topology
        self.data_parallel_size = max(self._topo.get_dim("data"), 1)
        self.pipe_parallel_size = max(self._topo.get_dim("pipe"), 1)
        self.model_parallel_size = max(self._topo.get_dim("model"), 1)
        self.slice_parallel_size = self.model_parallel_size
        assert self.world_size == self.data_parallel_size * self.pipe_parallel_size * self.model_parallel_size
        coord = self._topo.get_coord(self.global_rank)
        self.stage_id = getattr(coord, "pipe", 0)
        self.data_parallel_id = getattr(coord, "data", 0)

        def make(axis):
            mine, mine_ranks = None, None
            for ranks in self._topo.get_axis_comm_lists(axis):
                g = dist.new_group(ranks=ranks)
                if self.global_rank in ranks:
                    mine, mine_ranks = g, ranks
            return mine, mine_ranks

        self.dp_proc_group, self.dp_group = make("data")
        if self.dp_proc_group is None:
            self.dp_group = [self.global_rank]
        self.pp_proc_group, self.pp_group = make("pipe")
        if self.pp_group is None:
            self.pp_group = [self.global_rank]
        self.is_first_stage = self.stage_id == 0
        self.is_last_stage = self.stage_id == self.pipe_parallel_size - 1
        if "model" in self._topo.get_axis_names():
            self.slice_proc_group, self.slice_group = make("model")
        else:
            self.slice_proc_group, self.slice_group = None, [self.global_rank]
        # "model parallel" for gradient-norm purposes = everything that is not data parallel (pipe x model)
        self.ds_model_proc_group, self.ds_model_rank, self.ds_model_world_size = None, -1, -1
        for dp in range(self.data_parallel_size):
            ranks = sorted(self._topo.get_axis_list(axis="data", idx=dp))
            g = dist.new_group(ranks=ranks)
            if self.global_rank in ranks:
                self.ds_model_proc_group, self.ds_model_world_size = g, len(ranks)
                self.ds_model_rank = ranks.index(self.global_rank)
        self.p2p_groups = self._build_p2p_groups()

    def _build_p2p_groups(self):
        pairs = []
        for ranks in self._topo.get_axis_comm_lists("pipe"):
            for i, r in enumerate(ranks):
                pairs.append([r, ranks[(i + 1) % len(ranks)]])
        return pairs

    def get_stage_id(self):
        return self.stage_id

    def get_data_parallel_id(self):
        return self.data_parallel_id

    def stage_to_global(self, stage_id: int, **kwargs) -> int:
        me = self._topo.get_coord(self.global_rank)
        coord = me._replace(pipe=stage_id, **kwargs)
        return self._topo.get_rank(**coord._asdict())

    def topology(self):
        return self._topo

    # ---- mpu interface -------------------------------------------------------------------------------
    def get_global_rank(self):
        return self.global_rank

    def get_pipe_parallel_rank(self):
        return self.stage_id

    def get_pipe_parallel_world_size(self):
        return self.pipe_parallel_size

    def get_pipe_parallel_group(self):
        return self.pp_proc_group

    def get_data_parallel_rank(self):
        return self.data_parallel_id

    def get_data_parallel_world_size(self):
        return self.data_parallel_size

    def get_data_parallel_group(self):
        return self.dp_proc_group

    def get_model_parallel_rank(self):
        return self.ds_model_rank

    def get_model_parallel_world_size(self):
        return self.ds_model_world_size

    def get_model_parallel_group(self):
        return self.ds_model_proc_group

    def get_slice_parallel_rank(self):
        return getattr(self._topo.get_coord(self.global_rank), "model", 0)

    def get_slice_parallel_world_size(self):
        return self.slice_parallel_size

    def get_slice_parallel_group(self):
        return self.slice_proc_group
