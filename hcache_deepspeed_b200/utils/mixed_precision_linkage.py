"""Link low-precision (bf16 / fp16) parameters to the slice of the flat fp32 partition this rank owns
(``deepspeed/utils/mixed_precision_linkage.py:11-51``, ``tensor_fragment.py:312-337``).

A parameter occupying ``[lp_start, lp_start + numel)`` of its group's flat buffer overlaps the rank's partition
``[partition_start, partition_start + partition_size)`` in one contiguous fragment; ``_hp_mapping`` records where that
fragment sits in both and holds views of the fp32 master and (lazily) its optimizer state."""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import Dict, List, Optional, Tuple

import torch


@dataclass
class fragment_address:
    start: int
    numel: int


@dataclass
class tensor_fragment:
    lp_fragment: torch.Tensor
    lp_fragment_address: fragment_address
    hp_fragment: torch.Tensor
    hp_fragment_address: fragment_address
    param_group_index: int = 0
    optim_fragment: Dict[str, torch.Tensor] = field(default_factory=dict)

    def update_hp(self) -> None:
        self.hp_fragment.data.copy_(self.lp_fragment.data)

    def update_lp(self) -> None:
        self.lp_fragment.data.copy_(self.hp_fragment.data)

    def get_hp_fragment(self, optim_state_key: Optional[str] = None) -> torch.Tensor:
        return self.hp_fragment if optim_state_key is None else self.optim_fragment[optim_state_key]

    def set_optim_state_fragment(self, flat_hp_partition: torch.Tensor, optim_state: Dict[str, torch.Tensor]) -> None:
        a = self.hp_fragment_address
        self.optim_fragment = {k: v.narrow(0, a.start, a.numel) for k, v in optim_state.items()
                               if torch.is_tensor(v) and v.dim() > 0 and v.numel() == flat_hp_partition.numel()}


def _init_lp_to_hp_mapping(lp_param_list, partition_start: int, partition_size: int, dp_group) -> List[Tuple[torch.Tensor, int]]:
    out, offset, index = [], 0, 0
    end = partition_start + partition_size
    for p in lp_param_list:
        p._hp_mapping, p._dp_group = None, dp_group
        if offset < end and offset + p.numel() > partition_start:
            out.append((p, offset))
            p._index_in_param_group = index
            index += 1
        offset += p.numel()
    return out


def get_hp_fragment_mapping(lp_param, lp_start: int, flat_hp_partition: torch.Tensor, param_group_index: int,
                            partition_start: int, partition_size: int) -> tensor_fragment:
    start = max(lp_start, partition_start)
    end = min(lp_start + lp_param.numel(), partition_start + partition_size)
    assert start < end, f"fragment start {start} should be < fragment end {end}"
    hp = fragment_address(start - partition_start, end - start)
    lp = fragment_address(start - lp_start, end - start)
    return tensor_fragment(lp_param.view(-1).narrow(0, lp.start, lp.numel), lp, flat_hp_partition.narrow(0, hp.start, hp.numel),
                           hp, param_group_index)


def link_hp_params(lp_param_list, flat_hp_partition: torch.Tensor, gradient_dict=None, offload_gradient_dict=None,
                   use_offload: bool = False, param_group_index: int = 0, partition_start: int = 0,
                   partition_size: Optional[int] = None, dp_group=None) -> None:
    size = flat_hp_partition.numel() if partition_size is None else partition_size
    for p, lp_start in _init_lp_to_hp_mapping(lp_param_list, partition_start, size, dp_group):
        p._hp_mapping = get_hp_fragment_mapping(p, lp_start, flat_hp_partition, param_group_index, partition_start, size)


def lazy_init_hp_params_optimizer_state(lp_param_list, flat_hp_partition: torch.Tensor, optimizer_state) -> None:
    for p in lp_param_list:
        if getattr(p, "_hp_mapping", None) is not None:
            p._hp_mapping.set_optim_state_fragment(flat_hp_partition, optimizer_state[flat_hp_partition])
