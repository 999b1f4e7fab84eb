"""Contracting a (pipeline, tensor, data)-parallel checkpoint layout onto fewer ranks: which source files each target
rank has to read and how their optimizer states merge.

Reference: ``deepspeed/checkpoint/reshape_meg_2d.py:11-100``, ``reshape_3d_utils.py:16-111``,
``reshape_utils.py:13-113``, ``zero_checkpoint.py:16-142``, ``utils.py:13-67``. The reference keeps the maps as
dicts of ``"pp,tp"`` strings built by repeated list partitioning; here a layout is an index array ``[pp, tp]`` and a
contraction is a block view of it, which yields the same file order.
"""
from __future__ import annotations

import os
import re
from collections import OrderedDict
from typing import Dict, List, Optional, Sequence, Tuple

import numpy as np
import torch

from .constants import (BASE_OPTIMIZER_STATE, BF16_ZERO_FILE_PREFIX, FP16_ZERO_FILE_PREFIX, GROUP_PADDINGS, LAYER_FILE_PREFIX,
                        MODEL_FILE_PREFIX, MODEL_FILE_SUFFIX, OPTIM_FILE_SUFFIX, OPTIMIZER_STATE_DICT, PARTITION_COUNT,
                        ZERO_FILE_PREFIX)

PP_DIM, TP_DIM, DP_DIM = "PP", "TP", "DP"


# ---- files ---------------------------------------------------------------------------------------------------
def basic_folder_validation(dir_: str) -> None:
    assert os.path.exists(dir_), f"{dir_} path does not exist"
    assert os.path.isdir(dir_), f"{dir_} is not a folder"


def get_files(dir_: str) -> List[str]:
    return [os.path.join(root, f) for root, _, files in os.walk(dir_) for f in files]


def get_files_with_prefix(all_files: Sequence[str], prefix: str) -> List[str]:
    return sorted(f for f in all_files if os.path.basename(f).startswith(prefix))


def validate_files(file_list: Sequence[str]) -> None:
    for f in file_list:
        if not os.path.isfile(f):
            print(f"Error: {f} is not existent")


def sort_zero_files(files: Sequence[str], prefix: str) -> List[str]:
    """Order ``<prefix><dp>_mp_rank_<mp>...`` files by (dp rank, mp rank) numerically."""
    pat = re.compile(f"{re.escape(prefix)}([0-9]+)_{re.escape(MODEL_FILE_PREFIX)}([0-9]+)")
    keyed = []
    for f in files:
        m = pat.search(f)
        if m is None:
            raise ValueError(f"Cannot parse dp_rank and mp_rank from {f}")
        keyed.append((int(m.group(1)), int(m.group(2)), f))
    return [f for _, _, f in sorted(keyed)]


def get_zero_files(dir_: str) -> List[str]:
    files = get_files(dir_)
    for prefix in (ZERO_FILE_PREFIX, FP16_ZERO_FILE_PREFIX, BF16_ZERO_FILE_PREFIX):
        found = get_files_with_prefix(files, prefix)
        if found:
            return sort_zero_files(found, prefix)
    return []


def partition_data(data_list: Sequence, num_partitions: int) -> List[list]:
    n = len(data_list)
    assert n % num_partitions == 0, f"{n} items do not split into {num_partitions} partitions"
    size = n // num_partitions
    return [list(data_list[i:i + size]) for i in range(0, n, size)]


def get_model_ckpt_name_for_rank(base_folder: str, mp_rank_str: str) -> str:
    return os.path.join(base_folder, MODEL_FILE_PREFIX + mp_rank_str + MODEL_FILE_SUFFIX)


def get_zero_ckpt_name_for_rank(base_folder: str, dp_rank: int, mp_rank: int) -> str:
    return os.path.join(base_folder, f"{ZERO_FILE_PREFIX}{dp_rank}_{MODEL_FILE_PREFIX}{mp_rank:02d}{OPTIM_FILE_SUFFIX}")


def get_layer_ckpt_name_for_rank(base_folder: str, layer_id: str, tp_rank: int) -> str:
    return os.path.join(base_folder, f"{layer_id}-model_{tp_rank:02d}{MODEL_FILE_SUFFIX}")


def clone_tensors_for_torch_save(item, device=torch.device("cpu")):
    """Deep copy with every tensor detached and cloned onto ``device`` (so ``torch.save`` does not drag whole flat
    buffers along with the views into them)."""
    if torch.is_tensor(item):
        device = torch.device(device)
        return item.detach().clone() if item.device == device else item.detach().to(device)
    if isinstance(item, (list, tuple)):
        return type(item)(clone_tensors_for_torch_save(v, device) for v in item)
    if isinstance(item, dict):
        return type(item)({k: clone_tensors_for_torch_save(v, device) for k, v in item.items()})
    return item


# ---- state merging -------------------------------------------------------------------------------------------
def merge_state(a, b, key_list: Optional[List[str]] = None):
    """Merge two rank states: tensors concatenate along dim 0, containers merge element-wise, scalars keep ``a``."""
    keys = key_list or []
    if type(a) is not type(b):
        raise ValueError(f"Cannot merge states of types {type(a)} and {type(b)} at {'.'.join(keys)}")
    if isinstance(a, (dict, OrderedDict)):
        out = type(a)()
        for k, v in b.items():
            out[k] = merge_state(a[k], v, [str(k)]) if k in a else v
        return out
    if isinstance(a, (list, tuple)):
        if len(a) != len(b):
            raise ValueError(f"Cannot merge lists of different lengths ({len(a)} vs {len(b)}) at {'.'.join(keys)}")
        return type(a)(merge_state(x, y, keys) for x, y in zip(a, b))
    if torch.is_tensor(a):
        return torch.cat([a, b], 0)
    return a


merge_state_dict = merge_state
merge_state_list = merge_state


# ---- 2D (pipeline x tensor) layout -----------------------------------------------------------------------------
class meg_2d_parallel_map:
    """``(pp_index, tp_index) -> [source rank / file indices]``."""

    def __init__(self, pp_degree: int, tp_degree: int):
        self.pp_degree, self.tp_degree = pp_degree, tp_degree
        self.map: Dict[Tuple[int, int], list] = {}

    def simple_init(self) -> None:
        self.map = {(r // self.tp_degree, r % self.tp_degree): [r] for r in range(self.pp_degree * self.tp_degree)}

    def _check(self, pp_index, tp_index) -> None:
        assert pp_index is None or pp_index < self.pp_degree
        assert tp_index is None or tp_index < self.tp_degree

    def add_data(self, pp_index: int, tp_index: int, data: list) -> None:
        self._check(pp_index, tp_index)
        assert type(data) is list
        self.map.setdefault((pp_index, tp_index), []).extend(data)

    def get_data(self, pp_index: Optional[int] = None, tp_index: Optional[int] = None) -> list:
        self._check(pp_index, tp_index)
        pps = range(self.pp_degree) if pp_index is None else [pp_index]
        tps = range(self.tp_degree) if tp_index is None else [tp_index]
        return [x for i in pps for j in tps for x in self.map[(i, j)]]

    def print_data(self, tag: str) -> None:
        print(tag)
        for (i, j), v in self.map.items():
            print(f"{i},{j} = {v}")


def reshape_meg_2d_parallel(old_pp_degree: int, old_tp_degree: int, new_pp_degree: int, new_tp_degree: int,
                            verbose: bool = False) -> meg_2d_parallel_map:
    """Target cell (i, j) reads the block of source ranks ``[i*rp, (i+1)*rp) x [j*rt, (j+1)*rt)``, pipeline-major."""
    assert new_pp_degree <= old_pp_degree and new_tp_degree <= old_tp_degree
    assert old_pp_degree % new_pp_degree == 0 and old_tp_degree % new_tp_degree == 0
    grid = np.arange(old_pp_degree * old_tp_degree).reshape(old_pp_degree, old_tp_degree)
    rp, rt = old_pp_degree // new_pp_degree, old_tp_degree // new_tp_degree
    out = meg_2d_parallel_map(new_pp_degree, new_tp_degree)
    for i in range(new_pp_degree):
        for j in range(new_tp_degree):
            out.add_data(i, j, grid[i * rp:(i + 1) * rp, j * rt:(j + 1) * rt].reshape(-1).tolist())
    if verbose:
        out.print_data("final_2d_map:")
    return out


def get_mpu_ranks(tp_size: int = 1, pp_size: int = 1, dp_size: int = 1, virtual_pp_size=None):
    """Megatron rank layout (tp fastest, then dp, then pp): returns (tp groups, pp groups, dp groups)."""
    world = tp_size * pp_size * dp_size
    ranks = np.arange(world).reshape(pp_size, dp_size, tp_size)
    tp_groups = ranks.reshape(-1, tp_size).tolist()
    dp_groups = ranks.transpose(0, 2, 1).reshape(-1, dp_size).tolist()
    pp_groups = ranks.reshape(pp_size, -1).T.tolist()
    return tp_groups, pp_groups, dp_groups


# ---- 3D layout ---------------------------------------------------------------------------------------------------
def flatten_dp_dimension(meg_2d_map: meg_2d_parallel_map, src_2d_size: int, dp_degree: int) -> meg_2d_parallel_map:
    out = meg_2d_parallel_map(meg_2d_map.pp_degree, meg_2d_map.tp_degree)
    for (i, j), cell in meg_2d_map.map.items():
        out.add_data(i, j, [idx + k * src_2d_size for idx in cell for k in range(dp_degree)])
    return out


def unflatten_dp_dimension(meg_2d_map: meg_2d_parallel_map, dp_degree: int) -> List[meg_2d_parallel_map]:
    maps = [meg_2d_parallel_map(meg_2d_map.pp_degree, meg_2d_map.tp_degree) for _ in range(dp_degree)]
    for (i, j), cell in meg_2d_map.map.items():
        for m, part in zip(maps, partition_data(cell, dp_degree)):
            m.add_data(i, j, part)
    return maps


class model_3d_desc:

    def __init__(self, pp_degree: int = 1, tp_degree: int = 1, dp_degree: int = 1):
        self.pp_degree, self.tp_degree, self.dp_degree = pp_degree, tp_degree, dp_degree

    def world_size(self) -> int:
        return self.pp_degree * self.tp_degree * self.dp_degree

    def get_desc(self) -> str:
        return f"{PP_DIM},{TP_DIM},{DP_DIM} = ({self.pp_degree}, {self.tp_degree}, {self.dp_degree})"

    def is_valid(self, pp_index: int, tp_index: int, dp_index: int):
        errs = [f"{name} indexing error: index {idx} >= degree {deg}"
                for idx, deg, name in ((pp_index, self.pp_degree, PP_DIM), (tp_index, self.tp_degree, TP_DIM),
                                       (dp_index, self.dp_degree, DP_DIM)) if idx >= deg]
        return not errs, errs

    def can_reshape(self, target: "model_3d_desc"):
        errs = [f"Expansion reshape not supported - {name}: {src} ---> {tgt}"
                for src, tgt, name in ((self.pp_degree, target.pp_degree, PP_DIM), (self.tp_degree, target.tp_degree, TP_DIM),
                                       (self.dp_degree, target.dp_degree, DP_DIM)) if tgt > src]
        return not errs, errs

    def reshape(self, target: "model_3d_desc", verbose: bool = False) -> List[meg_2d_parallel_map]:
        ok, errs = self.can_reshape(target)
        assert ok, ",".join(errs)
        m2d = reshape_meg_2d_parallel(self.pp_degree, self.tp_degree, target.pp_degree, target.tp_degree, verbose)
        flat = flatten_dp_dimension(m2d, self.pp_degree * self.tp_degree, self.dp_degree)
        return unflatten_dp_dimension(flat, target.dp_degree)


def get_model_3d_descriptor(dir_: str) -> model_3d_desc:
    """Infer (pp, tp, dp) of a checkpoint folder from its file names."""
    files, zero_files = get_files(dir_), get_zero_files(dir_)
    n_layer01 = len(get_files_with_prefix(files, f"{LAYER_FILE_PREFIX}01"))
    n_model = len(get_files_with_prefix(files, MODEL_FILE_PREFIX))
    if n_layer01 > 0:
        tp, pp = n_layer01, n_model // n_layer01
    else:
        tp, pp = n_model, 1
    return model_3d_desc(pp, tp, max(1, len(zero_files) // (pp * tp)))


# ---- ZeRO optimizer-state files of a checkpoint --------------------------------------------------------------
GROUP_STATE_KEY = "state"


class ZeROCheckpoint:

    def __init__(self, dir_: str):
        basic_folder_validation(dir_)
        self.dir = dir_
        self.file_list = get_zero_files(dir_)
        self.num_files = len(self.file_list)
        assert self.num_files > 0, f"No ZeRO files found in {dir_}"
        self.src_3d = get_model_3d_descriptor(dir_)
        self.reshape(model_3d_desc(self.src_3d.pp_degree, self.src_3d.tp_degree, self.src_3d.dp_degree))

    def reshape(self, target_3d_desc: model_3d_desc) -> None:
        self.target_3d = target_3d_desc
        self._3d_file_map = self.src_3d.reshape(target_3d_desc)

    def get_src_world_size(self) -> int:
        return self.src_3d.world_size()

    def get_src_tp_degree(self) -> int:
        return self.src_3d.tp_degree

    def get_src_pp_degree(self) -> int:
        return self.src_3d.pp_degree

    def get_src_dp_degree(self) -> int:
        return self.src_3d.dp_degree

    def get_file_indices_for_rank(self, pp_index: int, tp_index: int, dp_index: int) -> List[int]:
        assert dp_index < len(self._3d_file_map), f"DP index {dp_index} >= DP degree {len(self._3d_file_map)}"
        return self._3d_file_map[dp_index].get_data(pp_index, tp_index)

    def get_files_for_rank(self, pp_index: int, tp_index: int, dp_index: int) -> List[str]:
        return [self.file_list[i] for i in self.get_file_indices_for_rank(pp_index, tp_index, dp_index)]

    def get_state_for_rank(self, pp_index: int, tp_index: int, dp_index: int, keys_to_ignore: Sequence[str] = (),
                           strip_tensor_paddings: bool = True):
        merged = None
        for path in self.get_files_for_rank(pp_index, tp_index, dp_index):
            sd = torch.load(path, map_location="cpu", weights_only=False)
            for k in keys_to_ignore:
                sd.pop(k, None)
            if strip_tensor_paddings:
                self._strip_tensor_paddings(sd)
            merged = sd if merged is None else merge_state(merged, sd)
            opt = merged.get(OPTIMIZER_STATE_DICT) or {}
            if opt.get(PARTITION_COUNT):
                opt[PARTITION_COUNT] = [self.target_3d.dp_degree] * len(opt[PARTITION_COUNT])
            if strip_tensor_paddings and opt.get(GROUP_PADDINGS):
                opt[GROUP_PADDINGS] = [0] * len(opt[GROUP_PADDINGS])
        return merged

    def print_3d_index_map(self, tag: Optional[str] = None) -> None:
        if tag:
            print(f"3D index map: {tag}")
        for dp, m in enumerate(self._3d_file_map):
            m.print_data(f"dp = {dp}")

    def print_3d_file_map(self, tag: Optional[str] = None) -> None:
        if tag:
            print(f"3D file map: {tag}")
        for dp, m in enumerate(self._3d_file_map):
            for (pp, tp), idx in m.map.items():
                print(f"{pp}, {tp}, {dp} => {[self.file_list[i] for i in idx]}")

    @staticmethod
    def _strip_tensor_paddings(sd) -> None:
        """Drop the alignment padding at the tail of every flat per-group optimizer state tensor."""
        opt = sd.get(OPTIMIZER_STATE_DICT) or {}
        states = (opt.get(BASE_OPTIMIZER_STATE) or {}).get(GROUP_STATE_KEY)
        pads = opt.get(GROUP_PADDINGS)
        if states is None or pads is None:
            return
        for key, group_state in states.items():
            if pads[key] == 0:
                continue
            for name, value in group_state.items():
                if name != "step" and torch.is_tensor(value):
                    group_state[name] = value.narrow(0, 0, value.numel() - pads[key]).clone()
