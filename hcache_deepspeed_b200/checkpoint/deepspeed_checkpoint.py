"""Inspect a checkpoint directory: discover the tensor / pipeline / data parallel degrees from the file names and give
access to per-rank files. Reference: ``deepspeed/checkpoint/deepspeed_checkpoint.py:38-330`` (plus ``reshape_utils``)."""
from __future__ import annotations

import os
import re
from typing import Dict, List

import torch

from .constants import (BF16_ZERO_FILE_PREFIX, LAYER_FILE_PREFIX, MODEL_FILE_PREFIX, MODEL_FILE_SUFFIX, OPTIM_FILE_SUFFIX,
                        ZERO_FILE_PREFIX)


def get_files_with_prefix(all_files: List[str], prefix: str) -> List[str]:
    return sorted(f for f in all_files if os.path.basename(f).startswith(prefix))


def get_files(dir_: str) -> List[str]:
    out = []
    for root, _, files in os.walk(dir_):
        out += [os.path.join(root, f) for f in files]
    return sorted(out)


def partition_data(data_list, num_partitions: int):
    assert len(data_list) % num_partitions == 0, f"{len(data_list)} items cannot be split into {num_partitions} partitions"
    n = len(data_list) // num_partitions
    return [data_list[i:i + n] for i in range(0, len(data_list), n)]


class DeepSpeedCheckpoint:

    def __init__(self, dir: str, tp_degree: int = None, pp_degree: int = None, dp_degree: int = None, final_layer_norm_idx=None):
        self.dir = dir
        files = get_files(dir)
        self.file_list = files
        self.layer_files = get_files_with_prefix(files, LAYER_FILE_PREFIX)
        self.mp_rank_files = get_files_with_prefix(files, MODEL_FILE_PREFIX)
        self.zero_files = [f for f in files if os.path.basename(f).endswith(OPTIM_FILE_SUFFIX)]
        self.original_tp_degree = max(1, len({re.search(r"mp_rank_(\d+)", os.path.basename(f)).group(1)
                                              for f in self.mp_rank_files + self.zero_files if re.search(r"mp_rank_(\d+)", os.path.basename(f))}))
        dp_ids = {int(m.group(1)) for f in self.zero_files for m in [re.search(r"zero_pp_rank_(\d+)", os.path.basename(f))] if m}
        self.original_dp_degree = max(1, len(dp_ids))
        layer_ids = sorted({re.search(r"layer_(\d+)", os.path.basename(f)).group(1) for f in self.layer_files})
        self.layer_keys = layer_ids
        self.original_pp_degree = max(1, len(self.mp_rank_files) // self.original_tp_degree) if self.layer_files else 1
        self.tp_degree = self.original_tp_degree if tp_degree is None else tp_degree
        self.pp_degree = self.original_pp_degree if pp_degree is None else pp_degree
        self.dp_degree = self.original_dp_degree if dp_degree is None else dp_degree
        self.original_world_size = self.original_tp_degree * self.original_pp_degree * self.original_dp_degree
        self.world_size = self.tp_degree * self.pp_degree * self.dp_degree
        self.global_state = {}

    def is_change_tp_degree(self):
        return self.tp_degree != self.original_tp_degree

    def is_change_pp_degree(self):
        return self.pp_degree != self.original_pp_degree

    def is_change_dp_degree(self):
        return self.dp_degree != self.original_dp_degree

    def show_file_map(self):
        for f in self.file_list:
            print(os.path.relpath(f, self.dir))

    def get_zero_checkpoint_state(self, pp_index: int = 0, tp_index: int = 0, dp_index: int = 0):
        for f in self.zero_files:
            b = os.path.basename(f)
            if f"zero_pp_rank_{dp_index}_" in b and f"mp_rank_{tp_index:02d}" in b:
                return torch.load(f, map_location="cpu", weights_only=False)
        raise FileNotFoundError(f"no zero checkpoint for dp={dp_index} tp={tp_index}")

    def get_zero_files(self, pp_index: int = 0, tp_index: int = 0, dp_index: int = 0) -> List[str]:
        return [f for f in self.zero_files if f"zero_pp_rank_{dp_index}_" in os.path.basename(f)
                and f"mp_rank_{tp_index:02d}" in os.path.basename(f)]

    def get_model_state(self, tp_index: int = 0, pp_index: int = 0):
        idx = pp_index * self.original_tp_degree + tp_index
        return torch.load(self.mp_rank_files[idx], map_location="cpu", weights_only=False)

    def get_iteration(self):
        sd = self.get_model_state()
        return sd.get("iteration", sd.get("global_steps", 0))

    def get_args(self):
        return self.get_model_state().get("args")

    def get_layer_state(self, layer_key: str, tp_index: int = 0):
        files = [f for f in self.layer_files if f"layer_{layer_key}" in os.path.basename(f)]
        return torch.load(files[min(tp_index, len(files) - 1)], map_location="cpu", weights_only=False)

    def validate_files(self):
        for f in self.file_list:
            if not os.path.isfile(f):
                raise FileNotFoundError(f"{f} is not existent")


# ---- file names of one rank's pieces of a checkpoint (ref: checkpoint/utils.py:11-37) ---------------------------------
def get_model_ckpt_name_for_rank(base_folder: str, mp_rank_str: str) -> str:
    return os.path.join(base_folder, f"mp_rank_{mp_rank_str}_model_states.pt")


def get_zero_ckpt_name_for_rank(base_folder: str, dp_rank: int, mp_rank: int) -> str:
    return os.path.join(base_folder, f"zero_pp_rank_{dp_rank}_mp_rank_{mp_rank:02d}_optim_states.pt")


def get_layer_ckpt_name_for_rank(base_folder: str, layer_id: str, tp_rank: int) -> str:
    return os.path.join(base_folder, f"{layer_id}-model_{tp_rank:02d}-model_states.pt")
