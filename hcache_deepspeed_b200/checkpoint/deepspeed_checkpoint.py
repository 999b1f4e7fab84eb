"""Inspect a checkpoint directory: discover the tensor / pipeline / data parallel degrees from the file names and give
access to per-rank files. Reference: ``deepspeed/checkpoint/deepspeed_checkpoint.py:38-330`` (plus ``reshape_utils``)."""
from __future__ import annotations

import os
import re
from typing import Dict, List

import torch

from .reshape import meg_2d_parallel_map, merge_state, reshape_meg_2d_parallel
from .constants import (BF16_ZERO_FILE_PREFIX, LAYER_FILE_PREFIX, MODEL_FILE_PREFIX, MODEL_FILE_SUFFIX, OPTIM_FILE_SUFFIX,
                        ZERO_FILE_PREFIX)


EMBEDDING_LAYER_INDEX = 0
FINAL_LAYER_NORM_INDEX = -1
CHECKPOINT_INFO_KEY = "checkpoint_info"
# Megatron parameter names that are replicated over tensor-parallel ranks / joined along dim 1 when slices are merged
SEQUENTIAL_LAYERS = ["input_layernorm.weight", "input_layernorm.bias", "self_attention.dense.bias",
                     "post_attention_layernorm.weight", "post_attention_layernorm.bias", "mlp.dense_4h_to_h.bias",
                     "position_embeddings.weight"]
LAYER_CONCAT_DIM = {"self_attention.dense.weight": 1, "mlp.dense_4h_to_h.weight": 1}


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
        if self.layer_files:
            # pipeline checkpoints number their mp_rank files over (pipe x tensor): the tensor degree is the number of
            # slices one layer was written in
            first = f"{LAYER_FILE_PREFIX}{min(layer_ids, key=int)}-"
            self.original_tp_degree = max(1, len(get_files_with_prefix(self.layer_files, first)))
        self.original_pp_degree = max(1, len(self.mp_rank_files) // self.original_tp_degree) if self.layer_files else 1
        self.tp_degree = self.original_tp_degree if tp_degree is None else tp_degree
        self.pp_degree = self.original_pp_degree if pp_degree is None else pp_degree
        self.dp_degree = self.original_dp_degree if dp_degree is None else dp_degree
        self.original_world_size = self.original_tp_degree * self.original_pp_degree * self.original_dp_degree
        self.world_size = self.tp_degree * self.pp_degree * self.dp_degree
        self.global_state = {}
        # ---- Megatron-DeepSpeed layout: layer 0 = embedding, last layer = final norm, transformer blocks in between ----
        self.final_layer_norm_idx = FINAL_LAYER_NORM_INDEX if final_layer_norm_idx is None else final_layer_norm_idx
        self.layer_file_keys = [f"{LAYER_FILE_PREFIX}{k}" for k in sorted(layer_ids, key=int)]
        self.old_2d_map = meg_2d_parallel_map(self.original_pp_degree, self.original_tp_degree)
        self.old_2d_map.simple_init()
        self.new_2d_map = reshape_meg_2d_parallel(self.original_pp_degree, self.original_tp_degree, self.pp_degree, self.tp_degree)
        if self.layer_file_keys:
            self.tp_to_embedding_map = self._tp_map_of_layer(EMBEDDING_LAYER_INDEX)
            self.tp_to_final_norm_map = self._tp_map_of_layer(self.final_layer_norm_idx)
            self.pp_to_transformer_map, self.transformer_file_map = self._transformer_maps()
        else:
            self.tp_to_embedding_map, self.tp_to_final_norm_map = {}, {}
            self.pp_to_transformer_map, self.transformer_file_map = {}, {}

    # ---- layer-file maps under the (possibly new) tp / pp degrees ----------------------------------------------------------
    def _files_of_layer(self, key: str) -> List[str]:
        return get_files_with_prefix(self.layer_files, key + "-")

    def _tp_map_of_layer(self, layer_index: int) -> Dict[int, List[str]]:
        """new tp rank -> the old tp slices of that layer it has to merge."""
        return dict(enumerate(partition_data(self._files_of_layer(self.layer_file_keys[layer_index]), self.tp_degree)))

    def _transformer_layer_keys(self) -> List[str]:
        end = self.final_layer_norm_idx if self.final_layer_norm_idx != 0 else None
        return self.layer_file_keys[1:end]

    def _transformer_maps(self):
        keys = self._transformer_layer_keys()
        per_stage = max(1, len(keys) // max(1, self.pp_degree))
        pp_map = {pp: keys[pp * per_stage:(pp + 1) * per_stage] for pp in range(self.pp_degree)}
        file_map: Dict[tuple, List[List[str]]] = {}
        for i, key in enumerate(keys):
            pp = min(i // per_stage, self.pp_degree - 1)
            for tp, files in enumerate(partition_data(self._files_of_layer(key), self.tp_degree)):
                file_map.setdefault((tp, pp), []).append(files)
        return pp_map, file_map

    @staticmethod
    def _load(path: str):
        return torch.load(path, map_location="cpu", weights_only=False)

    @staticmethod
    def _merge_state_dicts(sd_list: List[dict]) -> dict:
        """Merge tp slices of one layer: replicated tensors (norms, row-parallel biases, position embeddings) are taken
        once, row-parallel weights are joined along dim 1, everything else along dim 0."""
        if len(sd_list) == 1:
            return sd_list[0]
        out = {}
        for key, first in sd_list[0].items():
            if key in SEQUENTIAL_LAYERS or not torch.is_tensor(first):
                out[key] = first
            else:
                out[key] = torch.cat([sd[key] for sd in sd_list], dim=LAYER_CONCAT_DIM.get(key, 0))
        return out

    def get_embedding_layer_id(self) -> str:
        return self.layer_file_keys[EMBEDDING_LAYER_INDEX]

    def get_final_norm_layer_id(self) -> str:
        return self.layer_file_keys[self.final_layer_norm_idx]

    def get_embedding_files(self, tp_index: int) -> List[str]:
        assert tp_index in self.tp_to_embedding_map
        return self.tp_to_embedding_map[tp_index]

    def get_embedding_state(self, tp_index: int) -> Dict:
        return self._merge_state_dicts([self._load(f) for f in self.get_embedding_files(tp_index)])

    def get_final_norm_files(self, tp_index: int) -> List[str]:
        assert tp_index in self.tp_to_final_norm_map
        return self.tp_to_final_norm_map[tp_index]

    def get_final_norm_state(self, tp_index: int) -> Dict:
        return self._load(self.get_final_norm_files(tp_index)[0])

    def get_pp_transformer_map(self, pp_index: int) -> List[str]:
        assert pp_index < self.pp_degree
        return self.pp_to_transformer_map[pp_index]

    def get_transformer_state(self, tp_index: int, pp_index: int) -> List[dict]:
        """One merged state dict per transformer layer of stage ``pp_index``, for new tp rank ``tp_index``."""
        assert tp_index < self.tp_degree and pp_index < self.pp_degree
        return [self._merge_state_dicts([self._load(f) for f in files]) for files in self.transformer_file_map[(tp_index, pp_index)]]

    def get_2d_parallel_files(self, tp_index: int, pp_index: int) -> List[str]:
        assert tp_index < self.tp_degree and pp_index < self.pp_degree
        return [self.mp_rank_files[i] for i in self.new_2d_map.get_data(pp_index=pp_index, tp_index=tp_index)]

    def get_2d_parallel_state(self, tp_index: int, pp_index: int) -> dict:
        """The ``mp_rank_*`` model-state files that fold into new rank (tp, pp), merged."""
        merged = None
        for f in self.get_2d_parallel_files(tp_index, pp_index):
            sd = self._load(f)
            merged = sd if merged is None else merge_state(merged, sd)
        return merged

    def get_checkpoint_info(self, info_key: str = CHECKPOINT_INFO_KEY):
        return self.get_model_state().get(info_key)

    def show_2d_mapping(self) -> None:
        print("reshaped 2d map ---- begin")
        for pp in range(self.pp_degree):
            for tp in range(self.tp_degree):
                print(f"[{pp}, {tp}] = {self.get_2d_parallel_files(tp_index=tp, pp_index=pp)}")
        print("reshaped 2d map ---- end")

    @staticmethod
    def _dump_mapping(data_map: dict, tag: str) -> None:
        print(f"Dump mapping: {tag}")
        for k, v in data_map.items():
            print(f"{k} = {v}")

    def show_tp_embedding_map(self) -> None:
        self._dump_mapping(self.tp_to_embedding_map, "tp_to_embedding_layers")

    def show_tp_final_norm_map(self) -> None:
        self._dump_mapping(self.tp_to_final_norm_map, "tp_to_final_norm_layers")

    def show_pp_transformer_map(self) -> None:
        self._dump_mapping(self.pp_to_transformer_map, "pp_to_transformer_layers")

    def show_transformer_file_map(self) -> None:
        self._dump_mapping(self.transformer_file_map, "rank_to_transformer_files")

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
