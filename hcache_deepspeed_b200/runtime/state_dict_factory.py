"""Load Megatron-style tensor-parallel checkpoints under a different TP degree: merge (ckpt ranks > target ranks) or
split (ckpt ranks < target ranks) the sharded tensors by key pattern — column-parallel weights along dim 0,
row-parallel along dim 1, fused QKV per (q|k|v) third. Reference: ``deepspeed/runtime/state_dict_factory.py:22-430``."""
from __future__ import annotations

import json
import os
from abc import ABC, abstractmethod
from typing import List, Optional

import torch

AUTO_MODULE_KEY = "auto"


class SDLoaderFactory:

    @staticmethod
    def get_sd_loader_json(json_file, checkpoint_engine=None):
        data = json_file if isinstance(json_file, dict) else json.load(open(json_file))
        sd_type, ckpt_list, version = data["type"], data["checkpoints"], data.get("version", 0.0)
        if sd_type.lower() in ("bloom", "ds_model"):
            return data
        return SDLoaderFactory.get_sd_loader(ckpt_list, checkpoint_engine, sd_type, version)

    @staticmethod
    def get_sd_loader(ckpt_list, checkpoint_engine=None, sd_type="Megatron", version=None):
        if sd_type == "Megatron":
            return MegatronSDLoader(ckpt_list, version, checkpoint_engine)
        raise NotImplementedError(f"{sd_type} checkpoint type is not supported")


class SDLoaderBase(ABC):

    def __init__(self, ckpt_list: List[str], version, checkpoint_engine=None):
        self.module_key = None
        self.ckpt_list = ckpt_list
        self.version = version
        self.checkpoint_engine = checkpoint_engine
        self.check_ckpt_list()

    def _load_one(self, path):
        if self.checkpoint_engine is not None:
            return self.checkpoint_engine.load(path, map_location="cpu")
        return torch.load(path, map_location="cpu", weights_only=False)

    def load(self, mp_world_size: int, mp_rank: int, module_key=AUTO_MODULE_KEY, is_pipe_parallel=False, quantize=False,
             quantize_bits=8, quantize_groups=64, mlp_extra_grouping=True):
        self.module_key = module_key
        n = len(self.ckpt_list)
        idx = mp_rank * n // mp_world_size
        load_path = self.ckpt_list[idx]
        if n == mp_world_size:
            sd = self._load_one(load_path)
            return load_path, sd, (None, None)
        if n > mp_world_size:
            return (load_path, ) + (self.merge_state_dict(mp_world_size, mp_rank), (None, None))
        return (load_path, ) + (self.split_state_dict(mp_world_size, mp_rank), (None, None))

    def get_merge_state_dicts(self, mp_world_size, mp_rank):
        n = len(self.ckpt_list)
        assert n % mp_world_size == 0, "Invalid checkpoints and world size for sd merge"
        per = n // mp_world_size
        return [self._load_one(p) for p in self.ckpt_list[per * mp_rank:per * (mp_rank + 1)]]

    def get_split_state_dict(self, mp_world_size, mp_rank):
        n = len(self.ckpt_list)
        assert mp_world_size % n == 0, "Invalid checkpoints and world size for sd split"
        num_to_split = mp_world_size // n
        return self._load_one(self.ckpt_list[mp_rank // num_to_split]), num_to_split, mp_rank % num_to_split

    def _choose_module_key(self, sd):
        assert not ("module" in sd and "model" in sd), "checkpoint has both 'model' and 'module' keys, not sure how to proceed"
        assert "module" in sd or "model" in sd, "checkpoint contains neither 'model' or 'module' keys, not sure how to proceed"
        return "module" if "module" in sd else "model"

    def get_module(self, sd):
        if self.module_key is None:
            return sd
        if self.module_key == AUTO_MODULE_KEY:
            return sd[self._choose_module_key(sd)]
        return sd[self.module_key]

    def set_module(self, sd, module):
        if self.module_key is None:
            sd = module
        elif self.module_key == AUTO_MODULE_KEY:
            sd[self._choose_module_key(sd)] = module
        else:
            sd[self.module_key] = module
        return sd

    def check_ckpt_list(self):
        assert len(self.ckpt_list) > 0

    @abstractmethod
    def merge_state_dict(self, mp_world_size, mp_rank):
        ...

    @abstractmethod
    def split_state_dict(self, mp_world_size, mp_rank):
        ...


class MegatronSDLoader(SDLoaderBase):
    COLUMN = ("mlp.dense_h_to_4h.weight", "mlp.dense_h_to_4h.bias", "word_embeddings.weight", "final_linear.weight")
    ROW = ("attention.dense.weight", "mlp.dense_4h_to_h.weight")
    QKV = ("attention.query_key_value", )

    def merge_query_key_value(self, param_list, ckpt_ver):
        """version 0: [q_all|k_all|v_all] per shard -> regroup thirds; version >= 1: per-head interleave, plain cat."""
        if ckpt_ver == 0:
            thirds = [torch.chunk(p, 3, dim=0) for p in param_list]
            return torch.cat([torch.cat([t[i] for t in thirds], dim=0) for i in range(3)], dim=0)
        return torch.cat(param_list, dim=0)

    def split_query_key_value(self, param, num_to_split, offset, ckpt_ver):
        if ckpt_ver == 0:
            thirds = torch.chunk(param, 3, dim=0)
            return torch.cat([torch.chunk(t, num_to_split, dim=0)[offset] for t in thirds], dim=0)
        return torch.chunk(param, num_to_split, dim=0)[offset]

    def merge_state_dict(self, mp_world_size, mp_rank):
        sds = self.get_merge_state_dicts(mp_world_size, mp_rank)
        ver = self.version or 0
        out = dict(sds[0])
        mods = [self.get_module(s) for s in sds]
        new = {}
        for k in mods[0]:
            vals = [m[k] for m in mods]
            if any(q in k for q in self.QKV):
                new[k] = self.merge_query_key_value(vals, ver)
            elif any(c in k for c in self.COLUMN):
                new[k] = torch.cat(vals, dim=0)
            elif any(r in k for r in self.ROW):
                new[k] = torch.cat(vals, dim=1)
            else:
                new[k] = vals[0]
        return self.set_module(out, new)

    def split_state_dict(self, mp_world_size, mp_rank):
        sd, num_to_split, offset = self.get_split_state_dict(mp_world_size, mp_rank)
        ver = self.version or 0
        out = dict(sd)
        new = {}
        for k, v in self.get_module(sd).items():
            if any(q in k for q in self.QKV):
                new[k] = self.split_query_key_value(v, num_to_split, offset, ver)
            elif any(c in k for c in self.COLUMN):
                new[k] = torch.chunk(v, num_to_split, dim=0)[offset]
            elif any(r in k for r in self.ROW):
                new[k] = torch.chunk(v, num_to_split, dim=1)[offset]
            else:
                new[k] = v
        return self.set_module(out, new)
