"""HuggingFace checkpoint directory reader.

Behaviour of ``deepspeed/inference/v2/checkpoint/huggingface_engine.py:16-124``: prefer ``*.safetensors``
(optionally through ``model.safetensors.index.json``), fall back to ``pytorch_model*.bin``; when the path is
not a local directory try the hub (needs network and ``huggingface_hub``). Tensors are yielded one shard
file at a time so host memory stays bounded.
"""
from __future__ import annotations

import json
import os
from typing import Iterable, List, Optional, Tuple

import torch

from ....utils.logging import inference_logger
from .base_engine import CheckpointEngineBase


class HuggingFaceCheckpointEngine(CheckpointEngineBase):

    def __init__(self, model_name_or_path: str, auth_token: Optional[str] = None, **hf_kwargs) -> None:
        super().__init__()
        from transformers import AutoConfig
        self.model_name_or_path = model_name_or_path
        self.auth_token = auth_token
        self.model_config = AutoConfig.from_pretrained(model_name_or_path, **hf_kwargs)
        if not hasattr(self.model_config, "max_seq_length"):
            for attr in ("max_position_embeddings", "n_positions", "seq_length"):
                if hasattr(self.model_config, attr):
                    self.model_config.max_seq_length = getattr(self.model_config, attr)
                    break
        self._local_checkpoint_dir: Optional[str] = None
        self._all_ckpt_paths = self._fetch_checkpoint_files()

    def _fetch_checkpoint_files(self) -> List[str]:
        path = self.model_name_or_path
        if not os.path.isdir(path):
            try:
                from huggingface_hub import snapshot_download
                path = snapshot_download(path, allow_patterns=["*.safetensors", "*.bin", "*.json", "*.pt"],
                                         revision=None, token=self.auth_token)
            except Exception as e:  # offline image: a local directory is required
                raise FileNotFoundError(f"{self.model_name_or_path} is not a local directory and could not be "
                                        f"downloaded ({type(e).__name__}: {e})") from e
        self._local_checkpoint_dir = path
        files = os.listdir(path)

        def from_index(index_name: str) -> Optional[List[str]]:
            p = os.path.join(path, index_name)
            if not os.path.exists(p):
                return None
            with open(p) as f:
                weight_map = json.load(f)["weight_map"]
            return [os.path.join(path, n) for n in sorted(set(weight_map.values()))]

        for index, single, ext in (("model.safetensors.index.json", "model.safetensors", ".safetensors"),
                                   ("pytorch_model.bin.index.json", "pytorch_model.bin", ".bin")):
            got = from_index(index)
            if got:
                return got
            if single in files:
                return [os.path.join(path, single)]
            shards = sorted(f for f in files if f.endswith(ext) and not f.startswith("training_args"))
            if shards:
                return [os.path.join(path, f) for f in shards]
        raise FileNotFoundError(f"no *.safetensors or *.bin weights under {path}")

    def parameters(self) -> Iterable[Tuple[str, torch.Tensor]]:
        for ckpt in self._all_ckpt_paths:
            inference_logger().info(f"Loading checkpoint: {ckpt}")
            if ckpt.endswith(".safetensors"):
                from safetensors import safe_open
                with safe_open(ckpt, framework="pt", device="cpu") as f:
                    for name in f.keys():
                        yield name, f.get_tensor(name)
            else:
                sd = torch.load(ckpt, map_location="cpu", weights_only=True)
                if isinstance(sd, dict) and "state_dict" in sd and isinstance(sd["state_dict"], dict):
                    sd = sd["state_dict"]
                for name, tensor in sd.items():
                    yield name, tensor
                del sd
