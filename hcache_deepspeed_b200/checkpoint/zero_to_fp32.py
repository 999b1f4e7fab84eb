"""Consolidate a ZeRO (stage 1/2/3) checkpoint directory into one fp32 ``state_dict`` — no GPUs, no process group:
reads every ``*_optim_states.pt`` shard, concatenates the fp32 partitions in rank order and cuts them back into the
parameters named by ``param_shapes``. Usable as a library and as a script (``python -m ...zero_to_fp32 ckpt_dir out``).
Reference: ``deepspeed/utils/zero_to_fp32.py`` (get_fp32_state_dict_from_zero_checkpoint :484,
convert_zero_checkpoint_to_fp32_state_dict :541, load_state_dict_from_zero_checkpoint :604)."""
from __future__ import annotations

import argparse
import glob
import math
import os
import re
from collections import OrderedDict
from typing import Dict, Optional

import torch

from .constants import (BUFFER_NAMES, FP32_FLAT_GROUPS, OPTIMIZER_STATE_DICT, PARAM_SHAPES, PARTITION_COUNT,
                        SINGLE_PARTITION_OF_FP32_GROUPS, ZERO_STAGE)


def _natural_key(s: str):
    return [int(t) if t.isdigit() else t for t in re.split(r"(\d+)", s)]


def get_checkpoint_files(checkpoint_dir: str, pattern: str):
    files = sorted(glob.glob(os.path.join(checkpoint_dir, pattern)), key=_natural_key)
    if not files:
        raise FileNotFoundError(f"can't find {pattern} files in directory '{checkpoint_dir}'")
    return files


def get_model_state_file(checkpoint_dir: str, zero_stage: int) -> str:
    if not os.path.isdir(checkpoint_dir):
        raise FileNotFoundError(f"Directory '{checkpoint_dir}' doesn't exist")
    name = "mp_rank_00_model_states.pt" if zero_stage <= 2 else "zero_pp_rank_0_mp_rank_00_model_states.pt"
    path = os.path.join(checkpoint_dir, name)
    if not os.path.exists(path):
        raise FileNotFoundError(f"can't find model states file at '{path}'")
    return path


def _load(path):
    return torch.load(path, map_location="cpu", weights_only=False)


def _resolve_tag_dir(checkpoint_dir: str, tag: Optional[str]) -> str:
    if tag is None:
        latest = os.path.join(checkpoint_dir, "latest")
        if os.path.isfile(latest):
            with open(latest) as f:
                tag = f.read().strip()
        else:
            raise ValueError(f"Unable to find 'latest' file at {latest}")
    ds_dir = os.path.join(checkpoint_dir, tag)
    if not os.path.isdir(ds_dir):
        raise FileNotFoundError(f"Directory '{ds_dir}' doesn't exist")
    return ds_dir


def _merge_stage12(shards, param_shapes, world) -> Dict[str, torch.Tensor]:
    out = OrderedDict()
    n_groups = len(shards[0][SINGLE_PARTITION_OF_FP32_GROUPS])
    for gi in range(n_groups):
        flat = torch.cat([s[SINGLE_PARTITION_OF_FP32_GROUPS][gi].float().reshape(-1) for s in shards])
        off = 0
        for name, shape in param_shapes[gi].items():
            n = math.prod(shape) if len(shape) else 1
            out[name] = flat.narrow(0, off, n).view(shape).clone()
            off += n
        assert flat.numel() - off < 2 * world + 2 * world, f"group {gi}: {flat.numel() - off} unconsumed elements (padding only expected)"
    return out


def _merge_stage3(shards, param_shapes, world) -> Dict[str, torch.Tensor]:
    out = OrderedDict()
    n_groups = len(shards[0][FP32_FLAT_GROUPS])
    for gi in range(n_groups):
        offs = [0] * world
        for name, shape in param_shapes[gi].items():
            n = math.prod(shape) if len(shape) else 1
            psize = (n + world - 1) // world
            parts = [shards[r][FP32_FLAT_GROUPS][gi].float().narrow(0, offs[r], psize) for r in range(world)]
            out[name] = torch.cat(parts)[:n].view(shape).clone()
            offs = [o + psize for o in offs]
    return out


def get_fp32_state_dict_from_zero_checkpoint(checkpoint_dir: str, tag: Optional[str] = None,
                                             exclude_frozen_parameters: bool = False, lazy_mode: bool = False):
    """fp32 ``state_dict`` (parameters + buffers) reconstructed from the shards under ``checkpoint_dir/tag``."""
    ds_dir = _resolve_tag_dir(checkpoint_dir, tag)
    optim_files = get_checkpoint_files(ds_dir, "*_optim_states.pt")
    shards = [_load(f)[OPTIMIZER_STATE_DICT] for f in optim_files]
    stage = shards[0].get(ZERO_STAGE, 2)
    world = shards[0].get(PARTITION_COUNT, len(shards))
    if isinstance(world, (list, tuple)):
        world = max(world)
    if world != len(shards):
        raise ValueError(f"Expected {world} of '*_optim_states.pt' under '{ds_dir}' but found {len(shards)} files. "
                         "Possibly due to an overwrite of an old checkpoint, or a checkpoint didn't get saved by one or more processes.")
    model_state = _load(get_model_state_file(ds_dir, stage))
    # shapes live in the model-states file (the reference's layout: utils/zero_to_fp32.py:100-160 there); older checkpoints
    # of this code base kept them in the optimizer shard
    param_shapes = model_state.get(PARAM_SHAPES) or shards[0].get(PARAM_SHAPES)
    if param_shapes is None:
        raise KeyError(f"no '{PARAM_SHAPES}' in the model-states or optimizer files under {ds_dir}")
    if isinstance(param_shapes, dict):
        param_shapes = [param_shapes]
    sd = _merge_stage3(shards, param_shapes, world) if stage == 3 else _merge_stage12(shards, param_shapes, world)
    buffers = model_state.get("buffers") if stage == 3 else {k: v for k, v in model_state["module"].items()
                                                            if k in set(model_state.get(BUFFER_NAMES, []))}
    for k, v in (buffers or {}).items():
        sd[k] = v.float() if torch.is_floating_point(v) else v
    if stage != 3 and not exclude_frozen_parameters:
        for k, v in model_state["module"].items():  # frozen parameters are not in any optimizer group
            if k not in sd:
                sd[k] = v.float() if torch.is_floating_point(v) else v
    if stage == 3 and not exclude_frozen_parameters and model_state.get("frozen_param_shapes"):
        # every rank's model-states file holds its slice of each frozen parameter
        states = [model_state] + [_load(f) for f in sorted(glob.glob(os.path.join(ds_dir, "zero_pp_rank_*_model_states.pt")),
                                                            key=lambda f: int(re.search(r"zero_pp_rank_(\d+)_", f).group(1)))][1:]
        for name, shape in model_state["frozen_param_shapes"].items():
            n = math.prod(shape) if len(shape) else 1
            flat = torch.cat([st["frozen_param_fragments"][name].float().reshape(-1) for st in states])
            sd[name] = flat[:n].view(shape).clone()
    for alias, stored in (model_state.get("shared_params") or {}).items():
        if stored in sd and alias not in sd:
            sd[alias] = sd[stored]
    return sd


def convert_zero_checkpoint_to_fp32_state_dict(checkpoint_dir: str, output_dir: str, max_shard_size: str = "5GB",
                                               safe_serialization: bool = False, tag: Optional[str] = None,
                                               exclude_frozen_parameters: bool = False):
    sd = get_fp32_state_dict_from_zero_checkpoint(checkpoint_dir, tag, exclude_frozen_parameters)
    os.makedirs(output_dir, exist_ok=True) if not output_dir.endswith((".bin", ".pt")) else None
    if output_dir.endswith((".bin", ".pt")):
        os.makedirs(os.path.dirname(output_dir) or ".", exist_ok=True)
        torch.save(sd, output_dir)
        return output_dir
    if safe_serialization:
        from safetensors.torch import save_file
        path = os.path.join(output_dir, "model.safetensors")
        save_file({k: v.contiguous() for k, v in sd.items()}, path)
    else:
        path = os.path.join(output_dir, "pytorch_model.bin")
        torch.save(sd, path)
    return path


def load_state_dict_from_zero_checkpoint(model: torch.nn.Module, checkpoint_dir: str, tag: Optional[str] = None):
    sd = get_fp32_state_dict_from_zero_checkpoint(checkpoint_dir, tag)
    model = model.cpu()
    model.load_state_dict(sd, strict=False)
    return model


def main(args=None):
    p = argparse.ArgumentParser()
    p.add_argument("checkpoint_dir", type=str, help="path to the desired checkpoint folder, e.g., path/checkpoint-12")
    p.add_argument("output_dir", type=str, help="directory (or .bin/.pt file) for the consolidated fp32 weights")
    p.add_argument("--max_shard_size", type=str, default="5GB")
    p.add_argument("--safe_serialization", default=False, action="store_true")
    p.add_argument("-t", "--tag", type=str, default=None)
    p.add_argument("--exclude_frozen_parameters", action="store_true")
    a = p.parse_args(args)
    print(convert_zero_checkpoint_to_fp32_state_dict(a.checkpoint_dir, a.output_dir, a.max_shard_size, a.safe_serialization,
                                                     a.tag, a.exclude_frozen_parameters))


if __name__ == "__main__":
    main()
