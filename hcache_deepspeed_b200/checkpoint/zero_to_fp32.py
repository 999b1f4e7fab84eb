"""Consolidate a ZeRO (stage 1/2/3) checkpoint directory into one fp32 ``state_dict`` — no GPUs, no process group:
reads every ``*_optim_states.pt`` shard, concatenates the fp32 partitions in rank order and cuts them back into the
parameters named by ``param_shapes``. Usable as a library and as a script (``python -m ...zero_to_fp32 ckpt_dir out``).
Reference: ``deepspeed/utils/zero_to_fp32.py`` (get_fp32_state_dict_from_zero_checkpoint :484,
convert_zero_checkpoint_to_fp32_state_dict :541, load_state_dict_from_zero_checkpoint :604)."""
from __future__ import annotations

import argparse
import dataclasses
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


# ---- names of the reference's script (utils/zero_to_fp32.py:40-160, 348, 391, 513) -----------------------------------------
@dataclasses.dataclass
class zero_model_state:
    buffers: dict
    param_shapes: list
    shared_params: dict
    ds_version: Optional[str]
    frozen_param_shapes: Optional[dict]
    frozen_param_fragments: Optional[dict]


def atoi(text: str):
    return int(text) if text.isdigit() else text


def natural_keys(text: str):
    """Sort key that orders ``rank_2`` before ``rank_10``."""
    return [atoi(c) for c in re.split(r"(\d+)", text)]


def get_optim_files(checkpoint_dir: str):
    return get_checkpoint_files(checkpoint_dir, "*_optim_states.pt")


def get_model_state_files(checkpoint_dir: str):
    return get_checkpoint_files(checkpoint_dir, "*_model_states.pt")


def parse_model_states(files) -> list:
    """One :class:`zero_model_state` per model-states file: buffers (fp32), parameter shapes, shared-parameter aliases and
    the frozen-parameter records."""
    out = []
    for f in files:
        sd = _load(f)
        if BUFFER_NAMES not in sd:
            raise ValueError(f"{f} is not a model state checkpoint")
        names = set(sd[BUFFER_NAMES])
        source = sd.get("buffers") or {k: v for k, v in (sd.get("module") or {}).items() if k in names}
        buffers = {k: (v.float() if torch.is_floating_point(v) else v) for k, v in source.items() if k in names}
        shapes = sd.get(PARAM_SHAPES)
        out.append(zero_model_state(buffers=buffers, param_shapes=[shapes] if isinstance(shapes, dict) else shapes,
                                    shared_params=sd.get("shared_params") or {}, ds_version=sd.get("ds_version"),
                                    frozen_param_shapes=sd.get("frozen_param_shapes"),
                                    frozen_param_fragments=sd.get("frozen_param_fragments")))
    return out


def parse_optim_states(files, ds_checkpoint_dir: str):
    """(zero stage, world size, per-rank list of flat fp32 groups) from the optimizer shard files."""
    shards = [_load(f)[OPTIMIZER_STATE_DICT] for f in files]
    if ZERO_STAGE not in shards[0]:
        raise ValueError(f"{files[0]} is not a zero checkpoint")
    stage = shards[0][ZERO_STAGE]
    world = shards[0].get(PARTITION_COUNT, len(shards))
    if isinstance(world, (list, tuple)):
        world = max(world)
    if world != len(files):
        raise ValueError(f"Expected {world} of '*_optim_states.pt' under '{ds_checkpoint_dir}' but found {len(files)} files. "
                         "Possibly due to an overwrite of an old checkpoint, or a checkpoint didn't get saved by one or more processes.")
    key = FP32_FLAT_GROUPS if stage == 3 else SINGLE_PARTITION_OF_FP32_GROUPS
    return stage, world, [s[key] for s in shards]


def zero3_partitioned_param_info(unpartitioned_numel: int, world_size: int):
    """(elements each rank holds, padding elements at the end of the last rank's slice) of a ZeRO-3 parameter."""
    remainder = unpartitioned_numel % world_size
    return math.ceil(unpartitioned_numel / world_size), (world_size - remainder) if remainder else 0


class GatheredTensor:
    """A parameter described by where its pieces sit in the loaded shards; the full fp32 tensor is only built by
    ``contiguous()``. ``lazy_mode`` of ``get_fp32_state_dict_from_zero_checkpoint`` returns these, so a converter can
    materialise and write one tensor at a time instead of holding the whole model twice."""

    def __init__(self, pieces, shape):
        self.pieces = pieces                 # [(flat tensor, start, length), ...] in order
        self.shape = torch.Size(shape)
        self.dtype = torch.float32

    def numel(self) -> int:
        return math.prod(self.shape) if len(self.shape) else 1

    def contiguous(self) -> torch.Tensor:
        n = self.numel()
        flat = torch.cat([t.narrow(0, lo, ln).float().reshape(-1) for t, lo, ln in self.pieces])
        return flat[:n].view(self.shape).clone()


def to_torch_tensor(state_dict, return_empty_tensor: bool = False):
    """Materialise the ``GatheredTensor`` entries of a lazy state dict (shared entries stay shared). With
    ``return_empty_tensor`` only correctly shaped empty tensors are made (to plan shard files cheaply)."""
    out, done = OrderedDict(), {}
    for name, t in state_dict.items():
        if id(t) in done:
            out[name] = out[done[id(t)]]
            continue
        done[id(t)] = name
        if isinstance(t, GatheredTensor):
            out[name] = torch.empty(t.shape, dtype=t.dtype) if return_empty_tensor else t.contiguous()
        else:
            out[name] = t
    return out


def _merge_stage12(shards, param_shapes, world, lazy: bool = False) -> Dict[str, torch.Tensor]:
    out = OrderedDict()
    n_groups = len(shards[0][SINGLE_PARTITION_OF_FP32_GROUPS])
    for gi in range(n_groups):
        flat = torch.cat([s[SINGLE_PARTITION_OF_FP32_GROUPS][gi].float().reshape(-1) for s in shards])
        off = 0
        for name, shape in param_shapes[gi].items():
            n = math.prod(shape) if len(shape) else 1
            out[name] = GatheredTensor([(flat, off, n)], shape) if lazy else flat.narrow(0, off, n).view(shape).clone()
            off += n
        assert flat.numel() - off < 2 * world + 2 * world, f"group {gi}: {flat.numel() - off} unconsumed elements (padding only expected)"
    return out


def _merge_stage3(shards, param_shapes, world, lazy: bool = False) -> Dict[str, torch.Tensor]:
    out = OrderedDict()
    n_groups = len(shards[0][FP32_FLAT_GROUPS])
    for gi in range(n_groups):
        offs = [0] * world
        for name, shape in param_shapes[gi].items():
            n = math.prod(shape) if len(shape) else 1
            psize, _ = zero3_partitioned_param_info(n, world)
            if lazy:
                out[name] = GatheredTensor([(shards[r][FP32_FLAT_GROUPS][gi], offs[r], psize) for r in range(world)], shape)
            else:
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
    sd = _merge_stage3(shards, param_shapes, world, lazy_mode) if stage == 3 else _merge_stage12(shards, param_shapes, world, lazy_mode)
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
    """Write the consolidated fp32 weights under ``output_dir``: one ``pytorch_model.bin`` / ``model.safetensors`` when it
    fits ``max_shard_size``, else numbered shard files plus a HuggingFace-style ``*.index.json``. Tensors are materialised
    one shard at a time (lazy mode), so peak memory is the shards read plus one output shard. A path ending in ``.bin`` /
    ``.pt`` asks for a single file there (the older calling convention)."""
    if output_dir.endswith((".bin", ".pt")):
        sd = get_fp32_state_dict_from_zero_checkpoint(checkpoint_dir, tag, exclude_frozen_parameters)
        os.makedirs(os.path.dirname(output_dir) or ".", exist_ok=True)
        torch.save(sd, output_dir)
        return output_dir
    lazy = get_fp32_state_dict_from_zero_checkpoint(checkpoint_dir, tag, exclude_frozen_parameters, lazy_mode=True)
    os.makedirs(output_dir, exist_ok=True)
    base, ext = ("model", ".safetensors") if safe_serialization else ("pytorch_model", ".bin")
    limit = _size_to_bytes(max_shard_size)
    shards, cur, cur_bytes, seen = [], [], 0, {}
    for name, t in lazy.items():
        if id(t) in seen:                    # a shared tensor travels with its first name
            cur.append(name)
            continue
        seen[id(t)] = name
        n = (t.numel() if isinstance(t, GatheredTensor) else t.numel()) * (4 if isinstance(t, GatheredTensor) else t.element_size())
        if cur and cur_bytes + n > limit:
            shards.append(cur)
            cur, cur_bytes = [], 0
        cur.append(name)
        cur_bytes += n

    if cur:
        shards.append(cur)

    def write(names, path):
        part = to_torch_tensor(OrderedDict((k, lazy[k]) for k in names))
        if safe_serialization:
            from safetensors.torch import save_file
            uniq, done = {}, set()
            for k, v in part.items():            # safetensors refuses aliased storage: keep the first name of a shared tensor
                if id(v) not in done:
                    uniq[k] = v.contiguous()
                    done.add(id(v))
            save_file(uniq, path, metadata={"format": "pt"})
        else:
            torch.save(part, path)
        return sum(v.numel() * v.element_size() for v in part.values())

    if len(shards) <= 1:
        path = os.path.join(output_dir, base + ext)
        write(shards[0] if shards else [], path)
        return path
    import json
    weight_map, total = {}, 0
    for i, names in enumerate(shards, start=1):
        fname = f"{base}-{i:05d}-of-{len(shards):05d}{ext}"
        total += write(names, os.path.join(output_dir, fname))
        weight_map.update({k: fname for k in names})
    index = os.path.join(output_dir, base + ext + ".index.json")
    with open(index, "w") as f:
        json.dump({"metadata": {"total_size": total}, "weight_map": weight_map}, f, indent=2, sort_keys=True)
    return index


def _size_to_bytes(size) -> int:
    if isinstance(size, (int, float)):
        return int(size)
    m = re.fullmatch(r"\s*([0-9.]+)\s*([KMGT]?)(i?)B?\s*", str(size), flags=re.I)
    if not m:
        raise ValueError(f"cannot read a size from {size!r} (expected e.g. '5GB', '200MB')")
    base = 1024 if m.group(3) else 1000
    return int(float(m.group(1)) * base ** " KMGT".index(m.group(2).upper() or " "))


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
