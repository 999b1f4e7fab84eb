"""Convert a ZeRO checkpoint into the universal format: one directory per parameter holding the full fp32 weight and
its optimizer moments (``zero/<param>/{fp32,exp_avg,exp_avg_sq,step}.pt``), independent of the parallel layout it was
trained with; any other layout can then resume from it (``checkpoint.load_universal``).
Reference: ``deepspeed/checkpoint/ds_to_universal.py`` (extract_zero_shards :112, merge_tp_slices :232, main :469)."""
from __future__ import annotations

import argparse
import math
import os
import shutil
from typing import Dict, Optional

import torch

from .constants import (BASE_OPTIMIZER_STATE, OPTIMIZER_STATE_DICT, PARAM_SHAPES, PARTITION_COUNT,
                        SINGLE_PARTITION_OF_FP32_GROUPS, UNIVERSAL_CHECKPOINT_INFO, UNIVERSAL_CHECKPOINT_VERSION_KEY,
                        UNIVERSAL_CHECKPOINT_VERSION_VALUE, ZERO_STAGE)
from .zero_to_fp32 import _load, _resolve_tag_dir, get_checkpoint_files, get_model_state_file


def _flat_state(shards, key: str, gi: int):
    """Concatenate one optimizer-state tensor (e.g. exp_avg) of param group ``gi`` over the data-parallel shards."""
    parts = []
    for s in shards:
        st = s[BASE_OPTIMIZER_STATE]["state"]
        if gi not in st or key not in st[gi]:
            return None
        parts.append(st[gi][key].float().reshape(-1))
    return torch.cat(parts)


def convert_to_universal(input_folder: str, output_folder: str, tag: Optional[str] = None, num_extract_workers: int = 1,
                         num_merge_workers: int = 1, keep_temp_folder: bool = False, strict: bool = True) -> str:
    """Write ``output_folder`` (``zero/<param>/*.pt`` + the model-states file) from the ZeRO-1/2 checkpoint
    ``input_folder/tag``. Returns the output folder."""
    ds_dir = _resolve_tag_dir(input_folder, tag) if (tag is not None or os.path.isfile(os.path.join(input_folder, "latest"))) \
        else input_folder
    optim_files = get_checkpoint_files(ds_dir, "*_optim_states.pt")
    shards = [_load(f)[OPTIMIZER_STATE_DICT] for f in optim_files]
    stage = shards[0].get(ZERO_STAGE, 2)
    if stage == 3:
        raise NotImplementedError("universal conversion of ZeRO-3 checkpoints: consolidate with zero_to_fp32 first")
    world = shards[0].get(PARTITION_COUNT, len(shards))
    assert world == len(shards), f"expected {world} optimizer shards, found {len(shards)}"
    param_shapes = shards[0][PARAM_SHAPES]
    zero_dir = os.path.join(output_folder, "zero")
    os.makedirs(zero_dir, exist_ok=True)
    n_groups = len(shards[0][SINGLE_PARTITION_OF_FP32_GROUPS])
    for gi in range(n_groups):
        flats = {"fp32": torch.cat([s[SINGLE_PARTITION_OF_FP32_GROUPS][gi].float().reshape(-1) for s in shards])}
        for key in ("exp_avg", "exp_avg_sq"):
            f = _flat_state(shards, key, gi)
            if f is not None:
                flats[key] = f
        step = shards[0][BASE_OPTIMIZER_STATE]["state"].get(gi, {}).get("step")
        if step is None:
            pg = shards[0][BASE_OPTIMIZER_STATE]["param_groups"]
            step = pg[gi].get("step") if gi < len(pg) else None
        off = 0
        for name, shape in param_shapes[gi].items():
            n = math.prod(shape) if len(shape) else 1
            pdir = os.path.join(zero_dir, name)
            os.makedirs(pdir, exist_ok=True)
            for key, flat in flats.items():
                torch.save({"param": flat.narrow(0, off, n).view(shape).clone()}, os.path.join(pdir, f"{key}.pt"))
            if step is not None:
                torch.save(step, os.path.join(pdir, "step.pt"))
            off += n
    model_file = get_model_state_file(ds_dir, stage)
    msd = _load(model_file)
    msd[UNIVERSAL_CHECKPOINT_INFO] = {UNIVERSAL_CHECKPOINT_VERSION_KEY: UNIVERSAL_CHECKPOINT_VERSION_VALUE}
    torch.save(msd, os.path.join(output_folder, os.path.basename(model_file)))
    with open(os.path.join(os.path.dirname(output_folder.rstrip("/")) or ".", "latest_universal"), "w") as f:
        f.write(os.path.basename(output_folder.rstrip("/")))
    return output_folder


def main(args=None):
    p = argparse.ArgumentParser()
    p.add_argument("--input_folder", type=str, required=True, help="Input DeepSpeed Checkpoint folder")
    p.add_argument("--output_folder", type=str, required=True, help="Output DeepSpeed checkpoint folder")
    p.add_argument("--num_extract_workers", default=4, type=int)
    p.add_argument("--num_merge_workers", default=2, type=int)
    p.add_argument("--keep_temp_folder", action="store_true")
    p.add_argument("--no_strict", dest="strict", action="store_false")
    a = p.parse_args(args)
    print(convert_to_universal(a.input_folder, a.output_folder, None, a.num_extract_workers, a.num_merge_workers,
                               a.keep_temp_folder, a.strict))


if __name__ == "__main__":
    main()
