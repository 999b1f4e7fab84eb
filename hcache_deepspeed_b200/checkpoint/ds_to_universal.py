"""Convert a ZeRO checkpoint into the universal format: one directory per parameter holding the full fp32 weight and
its optimizer moments (``zero/<param>/{fp32,exp_avg,exp_avg_sq,step}.pt``), independent of the parallel layout it was
trained with; any other layout can then resume from it (``checkpoint.load_universal``).
Reference: ``deepspeed/checkpoint/ds_to_universal.py`` (extract_zero_shards :112, merge_tp_slices :232, main :469)."""
from __future__ import annotations

import argparse
import math
import os
import re
from typing import Dict, Optional

import torch

from .constants import (BASE_OPTIMIZER_STATE, CAT_DIM, FP32_FLAT_GROUPS, OPTIMIZER_STATE_DICT, ORIGINAL_VOCAB_SIZE,
                        PARAM_N_SUB_PARAMS, PARAM_SHAPES, PARAMETER_TO_AVERAGE_PATTERNS, PARAMETER_WITH_2_SUB_PARAMS_CAT_DIM_0,
                        PARAMETER_WITH_ROW_PARALLELISM_PATTERNS, PARAMETER_WITH_SUB_PARAMS, PARTITION_COUNT,
                        SINGLE_PARTITION_OF_FP32_GROUPS, SUB_PARAM_SHAPE, TP_REPLICATED_PARAMETER_PATTERNS,
                        UNIVERSAL_CHECKPOINT_INFO, UNIVERSAL_CHECKPOINT_VERSION_KEY, UNIVERSAL_CHECKPOINT_VERSION_VALUE,
                        VOCAB_TENSOR, VOCABULARY_PARAMETER_PATTERNS, ZERO_STAGE)
from .universal_checkpoint import SubparamShape
from .zero_to_fp32 import _load, _resolve_tag_dir, get_checkpoint_files, get_model_state_file


def atoi(text: str):
    return int(text) if text.isdigit() else text


def natural_keys(text: str):
    return [atoi(c) for c in re.split(r"(\d+)", text)]


def dp_index_to_str(dp_index: int) -> str:
    return f"{dp_index:0>2d}"


def _tp_rank_of(path: str) -> int:
    m = re.search(r"mp_rank_(\d+)", os.path.basename(path))
    return int(m.group(1)) if m else 0


def _dp_rank_of(path: str) -> int:
    m = re.search(r"zero_pp_rank_(\d+)", os.path.basename(path))
    return int(m.group(1)) if m else 0


def _inner(shard) -> Dict:
    """The wrapped torch optimizer's own state dict (stage 1/2 and stage 3 shards name it differently)."""
    return shard.get(BASE_OPTIMIZER_STATE) or shard[OPTIMIZER_STATE_DICT]


def _group_state(shard, gi: int, key: str):
    st = _inner(shard)["state"]
    return st[gi][key].float().reshape(-1) if gi in st and key in st[gi] and torch.is_tensor(st[gi][key]) else None


def _step_of(shard, gi: int):
    step = _inner(shard)["state"].get(gi, {}).get("step")
    if step is None:
        pg = _inner(shard)["param_groups"]
        step = pg[gi].get("step") if gi < len(pg) else None
    return step


def extract_zero_shards(shards, param_shapes) -> Dict[str, Dict[str, torch.Tensor]]:
    """ZeRO-1/2: every group's flat fp32 / moment buffers are the concatenation of the data-parallel shards; parameters lie
    back to back inside. Returns ``{param name: {"fp32" | "exp_avg" | "exp_avg_sq": tensor, "step": value}}``."""
    out: Dict[str, Dict[str, torch.Tensor]] = {}
    for gi in range(len(shards[0][SINGLE_PARTITION_OF_FP32_GROUPS])):
        flats = {"fp32": torch.cat([s[SINGLE_PARTITION_OF_FP32_GROUPS][gi].float().reshape(-1) for s in shards])}
        for key in ("exp_avg", "exp_avg_sq"):
            parts = [_group_state(s, gi, key) for s in shards]
            if all(p is not None for p in parts):
                flats[key] = torch.cat(parts)
        step, off = _step_of(shards[0], gi), 0
        for name, shape in param_shapes[gi].items():
            n = math.prod(shape) if len(shape) else 1
            out[name] = {k: f.narrow(0, off, n).view(shape).clone() for k, f in flats.items()}
            if step is not None:
                out[name]["step"] = step
            off += n
    return out


def extract_zero_shards_stage3(shards, param_shapes) -> Dict[str, Dict[str, torch.Tensor]]:
    """ZeRO-3: every parameter is cut into ``world`` equal (padded) slices, rank ``r`` keeps slice ``r`` of each at a
    running offset of its flat group; moments follow the same layout."""
    world = len(shards)
    out: Dict[str, Dict[str, torch.Tensor]] = {}
    for gi in range(len(shards[0][FP32_FLAT_GROUPS])):
        flats = {"fp32": [s[FP32_FLAT_GROUPS][gi].float().reshape(-1) for s in shards]}
        for key in ("exp_avg", "exp_avg_sq"):
            parts = [_group_state(s, gi, key) for s in shards]
            if all(p is not None for p in parts):
                flats[key] = parts
        step, off = _step_of(shards[0], gi), 0
        for name, shape in param_shapes[gi].items():
            n = math.prod(shape) if len(shape) else 1
            psize = (n + world - 1) // world
            out[name] = {k: merge_zero3_slices([f.narrow(0, off, psize) for f in per_rank], shape)
                         for k, per_rank in flats.items()}
            if step is not None:
                out[name]["step"] = step
            off += psize
    return out


def merge_zero3_slices(slices, shape) -> torch.Tensor:
    """Rank-ordered slices of one ZeRO-3 parameter -> the full tensor (padding of the last slice dropped)."""
    n = math.prod(shape) if len(shape) else 1
    return torch.cat([s.reshape(-1) for s in slices])[:n].view(shape).clone()


def _match(patterns, name: str):
    hits = [p for p in patterns if re.match(p, name)]
    assert len(hits) <= 1, f"Got more than one matching patterns={hits} for {name}"
    return hits[0] if hits else None


def merge_tp_slices(name: str, slices, info: Dict):
    """One parameter's tensor-parallel slices (rank order) -> (full tensor, metadata for the loader). ``info`` is the
    ``universal_checkpoint_info`` of the model-states file: patterns for parameters that are replicated, averaged,
    row-parallel (joined along dim 1), made of two stacked halves (gate/up), made of several sub-parameters, or
    vocabulary-sized (padding rows dropped). Everything else is column-parallel: joined along dim 0."""
    meta: Dict[str, object] = {}
    sub = next((SubparamShape(**d) for d in info.get(PARAMETER_WITH_SUB_PARAMS, [])
                if any(re.match(pat, name) for pat in SubparamShape(**d).patterns)), None)
    if len(slices) == 1 and not _match(info.get(VOCABULARY_PARAMETER_PATTERNS, []), name):
        param = slices[0]
    elif _match(info.get(TP_REPLICATED_PARAMETER_PATTERNS, []), name):
        param = slices[0]
    elif _match(info.get(PARAMETER_TO_AVERAGE_PATTERNS, []), name):
        param = sum(slices) / len(slices)
    elif _match(info.get(PARAMETER_WITH_2_SUB_PARAMS_CAT_DIM_0, []), name):
        halves = [torch.chunk(s, 2, dim=0) for s in slices]
        param = torch.cat([torch.cat([h[0] for h in halves], dim=0), torch.cat([h[1] for h in halves], dim=0)], dim=0)
        meta[CAT_DIM], meta[PARAM_N_SUB_PARAMS] = 0, 2
    elif sub is not None:
        dim, tp = sub.partition_dim, len(slices)
        sizes = sub.shape[dim] if isinstance(sub.shape[dim], (tuple, list)) else (sub.shape[dim], )
        part_shape = [(sum(d) if isinstance(d, (tuple, list)) else d) for d in sub.shape]
        part_shape[dim] //= tp
        views = [s.reshape(part_shape) for s in slices]
        chunks, off = [], 0
        for size in sizes:
            chunks.append(torch.cat([v.narrow(dim, off, size // tp) for v in views], dim=dim))
            off += size // tp
        param = torch.cat(chunks, dim=dim)
        meta[SUB_PARAM_SHAPE] = sub
    else:
        cat_dim = 1 if _match(info.get(PARAMETER_WITH_ROW_PARALLELISM_PATTERNS, []), name) else 0
        param = torch.cat(slices, dim=cat_dim) if slices[0].dim() > cat_dim else slices[0]
        meta[CAT_DIM] = cat_dim
    if _match(info.get(VOCABULARY_PARAMETER_PATTERNS, []), name) and ORIGINAL_VOCAB_SIZE in info:
        param = param[:info[ORIGINAL_VOCAB_SIZE]]
        meta[VOCAB_TENSOR] = True
    return param, meta


def dump_param_fragment(zero_dir: str, name: str, states: Dict[str, object], meta: Optional[Dict] = None) -> None:
    """Write one parameter's directory: ``{fp32, exp_avg, exp_avg_sq}.pt`` (each ``{"param": tensor, ...metadata}``) and
    ``step.pt``."""
    pdir = os.path.join(zero_dir, name)
    os.makedirs(pdir, exist_ok=True)
    for key, value in states.items():
        if key == "step":
            torch.save(value, os.path.join(pdir, "step.pt"))
        else:
            torch.save({"param": value.clone(), **(meta or {})}, os.path.join(pdir, f"{key}.pt"))


def convert_to_universal(input_folder: str, output_folder: str, tag: Optional[str] = None, num_extract_workers: int = 1,
                         num_merge_workers: int = 1, keep_temp_folder: bool = False, strict: bool = True) -> str:
    """Write ``output_folder`` (``zero/<param>/*.pt`` + the model-states file) from the ZeRO-1/2/3 checkpoint
    ``input_folder/tag``; tensor-parallel slices (``mp_rank_XX`` files) are merged by the rules of the checkpoint's
    ``universal_checkpoint_info``. Returns the output folder."""
    ds_dir = _resolve_tag_dir(input_folder, tag) if (tag is not None or os.path.isfile(os.path.join(input_folder, "latest"))) \
        else input_folder
    optim_files = get_checkpoint_files(ds_dir, "*_optim_states.pt")
    by_tp: Dict[int, list] = {}
    for f in optim_files:
        by_tp.setdefault(_tp_rank_of(f), []).append(f)
    per_tp, stage = {}, None
    for tp_rank, files in sorted(by_tp.items()):
        shards = [_load(f)[OPTIMIZER_STATE_DICT] for f in sorted(files, key=_dp_rank_of)]
        stage = shards[0].get(ZERO_STAGE, 2)
        world = shards[0].get(PARTITION_COUNT, len(shards))
        world = max(world) if isinstance(world, (list, tuple)) else world
        assert world == len(shards), f"expected {world} optimizer shards for mp_rank {tp_rank}, found {len(shards)}"
        model_file = get_model_state_file(ds_dir, stage) if tp_rank == 0 else \
            get_model_state_file(ds_dir, stage).replace("mp_rank_00", f"mp_rank_{tp_rank:02d}")
        shapes = shards[0].get(PARAM_SHAPES) or _load(model_file).get(PARAM_SHAPES)
        shapes = [shapes] if isinstance(shapes, dict) else shapes
        per_tp[tp_rank] = (extract_zero_shards_stage3 if stage == 3 else extract_zero_shards)(shards, shapes)
    model_file = get_model_state_file(ds_dir, stage)
    msd = _load(model_file)
    info = dict(msd.get(UNIVERSAL_CHECKPOINT_INFO) or {})
    zero_dir = os.path.join(output_folder, "zero")
    os.makedirs(zero_dir, exist_ok=True)
    tp_ranks = sorted(per_tp)
    for name in per_tp[tp_ranks[0]]:
        states, meta = {}, {}
        for key in per_tp[tp_ranks[0]][name]:
            if key == "step":
                states[key] = per_tp[tp_ranks[0]][name][key]
                continue
            slices = [per_tp[r][name][key] for r in tp_ranks if name in per_tp[r]]
            if strict and len(slices) != len(tp_ranks):
                raise KeyError(f"parameter {name} is missing from some tensor-parallel ranks")
            states[key], meta = merge_tp_slices(name, slices, info)
        dump_param_fragment(zero_dir, name, states, meta)
    info[UNIVERSAL_CHECKPOINT_VERSION_KEY] = UNIVERSAL_CHECKPOINT_VERSION_VALUE
    msd[UNIVERSAL_CHECKPOINT_INFO] = info
    torch.save(msd, os.path.join(output_folder, os.path.basename(model_file)))
    with open(os.path.join(os.path.dirname(output_folder.rstrip("/")) or ".", "latest_universal"), "w") as f:
        f.write(os.path.basename(output_folder.rstrip("/")))
    return output_folder


def parse_arguments(args=None):
    p = argparse.ArgumentParser()
    p.add_argument("--input_folder", type=str, required=True, help="Input DeepSpeed Checkpoint folder")
    p.add_argument("--output_folder", type=str, required=True, help="Output DeepSpeed checkpoint folder")
    p.add_argument("--num_extract_workers", default=4, type=int)
    p.add_argument("--num_merge_workers", default=2, type=int)
    p.add_argument("--keep_temp_folder", action="store_true")
    p.add_argument("--no_strict", dest="strict", action="store_false")
    p.add_argument("--inject_missing_state", action="store_true")
    return p.parse_args(args)


def main(args=None):
    a = parse_arguments(args)
    print(convert_to_universal(a.input_folder, a.output_folder, None, a.num_extract_workers, a.num_merge_workers,
                               a.keep_temp_folder, a.strict))


if __name__ == "__main__":
    main()
