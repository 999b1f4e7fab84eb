"""Flat-buffer (de)serialisation of kernel-ready weights.

Same on-disk contract as the reference (``deepspeed/inference/v2/model_implementations/flat_model_helpers.py:59-282``,
``engine_v2.py:284-301``): per TP rank one ``params_rank_{r}_of_{n}.pt`` holding a single uint8 tensor with every
parameter at a 256-byte-aligned offset, one ``metadata_rank_{r}_of_{n}.json`` describing name / shape / dtype /
offset, and a rank-0 ``ds_model_config.pkl``. The serialized form is specific to the TP degree, as in the reference.
"""
from __future__ import annotations

import os
from typing import Dict, Tuple

import torch

from .weights import ExpertWeights, LayerWeights, ModelWeights

ALIGN = 256
_DTYPES = {"bfloat16": torch.bfloat16, "float16": torch.float16, "float32": torch.float32}


def pad_to_aligned_offset(offset: int, alignment: int = ALIGN) -> int:
    return ((offset + alignment - 1) // alignment) * alignment


def make_param_filename(base: str, tp_rank: int, tp_size: int) -> str:
    return os.path.join(base, f"params_rank_{tp_rank}_of_{tp_size}.pt")


def make_metadata_filename(base: str, tp_rank: int, tp_size: int) -> str:
    return os.path.join(base, f"metadata_rank_{tp_rank}_of_{tp_size}.json")


def make_model_config_filename(base: str) -> str:
    return os.path.join(base, "ds_model_config.pkl")


def flatten_model_weights(w: ModelWeights) -> Tuple[torch.Tensor, Dict]:
    tensors = w.named_tensors()
    meta, offset = {}, 0
    seen = {}
    for name, t in tensors.items():
        key = (t.untyped_storage().data_ptr(), t.storage_offset(), tuple(t.shape))
        if key in seen:  # tied weights (lm_head is embed): store once
            meta[name] = dict(meta[seen[key]], alias_of=seen[key])
            continue
        seen[key] = name
        offset = pad_to_aligned_offset(offset)
        nbytes = t.numel() * t.element_size()
        meta[name] = {"offset": offset, "shape": list(t.shape), "dtype": str(t.dtype).replace("torch.", ""),
                      "nbytes": nbytes}
        offset += nbytes
    flat = torch.zeros(pad_to_aligned_offset(offset), dtype=torch.uint8, device=next(iter(tensors.values())).device)
    for name, m in meta.items():
        if "alias_of" in m:
            continue
        t = tensors[name].contiguous()
        flat[m["offset"]:m["offset"] + m["nbytes"]] = t.view(-1).view(torch.uint8)
    return flat, {"params": meta, "n_layers": len(w.layers)}


def restore_model_weights(flat: torch.Tensor, meta: Dict, device: torch.device) -> ModelWeights:
    """Views into ONE device buffer (the flat tensor is moved once; parameters are zero-copy slices)."""
    flat = flat.to(device)
    params = meta["params"]
    w = ModelWeights(layers=[LayerWeights() for _ in range(meta["n_layers"])])

    def view(m):
        dt = _DTYPES[m["dtype"]]
        return flat[m["offset"]:m["offset"] + m["nbytes"]].view(dt).view(m["shape"])

    experts: Dict[Tuple[int, int], Dict[str, torch.Tensor]] = {}
    shared: Dict[int, Dict[str, torch.Tensor]] = {}
    for name, m in params.items():
        t = view(m)
        parts = name.split(".")
        if parts[0] != "layers":
            setattr(w, name, t)
            continue
        li = int(parts[1])
        if parts[2] == "experts":
            experts.setdefault((li, int(parts[3])), {})[parts[4]] = t
        elif parts[2] == "shared_expert":
            shared.setdefault(li, {})[parts[3]] = t
        else:
            setattr(w.layers[li], parts[2], t)
    for (li, e), d in sorted(experts.items()):
        w.layers[li].experts.append(ExpertWeights(d["w_up"], d["w_down"]))
    for li, d in shared.items():
        w.layers[li].shared_expert = ExpertWeights(d["w_up"], d["w_down"])
    return w
