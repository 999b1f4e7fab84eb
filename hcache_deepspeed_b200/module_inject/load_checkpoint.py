"""Loading checkpoints into a (possibly meta-device) module before injection, and writing / reading the per-rank
"mp checkpoint" of an already tensor-sliced model.

Reference: ``deepspeed/module_inject/load_checkpoint.py:22-285`` (``load_model_with_checkpoint``: walks the injected
module tree and copies / slices tensors by hand, materialising meta tensors), ``inference/engine.py:366-475``
(``_load_checkpoint``: checkpoint json ``{"type", "checkpoints", "version"}``, DeepSpeed ``mp_rank_*`` folders with a
``latest`` tag, Megatron merge / split through ``SDLoaderFactory``) and ``replace_module.py:384-478`` (saving
``save_mp_checkpoint_path``). Here checkpoints are applied to the ORIGINAL module (``load_state_dict(assign=True)``
materialises meta parameters) and injection runs afterwards, so no per-container load code is needed.
"""
from __future__ import annotations

import json
import os
import re
from typing import Dict, Iterable, List, Optional, Union

import torch
from torch import nn

from ..utils.logging import logger


def _read(path: str, map_location="cpu") -> Dict[str, torch.Tensor]:
    if path.endswith(".safetensors"):
        from safetensors.torch import load_file
        return load_file(path, device=str(map_location))
    return torch.load(path, map_location=map_location, weights_only=False)


def choose_module_key(sd: dict) -> Optional[str]:
    assert not ("module" in sd and "model" in sd), "checkpoint has both 'model' and 'module' keys, not sure how to proceed"
    return "module" if "module" in sd else "model" if "model" in sd else None


def load_model_with_checkpoint(module: nn.Module, sd: Dict[str, torch.Tensor], strict: bool = False,
                               prefix_candidates: Iterable[str] = ("", "module.", "model.", "transformer.")) -> List[str]:
    """Apply one (partial) state dict. Keys are matched after trying to strip / add the usual wrappers' prefixes; meta
    parameters are replaced by the loaded tensors. Returns the keys that found no home."""
    own = dict(module.state_dict(keep_vars=True))
    remapped, unused = {}, []
    for k, v in sd.items():
        if not torch.is_tensor(v):
            continue
        hit = None
        for p in prefix_candidates:
            for cand in ((k[len(p):] if p and k.startswith(p) else None), p + k):
                if cand is not None and cand in own:
                    hit = cand
                    break
            if hit:
                break
        if hit is None:
            unused.append(k)
        else:
            remapped[hit] = v
    missing, unexpected = module.load_state_dict(remapped, strict=False, assign=True)
    if strict and (unused or [m for m in missing if not own[m].is_meta]):
        raise RuntimeError(f"checkpoint mismatch: unused keys {unused[:5]}, missing {list(missing)[:5]}")
    return unused


def tie_meta_lm_head(module: nn.Module) -> None:
    """A tied output embedding that is still on the meta device takes the (now loaded) input embedding."""
    emb = None
    for n, p in module.named_parameters():
        if re.search(r"(word_embeddings|embed_tokens|wte)\.", n):
            emb = p
    head = getattr(module, "lm_head", None)
    if emb is not None and head is not None and getattr(head, "weight", None) is not None and head.weight.is_meta:
        head.weight = emb


def resolve_checkpoint_files(checkpoint: Union[str, dict], base_dir: str = "", tag: Optional[str] = None):
    """-> (kind, files): ``"list"`` = plain state-dict files applied in order; ``"megatron"`` = an SD loader that merges or
    splits ``mp_rank_*`` files for the current tensor-parallel size."""
    from ..runtime.state_dict_factory import SDLoaderFactory
    if isinstance(checkpoint, str) and os.path.isdir(checkpoint):
        if tag is None and os.path.isfile(os.path.join(checkpoint, "latest")):
            tag = open(os.path.join(checkpoint, "latest")).read().strip()
        folder = os.path.join(checkpoint, tag) if tag else checkpoint
        mp = sorted(f for f in os.listdir(folder) if re.fullmatch(r"mp_rank_\d+_model_states\.pt", f))
        if mp:
            return "megatron", SDLoaderFactory.get_sd_loader([os.path.join(folder, f) for f in mp])
        files = sorted(f for f in os.listdir(folder) if f.endswith((".safetensors", ".bin", ".pt")))
        return "list", [os.path.join(folder, f) for f in files]
    data = checkpoint if isinstance(checkpoint, dict) else json.load(open(checkpoint))
    base = data.get("base_dir", base_dir) or base_dir
    files = [os.path.join(base, f) if base and not os.path.isabs(f) else f for f in data["checkpoints"]] \
        if isinstance(data["checkpoints"], list) else data["checkpoints"]
    if str(data.get("type", "")).lower() == "megatron":
        return "megatron", SDLoaderFactory.get_sd_loader(files, sd_type="Megatron", version=data.get("version"))
    if isinstance(files, dict):          # the layout written by save_mp_checkpoint: {"non_tp": [...], "tp": [...]}
        files = list(files.get("non_tp", [])) + list(files.get("tp", []))
    return "list", files


def load_checkpoint_into(module: nn.Module, checkpoint: Union[str, dict], base_dir: str = "", tp_size: int = 1, tp_rank: int = 0,
                         strict: bool = False, tag: Optional[str] = None) -> int:
    """Load ``checkpoint`` (folder / json path / json dict) into ``module``; returns the number of files applied."""
    kind, src = resolve_checkpoint_files(checkpoint, base_dir, tag)
    if kind == "megatron":
        _, sd, _ = src.load(tp_size, tp_rank)
        key = choose_module_key(sd)
        load_model_with_checkpoint(module, sd[key] if key else sd, strict)
        n = 1
    else:
        mine = [f for f in src if not re.search(r"tp_(\d+)_", os.path.basename(f)) or
                int(re.search(r"tp_(\d+)_", os.path.basename(f)).group(1)) == tp_rank]
        for i, f in enumerate(mine):
            logger.info(f"loading checkpoint ({i + 1}/{len(mine)}): {f}")
            sd = _read(f)
            key = choose_module_key(sd) if isinstance(sd, dict) and not all(torch.is_tensor(v) for v in sd.values()) else None
            load_model_with_checkpoint(module, sd[key] if key else sd, strict)
        n = len(mine)
    tie_meta_lm_head(module)
    if hasattr(module, "tie_weights"):      # assign=True replaced Parameter objects: re-establish HF weight tying
        try:
            module.tie_weights()
        except Exception as e:  # pragma: no cover - model-specific tie logic
            logger.warning(f"tie_weights() failed after checkpoint load: {e}")
    still_meta = [n_ for n_, p in module.named_parameters() if p.is_meta]
    if still_meta:
        raise RuntimeError(f"{len(still_meta)} parameters were not found in the checkpoint (e.g. {still_meta[:3]})")
    return n


def save_mp_checkpoint(module: nn.Module, path: str, tp_size: int = 1, tp_rank: int = 0, model_type: str = "ds_model",
                       dtype: Optional[torch.dtype] = None, version: float = 1.0) -> str:
    """Write this rank's (already sliced) weights as ``tp_<rank>_00.pt`` and, on rank 0, ``ds_inference_config.json``
    listing every rank's file — loading it back with the same ``tp_size`` needs no slicing."""
    os.makedirs(path, exist_ok=True)
    fname = f"tp_{tp_rank:0>2d}_00.pt"
    torch.save({k: v.detach().cpu() for k, v in module.state_dict().items()}, os.path.join(path, fname))
    cfg_path = os.path.join(path, "ds_inference_config.json")
    if tp_rank == 0:
        cfg = {"type": model_type, "base_dir": path, "checkpoints": {"non_tp": [], "tp": [f"tp_{r:0>2d}_00.pt" for r in range(tp_size)]},
               "version": version, "parallelization": "tp", "tp_size": tp_size,
               "dtype": str(dtype).replace("torch.", "") if dtype is not None else None}
        with open(cfg_path, "w") as f:
            json.dump(cfg, f)
    return cfg_path
