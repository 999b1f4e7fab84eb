"""Small helpers of the autotuner (``deepspeed/autotuning/utils.py``)."""
from __future__ import annotations

import copy
import itertools
import re
from typing import Dict, List, Optional


def search_error(filename: str) -> Optional[str]:
    """First line of a log that looks like an error, or ``None``."""
    try:
        with open(filename, errors="replace") as f:
            for line in f:
                for marker in ("Error", "error", "ERROR"):
                    i = line.find(marker)
                    if i != -1:
                        return line[i:].strip()
    except OSError:
        return "stderr.log does not exist"
    return None


def was_interruptted(filename: str) -> bool:
    try:
        with open(filename, errors="replace") as f:
            return any("KeyboardInterrupt" in line for line in f)
    except OSError:
        return False


def get_val_by_key(d: Dict, k: str):
    if k in d:
        return d[k]
    for v in d.values():
        if isinstance(v, dict):
            r = get_val_by_key(v, k)
            if r is not None:
                return r
    return None


def set_val_by_key(d: Dict, k: str, vv) -> None:
    if k in d:
        d[k] = vv
    for v in d.values():
        if isinstance(v, dict):
            set_val_by_key(v, k, vv)


def replace_dict(d: Dict, u: Dict, ignored_keys=()) -> Dict:
    """Overwrite ``d`` with ``u``, recursively; keys missing from ``u`` disappear unless ignored."""
    out = {k: v for k, v in d.items() if k in ignored_keys}
    for k, v in u.items():
        out[k] = replace_dict(d.get(k, {}), v, ignored_keys) if isinstance(v, dict) and isinstance(d.get(k), dict) else copy.deepcopy(v)
    return out


def get_all_configs(tuning_space: Dict, ignore_keys=None) -> List[Dict]:
    """Every combination of the list-valued leaves of a tuning space."""
    ignore_keys = set(ignore_keys or ())

    def leaves(d, prefix=()):
        for k, v in d.items():
            if isinstance(v, dict):
                yield from leaves(v, prefix + (k,))
            else:
                yield prefix + (k,), (v if isinstance(v, list) and k not in ignore_keys else [v])

    items = list(leaves(tuning_space))
    out = []
    for combo in itertools.product(*(vals for _, vals in items)):
        cfg: Dict = {}
        for (path, _), val in zip(items, combo):
            cur = cfg
            for part in path[:-1]:
                cur = cur.setdefault(part, {})
            cur[path[-1]] = val
        out.append(cfg)
    return out


def prune_configs(configs: List[Dict], ignored_keys=()) -> List[Dict]:
    """Drop duplicates (ignoring ``ignored_keys`` at any depth)."""
    def strip(d):
        return {k: strip(v) if isinstance(v, dict) else v for k, v in sorted(d.items()) if k not in ignored_keys}

    seen, out = set(), []
    for c in configs:
        key = repr(strip(c))
        if key not in seen:
            seen.add(key)
            out.append(c)
    return out


def canonical_name(config: Dict, tuning_keys=None, prefix: str = "", omit_val: bool = False) -> str:
    """Short experiment name from the tuned keys: ``z2_tmbspg8_gas1`` style."""
    abbrev = {"zero_optimization.stage": "z", "train_micro_batch_size_per_gpu": "tmbspg", "gradient_accumulation_steps": "gas"}
    parts = [prefix] if prefix else []

    def walk(d, path=""):
        for k, v in d.items():
            full = f"{path}.{k}" if path else k
            if isinstance(v, dict):
                walk(v, full)
            elif tuning_keys is None or k in tuning_keys or full in tuning_keys:
                short = abbrev.get(full) or "".join(w[0] for w in re.split(r"[_.]", k) if w)
                parts.append(short if omit_val else f"{short}{v}")

    walk(config)
    return "_".join(parts)


def memory_to_string(n: float, postfix: str = "", units=None, precision: int = 2) -> str:
    for unit, scale in (("T", 1 << 40), ("G", 1 << 30), ("M", 1 << 20), ("K", 1 << 10)):
        if (units is None and n >= scale) or units == unit:
            return f"{round(n / scale, precision)} {unit}{postfix}"
    return f"{n} {postfix}".strip()


def number_to_string(n: float, postfix: str = "", units=None, precision: int = 2) -> str:
    for unit, scale in (("B", 1e9), ("M", 1e6), ("K", 1e3)):
        if (units is None and n >= scale) or units == unit:
            return f"{round(n / scale, precision)} {unit}{postfix}"
    return f"{n} {postfix}".strip()
