"""Small helpers of the autotuner (``deepspeed/autotuning/utils.py``)."""
from __future__ import annotations

import copy
import itertools
import json
import os
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


# ---- template substitution, config merging / pruning, experiment files (ref: autotuning/utils.py:41-130, 181-290, 387-415) ----
def find_replace_str(value, replace_dict: Dict):
    """Substitute ``$name`` placeholders of a string from ``replace_dict`` (case-insensitive names; ``$nvme_path`` is left
    for the runtime). A placeholder whose value is not a string replaces the WHOLE value (it must then stand alone)."""
    if not isinstance(value, str):
        return str(value)
    found = re.findall(r"\$[\w]+", value)
    for var in found:
        key = var[1:].lower()
        if key == "nvme_path":
            continue
        assert key in replace_dict, f"unknown var key: {key}, in {replace_dict}"
        if isinstance(replace_dict[key], str):
            value = value.replace(var, replace_dict[key])
        else:
            assert len(found) == 1, "unable to replace multiple non-string matches"
            value = replace_dict[key]
    return value


def find_replace(target, replace_dict: Dict) -> None:
    """In-place placeholder substitution through nested dicts / lists."""
    if isinstance(target, dict):
        for k, v in target.items():
            if isinstance(v, str):
                target[k] = find_replace_str(v, replace_dict)
            elif isinstance(v, list):
                for i in range(len(v)):
                    v[i] = find_replace_str(v[i], replace_dict)
            elif isinstance(v, dict):
                find_replace(v, replace_dict)
    elif isinstance(target, list):
        for i in range(len(target)):
            target[i] = str(find_replace_str(target[i], replace_dict))


def get_list(val) -> list:
    return val if isinstance(val, list) else [val]


def combine_dict(d: Dict, u: Dict) -> Dict:
    """Merge ``u`` into ``d``: nested dicts merge, a key present in both becomes the list of distinct values (a tuning axis)."""
    for k, v in u.items():
        if isinstance(v, dict):
            d[k] = combine_dict(d.get(k, {}), v)
        elif k not in d:
            d[k] = v
        else:
            if not isinstance(d[k], list):
                d[k] = [d[k]]
            d[k].extend(x for x in get_list(v) if x not in d[k])
    return d


def del_if_exists(t: str, d: Dict) -> None:
    """Delete key ``t`` at the first level it occurs on each branch of a nested dict."""
    if t in d:
        del d[t]
        return
    for v in d.values():
        if isinstance(v, dict):
            del_if_exists(t, v)


def fetch_hostfile(hostfile_path: str):
    """``{host: slots}`` of a hostfile (``None`` when the file does not exist)."""
    from ..launcher.runner import fetch_hostfile as _fetch
    return _fetch(hostfile_path)


def validate_ds_config(config: Dict) -> bool:
    """Screens out configs that cannot run: ZeRO-2 with both legacy offload switches, ZeRO-3 offload without a
    DeepSpeed-built optimizer."""
    zero = config.get("zero_optimization") or {}
    stage = zero.get("stage")
    if not zero or stage not in (2, 3):
        return True
    if stage == 2:
        return not (zero.get("cpu_offload") and zero.get("cpu_offload_params"))
    offload = any((zero.get(k) or {}).get("device") in ("cpu", "nvme") for k in ("offload_optimizer", "offload_param"))
    return not (offload and not config.get("optimizer"))


def remove_dupe_dicts(items: List[Dict]) -> List[Dict]:
    """Distinct (nested) dicts of a list, first occurrences kept in order."""
    seen, out = set(), []
    for d in items:
        key = json.dumps(d, sort_keys=True)
        if key not in seen:
            seen.add(key)
            out.append(json.loads(key))
    return out


def prune_config(config: Dict, ignored_keys=()) -> None:
    """Delete the sections named in ``ignored_keys`` wherever they first occur (in place)."""
    for k in ignored_keys or ():
        del_if_exists(k, config)


def get_tuning_keys(tuning_space: Dict) -> List[str]:
    """Names of the entries that offer more than one value."""
    keys = []
    for k, v in tuning_space.items():
        if isinstance(v, dict):
            keys.extend(get_tuning_keys(v))
        elif isinstance(v, list) and len(v) > 1:
            keys.append(k)
    return keys


def get_first_config(config: Dict):
    """The config obtained by taking the first value of every list (the ``optimizer`` section is kept as given)."""
    if not config:
        return None
    cfg = copy.deepcopy(config)
    for k, v in cfg.items():
        if isinstance(v, dict) and k != "optimizer":
            cfg[k] = get_first_config(v)
        elif isinstance(v, list) and v:
            cfg[k] = v[0]
    return cfg


def write_experiments(exps: List[Dict], exps_dir: str) -> List[str]:
    """One ``<name>.json`` per experiment description; returns the paths."""
    os.makedirs(exps_dir, exist_ok=True)
    paths = []
    for exp in exps:
        path = os.path.join(exps_dir, f"{exp['name']}.json")
        with open(path, "w") as f:
            json.dump(exp, f)
        paths.append(path)
    return paths
