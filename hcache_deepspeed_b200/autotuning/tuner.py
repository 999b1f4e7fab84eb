"""Search strategies over a list of experiment configs: exhaustive grid, random, and model-based (a ridge-regression
cost model over the numeric config features picks the next most promising untried config, with epsilon exploration).
Reference: ``deepspeed/autotuning/tuner/{base_tuner,index_based_tuner,model_based_tuner,cost_model}.py``."""
from __future__ import annotations

import random
from typing import Callable, Dict, List, Optional, Tuple

import numpy as np


def flatten(d: Dict, prefix: str = "") -> Dict[str, float]:
    out = {}
    for k, v in d.items():
        key = f"{prefix}{k}"
        if isinstance(v, dict):
            out.update(flatten(v, key + "."))
        elif isinstance(v, bool):
            out[key] = float(v)
        elif isinstance(v, (int, float)):
            out[key] = float(v)
    return out


class BaseTuner:

    def __init__(self, exps: List[Dict], run_fn: Callable[[Dict], Optional[float]], metric: str = "throughput"):
        self.all_exps = list(exps)
        self.run_fn = run_fn
        self.metric = metric
        self.best_iter, self.best_exp, self.best_metric_val = 0, None, None
        self.records: List[Tuple[Dict, Optional[float]]] = []

    def has_next(self) -> bool:
        return len(self.all_exps) > 0

    def next_batch(self, sample_size: int) -> List[Dict]:
        raise NotImplementedError

    def update(self):
        pass

    def better(self, a: float, b: Optional[float]) -> bool:
        if b is None:
            return True
        return a < b if self.metric == "latency" else a > b

    def tune(self, sample_size: int = 1, n_trials: int = 1000, early_stopping: Optional[int] = None) -> int:
        i = 0
        while i < n_trials and self.has_next():
            batch = self.next_batch(sample_size)
            for exp in batch:
                val = self.run_fn(exp)
                self.records.append((exp, val))
                if val is not None and self.better(val, self.best_metric_val):
                    self.best_exp, self.best_metric_val, self.best_iter = exp, val, i
                i += 1
            self.update()
            if early_stopping and i >= self.best_iter + early_stopping:
                break
        return i


class GridSearchTuner(BaseTuner):
    def next_batch(self, sample_size=1):
        batch, self.all_exps = self.all_exps[:sample_size], self.all_exps[sample_size:]
        return batch


class RandomTuner(BaseTuner):
    def __init__(self, exps, run_fn, metric="throughput", seed: int = 0):
        super().__init__(exps, run_fn, metric)
        self.rng = random.Random(seed)

    def next_batch(self, sample_size=1):
        n = min(sample_size, len(self.all_exps))
        picks = self.rng.sample(range(len(self.all_exps)), n)
        batch = [self.all_exps[i] for i in picks]
        self.all_exps = [e for i, e in enumerate(self.all_exps) if i not in set(picks)]
        return batch


class ModelBasedTuner(BaseTuner):

    def __init__(self, exps, run_fn, metric="throughput", seed: int = 0, random_exploration_ratio: float = 0.2, num_init: int = 2):
        super().__init__(exps, run_fn, metric)
        self.rng = random.Random(seed)
        self.ratio, self.num_init = random_exploration_ratio, num_init
        keys = sorted({k for e in exps for k in flatten(e)})
        self.keys = keys
        self.w = None

    def _x(self, exp):
        f = flatten(exp)
        return np.array([f.get(k, 0.0) for k in self.keys] + [1.0])

    def update(self):
        done = [(e, v) for e, v in self.records if v is not None]
        if len(done) < 2:
            return
        X = np.stack([self._x(e) for e, _ in done])
        y = np.array([v for _, v in done])
        mu, sd = X.mean(0), X.std(0) + 1e-9
        self._mu, self._sd = mu, sd
        Xn = (X - mu) / sd
        Xn[:, -1] = 1.0
        self.w = np.linalg.solve(Xn.T @ Xn + 1e-3 * np.eye(Xn.shape[1]), Xn.T @ y)

    def next_batch(self, sample_size=1):
        batch = []
        for _ in range(min(sample_size, len(self.all_exps))):
            if self.w is None or len(self.records) < self.num_init or self.rng.random() < self.ratio:
                i = self.rng.randrange(len(self.all_exps))
            else:
                preds = []
                for e in self.all_exps:
                    x = (self._x(e) - self._mu) / self._sd
                    x[-1] = 1.0
                    preds.append(float(x @ self.w))
                i = int(np.argmin(preds) if self.metric == "latency" else np.argmax(preds))
            batch.append(self.all_exps.pop(i))
        return batch


# `autotuning.tuner` is a package in the reference; its sub-module paths import through this module (see _import_paths.py)
__path__: list = []


# ---- index <-> feature helpers of the model-based tuner (ref: autotuning/tuner/utils.py) -----------------------------------
def index_to_feature(p: int, dims) -> List[int]:
    """A flat index into a grid of sizes ``dims`` -> the coordinate vector (first dimension fastest)."""
    out = []
    for d in dims:
        out.append(p % d)
        p //= d
    return out


def feature_to_index(feature, dims) -> int:
    p, stride = 0, 1
    for k, d in zip(feature, dims):
        p += stride * int(k)
        stride *= d
    return p


def dict_to_dims(tuning_space: Dict) -> List[int]:
    """Number of choices of every leaf of a tuning space, depth first."""
    dims = []
    for v in tuning_space.values():
        if isinstance(v, dict):
            dims.extend(dict_to_dims(v))
        else:
            dims.append(len(v) if isinstance(v, list) else 1)
    return dims


def gen_combinations(d: Dict):
    """Every assignment of a (nested) space of lists, as dicts, lazily."""
    import itertools
    keys = list(d.keys())
    choices = [list(gen_combinations(v)) if isinstance(v, dict) else (v if isinstance(v, list) else [v]) for v in d.values()]
    for comb in itertools.product(*choices):
        yield dict(zip(keys, comb))


def dict_to_feature(feature_dict: Dict, keys, max_value=None) -> List[float]:
    """Numeric feature vector of a config: the values of ``keys`` wherever they occur (booleans as 0/1, strings hashed to
    their position among the sorted distinct values seen); with ``max_value`` each entry is scaled into [0, 1]."""
    out = []
    for k, v in feature_dict.items():
        if isinstance(v, dict):
            out.extend(dict_to_feature(v, keys, None))
        elif k in keys:
            out.append(float(v) if isinstance(v, (int, float, bool)) else float(abs(hash(str(v))) % 1000))
    if max_value is not None:
        out = [x / m if m else x for x, m in zip(out, max_value)]
    return out


class XGBoostCostModel:
    """Name kept for configs that ask for the reference's cost model; the model-based tuner here fits a ridge regression
    on the flattened config features (no xgboost in this image), behind the same ``fit`` / ``predict`` pair."""

    def __init__(self, loss_type: str = "reg", num_threads=None, log_interval: int = 25, upper_model=None):
        self.w = None

    def fit(self, xs, ys):
        x = np.asarray(xs, dtype=np.float64)
        y = np.asarray(ys, dtype=np.float64)
        x1 = np.concatenate([x, np.ones((x.shape[0], 1))], axis=1)
        self.scale = max(float(np.abs(y).max()), 1e-9)      # targets are fitted in units of their largest magnitude
        self.w = np.linalg.solve(x1.T @ x1 + 1e-6 * np.eye(x1.shape[1]), x1.T @ (y / self.scale))

    def predict(self, xs):
        x = np.asarray(xs, dtype=np.float64)
        return (np.concatenate([x, np.ones((x.shape[0], 1))], axis=1) @ self.w) * self.scale
