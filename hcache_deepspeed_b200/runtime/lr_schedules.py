"""Learning-rate schedules: LRRangeTest, OneCycle, WarmupLR, WarmupDecayLR, WarmupCosineLR.

Same names, parameters and curves as the reference (``deepspeed/runtime/lr_schedules.py``: LRRangeTest :267,
OneCycle :370, WarmupLR :634, WarmupDecayLR :723, WarmupCosineLR :774); each schedule is a pure function of
the step, evaluated per param group, so ``state_dict`` is only ``last_batch_iteration``.
"""
from __future__ import annotations

import argparse
import math
from typing import List, Union

from torch.optim import Optimizer

LR_SCHEDULE = "lr_schedule"
LR_RANGE_TEST, ONE_CYCLE, WARMUP_LR, WARMUP_DECAY_LR, WARMUP_COSINE_LR = ("LRRangeTest", "OneCycle", "WarmupLR",
                                                                         "WarmupDecayLR", "WarmupCosineLR")
VALID_LR_SCHEDULES = [LR_RANGE_TEST, ONE_CYCLE, WARMUP_LR, WARMUP_DECAY_LR, WARMUP_COSINE_LR]
WARMUP_LOG_RATE, WARMUP_LINEAR_RATE = "log", "linear"


def _as_list(optimizer: Optimizer, value, name: str) -> List[float]:
    n = len(optimizer.param_groups)
    if isinstance(value, (list, tuple)):
        if len(value) != n:
            raise ValueError(f"Expected {n} values for {name}, got {len(value)}")
        return list(value)
    return [value] * n


def get_torch_optimizer(optimizer):
    if isinstance(optimizer, Optimizer):
        return optimizer
    if hasattr(optimizer, "optimizer") and isinstance(optimizer.optimizer, Optimizer):
        return optimizer.optimizer
    raise TypeError(f"{type(optimizer).__name__} is not a subclass of torch.optim.Optimizer")


class _Schedule:
    def __init__(self, optimizer, last_batch_iteration: int = -1):
        self.optimizer = get_torch_optimizer(optimizer)
        self.last_batch_iteration = last_batch_iteration
        self._last_lr = [g["lr"] for g in self.optimizer.param_groups]

    def get_lr(self) -> List[float]:
        raise NotImplementedError

    def get_last_lr(self) -> List[float]:
        return self._last_lr

    def step(self, last_batch_iteration: int = None):
        if last_batch_iteration is None:
            last_batch_iteration = self.last_batch_iteration + 1
        self.last_batch_iteration = last_batch_iteration
        lrs = self.get_lr()
        for group, lr in zip(self.optimizer.param_groups, lrs):
            group["lr"] = lr
        self._last_lr = [g["lr"] for g in self.optimizer.param_groups]

    def state_dict(self):
        return {"last_batch_iteration": self.last_batch_iteration}

    def load_state_dict(self, sd):
        self.last_batch_iteration = sd["last_batch_iteration"]


class LRRangeTest(_Schedule):
    """lr = min_lr * (1 + step_rate * interval(step)); interval is continuous or a staircase of ``step_size``."""

    def __init__(self, optimizer, lr_range_test_min_lr: Union[float, list] = 1e-3, lr_range_test_step_size: int = 2000,
                 lr_range_test_step_rate: float = 1.0, lr_range_test_staircase: bool = False,
                 last_batch_iteration: int = -1):
        super().__init__(optimizer, last_batch_iteration)
        self.min_lr = _as_list(self.optimizer, lr_range_test_min_lr, "lr_range_test_min_lr")
        self.step_size, self.step_rate, self.staircase = lr_range_test_step_size, lr_range_test_step_rate, lr_range_test_staircase
        if last_batch_iteration == -1:
            for g, lr in zip(self.optimizer.param_groups, self.min_lr):
                g["lr"] = lr

    def get_lr(self):
        x = float(self.last_batch_iteration + 1) / self.step_size
        interval = math.floor(x) if self.staircase else x
        return [lr * (1 + self.step_rate * interval) for lr in self.min_lr]


class OneCycle(_Schedule):
    """Two-phase cycle (min -> max -> min) followed by a decay phase; momentum cycles inversely when enabled."""

    def __init__(self, optimizer, cycle_min_lr, cycle_max_lr, decay_lr_rate: float = 0.0, cycle_first_step_size: int = 2000,
                 cycle_second_step_size: int = None, cycle_first_stair_count: int = 0, cycle_second_stair_count: int = None,
                 decay_step_size: int = 0, cycle_momentum: bool = True, cycle_min_mom: float = 0.8,
                 cycle_max_mom: float = 0.9, decay_mom_rate: float = 0.0, last_batch_iteration: int = -1):
        super().__init__(optimizer, last_batch_iteration)
        first = float(cycle_first_step_size)
        second = float(cycle_second_step_size) if cycle_second_step_size is not None else first
        self.total_size = first + second
        self.step_ratio = first / self.total_size
        self.first_stair_count = cycle_first_stair_count
        self.second_stair_count = cycle_first_stair_count if cycle_second_stair_count is None else cycle_second_stair_count
        self.decay_step_size = decay_step_size
        self.min_lrs = _as_list(self.optimizer, cycle_min_lr, "cycle_min_lr")
        self.max_lrs = _as_list(self.optimizer, cycle_max_lr, "cycle_max_lr")
        self.decay_lr_rate = decay_lr_rate
        self.cycle_momentum = cycle_momentum
        self.min_moms = [(cycle_min_mom, 0.99)] * len(self.optimizer.param_groups)
        self.max_moms = [(cycle_max_mom, 0.99)] * len(self.optimizer.param_groups)
        self.decay_mom_rate = decay_mom_rate
        if cycle_momentum and "betas" not in self.optimizer.defaults and "momentum" not in self.optimizer.defaults:
            self.cycle_momentum = False
        if last_batch_iteration == -1:
            for g, lr in zip(self.optimizer.param_groups, self.min_lrs):
                g["lr"] = lr

    def _scale(self):
        it = self.last_batch_iteration + 1
        cycle = math.floor(1 + it / self.total_size)
        x = 1.0 + it / self.total_size - cycle
        return x / self.step_ratio if x <= self.step_ratio else (x - 1) / (self.step_ratio - 1)

    def get_lr(self):
        it = self.last_batch_iteration
        if it < self.total_size:
            s = self._scale()
            lrs = [lo + (hi - lo) * s for lo, hi in zip(self.min_lrs, self.max_lrs)]
            if self.cycle_momentum:
                self._set_momentum([(hi[0] - (hi[0] - lo[0]) * s, hi[1]) for lo, hi in zip(self.min_moms, self.max_moms)])
            return lrs
        decay_it = it - self.total_size + 1
        interval = decay_it / self.decay_step_size if self.decay_step_size else 0.0
        if self.cycle_momentum:
            self._set_momentum([(hi[0] * (1 + self.decay_mom_rate * interval), hi[1]) for hi in self.max_moms])
        return [lo / (1 + self.decay_lr_rate * interval) for lo in self.min_lrs]

    def _set_momentum(self, moms):
        for g, m in zip(self.optimizer.param_groups, moms):
            if "betas" in g:
                g["betas"] = (m[0], g["betas"][1])
            elif "momentum" in g:
                g["momentum"] = m[0]

    def get_mom(self):
        return [g.get("betas", (g.get("momentum"), ))[0] for g in self.optimizer.param_groups]


class WarmupLR(_Schedule):
    """Increase lr from ``warmup_min_lr`` to ``warmup_max_lr`` over ``warmup_num_steps`` (log or linear), then hold."""

    def __init__(self, optimizer, warmup_min_lr: Union[float, list] = 0.0, warmup_max_lr: Union[float, list] = 0.001,
                 warmup_num_steps: int = 1000, warmup_type: str = WARMUP_LOG_RATE, last_batch_iteration: int = -1):
        super().__init__(optimizer, last_batch_iteration)
        self.min_lrs = _as_list(self.optimizer, warmup_min_lr, "min_lr")
        self.max_lrs = _as_list(self.optimizer, warmup_max_lr, "max_lr")
        self.delta_lrs = [hi - lo for hi, lo in zip(self.max_lrs, self.min_lrs)]
        self.warmup_num_steps = max(2, warmup_num_steps)
        if warmup_type not in (WARMUP_LOG_RATE, WARMUP_LINEAR_RATE):
            warmup_type = WARMUP_LOG_RATE
        self.warmup_type = warmup_type
        self.inverse_log_warm_up = 1.0 / math.log(self.warmup_num_steps)
        if last_batch_iteration == -1:
            self._last_lr = self.get_lr_at(0)
            for g, lr in zip(self.optimizer.param_groups, self._last_lr):
                g["lr"] = lr

    def _gamma(self, it: int) -> float:
        if it < self.warmup_num_steps:
            if self.warmup_type == WARMUP_LOG_RATE:
                return self.inverse_log_warm_up * math.log(it + 1)
            return it / self.warmup_num_steps
        return 1.0

    def get_lr_at(self, it: int):
        g = self._gamma(it)
        return [lo + d * g for lo, d in zip(self.min_lrs, self.delta_lrs)]

    def get_lr(self):
        if self.last_batch_iteration < 0:
            return [0.0] * len(self.min_lrs)
        return self.get_lr_at(self.last_batch_iteration)


class WarmupDecayLR(WarmupLR):
    """WarmupLR followed by a linear decay to zero at ``total_num_steps``."""

    def __init__(self, optimizer, total_num_steps: int, warmup_min_lr=0.0, warmup_max_lr=0.001, warmup_num_steps: int = 1000,
                 warmup_type: str = WARMUP_LOG_RATE, last_batch_iteration: int = -1):
        self.total_num_steps = total_num_steps
        super().__init__(optimizer, warmup_min_lr, warmup_max_lr, warmup_num_steps, warmup_type, last_batch_iteration)

    def _gamma(self, it: int) -> float:
        if it < self.warmup_num_steps:
            return super()._gamma(it)
        return max(0.0, float(self.total_num_steps - it) / float(max(1.0, self.total_num_steps - self.warmup_num_steps)))


class WarmupCosineLR(_Schedule):
    """Linear/log warm-up of a ratio from ``warmup_min_ratio`` to 1, then cosine decay to ``cos_min_ratio``; applied to
    each group's initial lr."""

    def __init__(self, optimizer, total_num_steps: int, warmup_min_ratio: float = 0.0, warmup_num_steps: int = 1000,
                 cos_min_ratio: float = 0.0001, warmup_type: str = WARMUP_LOG_RATE, last_batch_iteration: int = -1):
        super().__init__(optimizer, last_batch_iteration)
        self.total_num_steps = total_num_steps
        self.warmup_min_ratio, self.cos_min_ratio = warmup_min_ratio, cos_min_ratio
        self.warmup_num_steps = max(2, warmup_num_steps)
        self.warmup_type = warmup_type if warmup_type in (WARMUP_LOG_RATE, WARMUP_LINEAR_RATE) else WARMUP_LOG_RATE
        self.inverse_log_warm_up = 1.0 / math.log(self.warmup_num_steps)
        self.org_lrs = [g["lr"] for g in self.optimizer.param_groups]

    def get_lr_ratio(self) -> float:
        it = self.last_batch_iteration
        if it < 0:
            return 0.0
        if it < self.warmup_num_steps:
            r = self.inverse_log_warm_up * math.log(it + 1) if self.warmup_type == WARMUP_LOG_RATE else it / self.warmup_num_steps
            return self.warmup_min_ratio + (1.0 - self.warmup_min_ratio) * r
        frac = min(1.0, (it - self.warmup_num_steps + 1) / max(1, self.total_num_steps - self.warmup_num_steps))
        cos = (1 + math.cos(math.pi * frac)) / 2
        return max(0.0, self.cos_min_ratio + (1.0 - self.cos_min_ratio) * cos)

    def get_lr(self):
        r = self.get_lr_ratio()
        return [lr * r for lr in self.org_lrs]


SCHEDULES = {LR_RANGE_TEST: LRRangeTest, ONE_CYCLE: OneCycle, WARMUP_LR: WarmupLR, WARMUP_DECAY_LR: WarmupDecayLR,
             WARMUP_COSINE_LR: WarmupCosineLR}


def get_lr_schedule_class(name: str):
    if name in SCHEDULES:
        return SCHEDULES[name]
    import torch.optim.lr_scheduler as tls
    if hasattr(tls, name):
        return getattr(tls, name)
    raise ValueError(f"{name} is not a supported LR schedule (valid: {VALID_LR_SCHEDULES} or torch.optim.lr_scheduler.*)")


def add_tuning_arguments(parser: argparse.ArgumentParser):
    """Command-line arguments for the convergence-tuning schedules (reference ``lr_schedules.py:54-168``)."""
    g = parser.add_argument_group("Convergence Tuning", "Convergence tuning configurations")
    g.add_argument("--lr_schedule", type=str, default=None, help="LR schedule for training.")
    g.add_argument("--lr_range_test_min_lr", type=float, default=0.001)
    g.add_argument("--lr_range_test_step_rate", type=float, default=1.0)
    g.add_argument("--lr_range_test_step_size", type=int, default=1000)
    g.add_argument("--lr_range_test_staircase", type=bool, default=False)
    g.add_argument("--cycle_first_step_size", type=int, default=1000)
    g.add_argument("--cycle_first_stair_count", type=int, default=-1)
    g.add_argument("--cycle_second_step_size", type=int, default=-1)
    g.add_argument("--cycle_second_stair_count", type=int, default=-1)
    g.add_argument("--decay_step_size", type=int, default=1000)
    g.add_argument("--cycle_min_lr", type=float, default=0.01)
    g.add_argument("--cycle_max_lr", type=float, default=0.1)
    g.add_argument("--decay_lr_rate", type=float, default=0.0)
    g.add_argument("--cycle_momentum", default=False, action="store_true")
    g.add_argument("--cycle_min_mom", type=float, default=0.8)
    g.add_argument("--cycle_max_mom", type=float, default=0.9)
    g.add_argument("--decay_mom_rate", type=float, default=0.0)
    g.add_argument("--warmup_min_lr", type=float, default=0)
    g.add_argument("--warmup_max_lr", type=float, default=0.001)
    g.add_argument("--warmup_num_steps", type=int, default=1000)
    g.add_argument("--warmup_type", type=str, default=WARMUP_LOG_RATE)
    return parser


# ---- schedule configs from command-line arguments (ref: lr_schedules.py:124-253) --------------------------------------------
_ARG_NAMES = {
    LR_RANGE_TEST: ("lr_range_test_min_lr", "lr_range_test_step_rate", "lr_range_test_step_size", "lr_range_test_staircase"),
    ONE_CYCLE: ("cycle_first_step_size", "cycle_first_stair_count", "cycle_second_step_size", "cycle_second_stair_count",
                "decay_step_size", "cycle_min_lr", "cycle_max_lr", "decay_lr_rate", "cycle_min_mom", "cycle_max_mom",
                "decay_mom_rate"),
    WARMUP_LR: ("warmup_min_lr", "warmup_max_lr", "warmup_num_steps", "warmup_type"),
}


def parse_arguments():
    """(namespace of the tuning arguments, the arguments it did not recognise) from ``sys.argv``."""
    parser = add_tuning_arguments(argparse.ArgumentParser())
    return parser.parse_known_args()


def _override(args, params: dict, names) -> None:
    for name in names:
        value = getattr(args, name, None)
        if value is not None:
            params[name] = value


def override_lr_range_test_params(args, params: dict) -> None:
    _override(args, params, _ARG_NAMES[LR_RANGE_TEST])


def override_1cycle_params(args, params: dict) -> None:
    _override(args, params, _ARG_NAMES[ONE_CYCLE])


def override_warmupLR_params(args, params: dict) -> None:
    _override(args, params, _ARG_NAMES[WARMUP_LR])


def override_params(args, params: dict) -> None:
    """Every tuning argument present on ``args`` replaces the same-named entry of ``params``."""
    for names in _ARG_NAMES.values():
        _override(args, params, names)


def get_config_from_args(args):
    """(``{"type": ..., "params": {...}}``, None) for ``--lr_schedule``, or (None, reason)."""
    name = getattr(args, LR_SCHEDULE, None)
    if name is None:
        return None, f"--{LR_SCHEDULE} not specified on command line"
    if name not in VALID_LR_SCHEDULES:
        return None, f"{name} is not supported LR schedule"
    config = {"type": name, "params": {}}
    _override(args, config["params"], _ARG_NAMES.get(name, _ARG_NAMES[WARMUP_LR]))
    return config, None


def get_lr_from_config(config):
    """The learning rate a schedule config starts from / peaks at: (lr, "") or (None, reason)."""
    if "type" not in config:
        return None, "LR schedule type not defined in config"
    if "params" not in config:
        return None, "LR schedule params not defined in config"
    name, params = config["type"], config["params"]
    if name not in VALID_LR_SCHEDULES:
        return None, f"{name} is not a valid LR schedule"
    if name == LR_RANGE_TEST:
        return params["lr_range_test_min_lr"], ""
    if name == ONE_CYCLE:
        return params["cycle_max_lr"], ""
    return params["warmup_max_lr"], ""


def update_lr(param_groups, lrs):
    for group, lr in zip(param_groups, lrs):
        group["lr"] = lr
    return [group["lr"] for group in param_groups]
