"""Static and dynamic loss scaling (reference ``deepspeed/runtime/fp16/loss_scaler.py``: LossScaler :72,
DynamicLossScaler :99). Dynamic policy: on overflow halve the scale (after ``delayed_shift`` hysteresis credits are
used up), after ``scale_window`` clean steps double it."""
from __future__ import annotations

import torch

INITIAL_LOSS_SCALE, SCALE_WINDOW, DELAYED_SHIFT, CONSECUTIVE_HYSTERESIS, MIN_LOSS_SCALE = (
    "init_scale", "scale_window", "delayed_shift", "consecutive_hysteresis", "min_scale")


class LossScalerBase:
    def __init__(self, cur_scale: float):
        self.cur_scale = cur_scale
        self.dynamic = False

    @property
    def loss_scale(self) -> float:
        return self.cur_scale

    def scale_gradient(self, module, grad_in, grad_out):
        return tuple(self.loss_scale * g for g in grad_in)

    def update_scale(self, overflow: bool):
        pass

    def backward(self, loss: torch.Tensor, retain_graph: bool = False):
        (loss * self.loss_scale).backward(retain_graph=retain_graph)


class LossScaler(LossScalerBase):
    """Static scale."""

    def __init__(self, scale: float = 1.0):
        super().__init__(scale)

    def has_overflow(self, params) -> bool:
        return False


class DynamicLossScaler(LossScalerBase):
    def __init__(self, init_scale: float = 2 ** 32, scale_factor: float = 2.0, scale_window: int = 1000,
                 min_scale: float = 1.0, delayed_shift: int = 1, consecutive_hysteresis: bool = False,
                 raise_error_at_min_scale: bool = True, dtype=torch.half):
        super().__init__(init_scale)
        self.cur_iter = 0
        self.last_overflow_iter = -1
        self.scale_factor, self.scale_window, self.min_scale = scale_factor, scale_window, min_scale
        self.delayed_shift = self.cur_hysteresis = delayed_shift
        self.consecutive_hysteresis = consecutive_hysteresis
        self.raise_error_at_min_scale = raise_error_at_min_scale
        self.dynamic = True
        self.dtype = dtype

    @staticmethod
    def _has_inf_or_nan(x: torch.Tensor) -> bool:
        s = float(x.float().sum())
        return s in (float("inf"), -float("inf")) or s != s

    def has_overflow_serial(self, params) -> bool:
        return any(p.grad is not None and self._has_inf_or_nan(p.grad.data) for p in params)

    has_overflow = has_overflow_serial

    def update_scale(self, overflow: bool):
        if overflow:
            if self.delayed_shift == 1 or self.cur_hysteresis == 1:
                if self.cur_scale == self.min_scale and self.raise_error_at_min_scale:
                    raise Exception("Current loss scale already at minimum - cannot decrease scale anymore. Exiting run.")
                self.cur_scale = max(self.cur_scale / self.scale_factor, self.min_scale)
            else:
                self.cur_hysteresis -= 1
            self.last_overflow_iter = self.cur_iter
        else:
            if self.consecutive_hysteresis:
                self.cur_hysteresis = self.delayed_shift
            if (self.cur_iter - self.last_overflow_iter) % self.scale_window == 0:
                if not self.consecutive_hysteresis:
                    self.cur_hysteresis = self.delayed_shift
                self.cur_scale *= self.scale_factor
        self.cur_iter += 1

    def state_dict(self):
        return {k: getattr(self, k) for k in ("cur_scale", "cur_iter", "last_overflow_iter", "cur_hysteresis")}

    def load_state_dict(self, sd):
        for k, v in sd.items():
            setattr(self, k, v)


def CreateLossScaler(dtype, static_loss_scale, dynamic_scaling, dynamic_loss_args):
    if dtype == torch.half and dynamic_scaling:
        return DynamicLossScaler(dtype=dtype, **(dynamic_loss_args or {}))
    return LossScaler(scale=static_loss_scale if dtype == torch.half else 1.0)


def to_python_float(t):
    return t.item() if hasattr(t, "item") else t[0]
