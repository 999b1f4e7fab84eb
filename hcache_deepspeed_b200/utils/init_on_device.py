"""``OnDevice``: construct a model directly on a device / as meta tensors with a given dtype
(``deepspeed/utils/init_on_device.py:12-80``)."""
from typing import Callable

import torch


class OnDevice:
    _orig_torch_empty, _orig_torch_zeros, _orig_torch_ones, _orig_torch_full = torch.empty, torch.zeros, torch.ones, torch.full

    def __init__(self, dtype, device: str = "meta", enabled: bool = True):
        self.dtype, self.enabled, self.device = dtype, enabled, device

    def fp_tensor_constructor(self, fn: Callable, target_fp_dtype: torch.dtype) -> Callable:
        def wrapped(*args, **kwargs):
            if kwargs.get("device", None) is None:
                kwargs["device"] = self.device
            t = fn(*args, **kwargs)
            return t.to(target_fp_dtype) if t.is_floating_point() else t
        return wrapped

    def __enter__(self):
        if not self.enabled:
            return
        torch.empty = self.fp_tensor_constructor(self._orig_torch_empty, self.dtype)
        torch.zeros = self.fp_tensor_constructor(self._orig_torch_zeros, self.dtype)
        torch.ones = self.fp_tensor_constructor(self._orig_torch_ones, self.dtype)
        torch.full = self.fp_tensor_constructor(self._orig_torch_full, self.dtype)

    def __exit__(self, exc_type, exc_value, traceback):
        if not self.enabled:
            return
        torch.empty, torch.zeros, torch.ones, torch.full = (self._orig_torch_empty, self._orig_torch_zeros,
                                                           self._orig_torch_ones, self._orig_torch_full)
