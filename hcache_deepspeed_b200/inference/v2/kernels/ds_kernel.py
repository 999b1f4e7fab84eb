"""Base of the kernel wrapper objects (``deepspeed/inference/v2/kernels/ds_kernel.py:11-32``): construction validates
the static configuration (dtype, sizes); ``__call__`` runs on pre-allocated outputs."""
from __future__ import annotations

from abc import ABC, abstractmethod

import torch


class DSKernelBase(ABC):
    supported_dtypes = (torch.float16, torch.bfloat16, torch.float32)

    @abstractmethod
    def __init__(self, *args, **kwargs):
        raise NotImplementedError()

    @abstractmethod
    def __call__(self, *args, **kwargs):
        raise NotImplementedError()

    @staticmethod
    def _ops(t: torch.Tensor):
        """sm_100a kernels for CUDA tensors, the fp32 reference formulation for host tensors."""
        from ....ops.backend import get_ops
        return get_ops(t.device)

    @classmethod
    def _check_dtype(cls, dtype) -> None:
        if dtype not in cls.supported_dtypes:
            raise ValueError(f"Unsupported data type: {dtype}, supported data types are {list(cls.supported_dtypes)}")
