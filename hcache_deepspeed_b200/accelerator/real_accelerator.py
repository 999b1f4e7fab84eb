"""``get_accelerator()``: CUDA when a GPU is visible, otherwise the CPU stand-in; ``DS_ACCELERATOR=cpu|cuda`` overrides.
(The reference probes eight vendors, ``deepspeed/accelerator/real_accelerator.py:51-260``; here that set is {cuda, cpu}.)"""
import os

ds_accelerator = None
SUPPORTED = ("cuda", "cpu")


def is_current_accelerator_supported():
    return get_accelerator()._name in SUPPORTED


def get_accelerator():
    global ds_accelerator
    if ds_accelerator is not None:
        return ds_accelerator
    import torch
    from .cuda_accelerator import CPU_Accelerator, CUDA_Accelerator
    name = os.environ.get("DS_ACCELERATOR")
    if name is not None and name not in SUPPORTED:
        raise ValueError(f"DS_ACCELERATOR must be one of {SUPPORTED}. Value '{name}' is not supported")
    if name is None:
        name = "cuda" if torch.cuda.is_available() else "cpu"
    ds_accelerator = CUDA_Accelerator() if name == "cuda" else CPU_Accelerator()
    return ds_accelerator


def set_accelerator(accel_obj):
    global ds_accelerator
    ds_accelerator = accel_obj
