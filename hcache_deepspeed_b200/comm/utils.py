"""Launcher-environment and call-introspection helpers of the comm layer (``deepspeed/comm/utils.py``)."""
from __future__ import annotations

import inspect
import os

import torch


def _from_env(names, default=None):
    for n in names:
        if os.environ.get(n) is not None:
            return int(os.environ[n])
    return default


def get_local_rank_from_launcher() -> int:
    return _from_env(["LOCAL_RANK", "OMPI_COMM_WORLD_LOCAL_RANK", "MPI_LOCALRANKID", "SLURM_LOCALID"], 0)


def get_world_rank_from_launcher() -> int:
    return _from_env(["RANK", "OMPI_COMM_WORLD_RANK", "PMI_RANK", "SLURM_PROCID"], 0)


def get_world_size_from_launcher() -> int:
    return _from_env(["WORLD_SIZE", "OMPI_COMM_WORLD_SIZE", "PMI_SIZE", "SLURM_NTASKS"], 1)


def get_default_args(func) -> dict:
    return {k: v.default for k, v in inspect.signature(func).parameters.items() if v.default is not inspect.Parameter.empty}


def get_tensor_position(func) -> int:
    """Index of the argument that carries the payload tensor of a collective."""
    names = list(inspect.signature(func).parameters)
    for cand in ("tensor", "tensors", "input_list", "input_tensor_list", "input_tensors"):
        if cand in names:
            return names.index(cand)
    return -1


def get_tensor_kwarg(func, kwargs):
    merged = {**get_default_args(func), **kwargs}
    for cand in ("tensor", "tensors", "input_list", "input_tensor_list", "input_tensors"):
        if merged.get(cand) is not None:
            return merged[cand]
    return None


def get_msg_size_from_args(func, *args, **kwargs) -> int:
    pos = get_tensor_position(func)
    t = args[pos] if 0 <= pos < len(args) else get_tensor_kwarg(func, kwargs)
    if t is None:
        return 0
    if isinstance(t, (list, tuple)):
        return sum(x.element_size() * x.nelement() for x in t if torch.is_tensor(x))
    return t.element_size() * t.nelement()


def get_debug_log_name(func_args: dict, debug: bool) -> str:
    return func_args["log_name"] + " | [Caller Func: " + inspect.stack()[2][3] + "]" if debug else func_args["log_name"]
