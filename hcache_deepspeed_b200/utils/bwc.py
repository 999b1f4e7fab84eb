"""Querying tensor / pipeline parallel coordinates from ``mpu`` objects of any vintage
(``deepspeed/utils/bwc.py:7-104``): current Megatron names first, then the "slice" names, then the legacy ones."""


def _first(mpu, names, default):
    if mpu is None:
        return default
    for n in names:
        if hasattr(mpu, n):
            return getattr(mpu, n)()
    raise AttributeError(f"{type(mpu).__name__} has none of {names}")


def bwc_tensor_model_parallel_rank(mpu=None) -> int:
    return _first(mpu, ("get_tensor_model_parallel_rank", "get_slice_parallel_rank", "get_model_parallel_rank"), 0)


def bwc_tensor_model_parallel_world_size(mpu=None) -> int:
    return _first(mpu, ("get_tensor_model_parallel_world_size", "get_slice_parallel_world_size",
                        "get_model_parallel_world_size"), 1)


def bwc_tensor_model_parallel_group(mpu=None):
    return _first(mpu, ("get_tensor_model_parallel_group", "get_slice_parallel_group", "get_model_parallel_group"), None)


def bwc_pipeline_parallel_world_size(mpu=None) -> int:
    return _first(mpu, ("get_pipeline_model_parallel_world_size", "get_pipe_parallel_world_size"), 1)


def bwc_pipeline_parallel_group(mpu=None):
    return _first(mpu, ("get_pipeline_model_parallel_group", "get_pipe_parallel_group"), None)
