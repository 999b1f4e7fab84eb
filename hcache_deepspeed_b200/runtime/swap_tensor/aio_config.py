"""``"aio"`` section defaults (``deepspeed/runtime/swap_tensor/aio_config.py`` / ``constants.py``)."""
AIO_DEFAULT_DICT = {"block_size": 1048576, "queue_depth": 8, "intra_op_parallelism": 1, "single_submit": False,
                    "overlap_events": True, "use_gds": False}


def get_aio_config(param_dict: dict) -> dict:
    return {**AIO_DEFAULT_DICT, **(param_dict.get("aio") or {})}
