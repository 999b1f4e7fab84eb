"""Logging helpers (ref: deepspeed/utils/logging.py:1-163, deepspeed/inference/v2/logging.py:13-26)."""
from __future__ import annotations

import functools
import logging
import os
import sys
from typing import Iterable, Optional

_LOG_LEVELS = {"debug": logging.DEBUG, "info": logging.INFO, "warning": logging.WARNING, "error": logging.ERROR,
               "critical": logging.CRITICAL}


class LoggerFactory:

    @staticmethod
    def create_logger(name: Optional[str] = None, level=logging.INFO) -> logging.Logger:
        if name is None:
            raise ValueError("name for logger cannot be None")
        formatter = logging.Formatter("[%(asctime)s] [%(levelname)s] [%(filename)s:%(lineno)d:%(funcName)s] %(message)s")
        lg = logging.getLogger(name)
        lg.setLevel(level)
        lg.propagate = False
        if not lg.handlers:
            ch = logging.StreamHandler(stream=sys.stderr)
            ch.setLevel(level)
            ch.setFormatter(formatter)
            lg.addHandler(ch)
        return lg


logger = LoggerFactory.create_logger(name="hcache_deepspeed_b200",
                                     level=_LOG_LEVELS.get(os.environ.get("HDS_LOG_LEVEL", "warning").lower(),
                                                           logging.WARNING))

_inf_logger = None


def inference_logger(level: int = logging.WARNING) -> logging.Logger:
    """Logger of the ragged engine. A disabled log call still costs microseconds: keep it out of hot loops."""
    global _inf_logger
    if _inf_logger is None:
        _inf_logger = LoggerFactory.create_logger(name="hds-inference", level=_LOG_LEVELS.get(
            os.environ.get("HDS_LOG_LEVEL", "").lower(), level))
    return _inf_logger


def _rank() -> int:
    try:
        import torch.distributed as dist
        if dist.is_available() and dist.is_initialized():
            return dist.get_rank()
    except Exception:
        pass
    return int(os.environ.get("RANK", 0))


def log_dist(message: str, ranks: Optional[Iterable[int]] = None, level: int = logging.INFO) -> None:
    """Log on the listed ranks only (``[-1]`` or None-with-rank-0 semantics as in the reference)."""
    my_rank = _rank()
    ranks = list(ranks) if ranks is not None else [0]
    if -1 in ranks or my_rank in ranks:
        logger.log(level, f"[Rank {my_rank}] {message}")


@functools.lru_cache(None)
def warning_once(*args, **kwargs) -> None:
    logger.warning(*args, **kwargs)


def print_json_dist(message: dict, ranks: Optional[Iterable[int]] = None, path: Optional[str] = None) -> None:
    import json
    my_rank = _rank()
    ranks = list(ranks) if ranks is not None else [0]
    if -1 in ranks or my_rank in ranks:
        message["rank"] = my_rank
        with open(path, "w") as f:
            json.dump(message, f)
            os.fsync(f)


def get_current_level() -> int:
    return logger.getEffectiveLevel()


def should_log_le(max_log_level_str: str) -> bool:
    if not isinstance(max_log_level_str, str):
        raise ValueError("max_log_level_str must be a string")
    key = max_log_level_str.lower()
    if key not in _LOG_LEVELS:
        raise ValueError(f"{max_log_level_str} is not one of the logging levels")
    return get_current_level() <= _LOG_LEVELS[key]


def print_configuration(args, name: str) -> None:
    """Log every attribute of an argparse namespace / config object, one dotted line each."""
    logger.info(f"{name}:")
    for arg in sorted(vars(args)):
        logger.info(f"  {arg} {'.' * max(1, 29 - len(arg))} {getattr(args, arg)}")
