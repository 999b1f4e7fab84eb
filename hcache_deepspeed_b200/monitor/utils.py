"""``deepspeed/monitor/utils.py``: is an optional experiment-tracking package importable?"""
import importlib.util


def _available(name: str) -> bool:
    return importlib.util.find_spec(name) is not None


def check_tb_availability():
    if not _available("tensorboard") and not _available("torch.utils.tensorboard"):
        raise ImportError("If you want to use tensorboard logging, please `pip install tensorboard`")


def check_wandb_availability():
    if not _available("wandb"):
        raise ImportError("If you want to use wandb logging, please `pip install wandb` and follow the instructions at https://docs.wandb.ai/quickstart")


def check_comet_availability():
    if not _available("comet_ml"):
        raise ImportError("If you want to use comet logging, please `pip install \"comet_ml>=3.41.0\"`")
