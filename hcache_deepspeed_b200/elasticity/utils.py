"""``deepspeed/elasticity/utils.py``: is the elastic agent usable with this torch?"""
from packaging import version as pkg_version


def is_torch_elastic_compatible() -> bool:
    """The agent builds on ``torch.distributed.elastic`` (torch >= 1.11)."""
    import torch
    return pkg_version.parse(torch.__version__.split("+")[0]) >= pkg_version.parse("1.11.0")
