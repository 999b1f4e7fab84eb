"""torch version gate and gradient-hook shim (``deepspeed/utils/torch.py:11-31``)."""
from packaging import version as pkg_version

import torch


def required_torch_version(min_version=None, max_version=None) -> bool:
    assert min_version or max_version, "Must provide a min_version or max_version argument"
    v = pkg_version.parse(torch.__version__.split("+")[0])
    if min_version and pkg_version.parse(str(min_version)) > v:
        return False
    if max_version and pkg_version.parse(str(max_version)) < v:
        return False
    return True


def register_grad_hook(param, hook):
    """Fire ``hook(param)`` once the parameter's gradient has been accumulated."""
    return param.register_post_accumulate_grad_hook(hook)
