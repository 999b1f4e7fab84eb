"""``policy_to_ds_container`` (``deepspeed/module_inject/utils.py``): the weight-extraction policy for an HF block."""
from __future__ import annotations

from .containers import families  # noqa: F401  (registers the policies)
from .containers.base import LAYER_CLASS_NAMES, POLICIES


def policy_to_ds_container(policy=None, block=None, **kwargs):
    """The registered extraction function (``(block, hf_config) -> LayerTensors``) for a block instance / class / name."""
    key = block if block is not None else policy
    name = key if isinstance(key, str) else (key.__name__ if isinstance(key, type) else type(key).__name__)
    if name not in POLICIES:
        raise ValueError(f"no kernel-injection policy for {name}; known: {sorted(POLICIES)}")
    return POLICIES[name]


def family_of(block) -> str:
    return LAYER_CLASS_NAMES[type(block).__name__]
