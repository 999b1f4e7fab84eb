"""v1 kernel-injection containers: per-family policies (``families.py``) that read a HuggingFace block into
:class:`LayerTensors`, and the machinery that swaps in the fused layer (``base.py``). Families the v2 ragged executor
covers are listed in :mod:`hcache_deepspeed_b200.module_inject.policy` and lifted whole instead."""
from ..policy import InjectionPolicy, policy_for, replace_policies  # noqa: F401
from .base import (LAYER_CLASS_NAMES, POLICIES, InjectedLayer, LayerTensors, build_fused_layer, inject_fused_layers,  # noqa: F401
                   register_policy)
from . import families  # noqa: F401,E402
