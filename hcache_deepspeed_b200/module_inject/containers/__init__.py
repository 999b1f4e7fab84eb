"""Per-family containers of the reference (``module_inject/containers/*.py``) are folded into
``models/families.py`` + ``models/arch.py``; see :mod:`hcache_deepspeed_b200.module_inject.policy`."""
from ..policy import InjectionPolicy, policy_for, replace_policies  # noqa: F401
