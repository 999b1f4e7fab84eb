"""Policy tables under their reference names (``deepspeed/module_inject/replace_policy.py``): ``replace_policies`` — the
kernel-injection policies (one per supported HF block class); ``generic_policies`` — diffusers sub-models."""
from .containers.base import LAYER_CLASS_NAMES, POLICIES
from .containers import families  # noqa: F401  (registers the policies)
from .policy import InjectionPolicy, policy_for  # noqa: F401

replace_policies = list(POLICIES.values())
policy_class_names = dict(LAYER_CLASS_NAMES)
generic_policies = ["UNetPolicy", "VAEPolicy"]
