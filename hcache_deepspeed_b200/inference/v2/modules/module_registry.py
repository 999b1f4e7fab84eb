"""``DSModuleRegistryBase`` under its reference module name (``deepspeed/inference/v2/modules/module_registry.py``)."""
from .ds_module import ConfigBundle, DSModuleRegistryBase  # noqa: F401
