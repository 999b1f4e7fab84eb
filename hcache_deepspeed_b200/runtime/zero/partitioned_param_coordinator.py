"""Module-name parity with ``deepspeed/runtime/zero/partitioned_param_coordinator.py``: fetch / release is driven by the
module hooks of :class:`.stage3.DeepSpeedZeRoOffload` (``get_param_coordinator`` returns it)."""
from .partition_parameters import all_gather_params  # noqa: F401
from .stage3 import DeepSpeedZeRoOffload as PartitionedParameterCoordinator  # noqa: F401
