"""ZeRO: optimizer-state (stage 1), gradient (stage 2) and parameter (stage 3) partitioning over the data-parallel
group. Public names mirror ``deepspeed.zero``."""
from .config import ZeroStageEnum  # noqa: F401
from .partition_parameters import (GatheredParameters, Init, ZeroParamStatus, ZeroParamType, register_external_parameter,  # noqa: F401
                                   unregister_external_parameter)
from .stage3 import unwrap_model_for_generation  # noqa: F401
from .tiling import TiledLinear, TiledLinearReturnBias  # noqa: F401
from .mics import MiCS_Init, MiCS_Optimizer  # noqa: F401
