from .config_v2 import DeepSpeedTPConfig, HCacheConfig, QuantizationConfig, RaggedInferenceEngineConfig  # noqa: F401
from .engine_factory import (build_engine_from_arch, build_engine_from_ds_checkpoint, build_engine_from_module,  # noqa: F401
                             build_hf_engine)
from .engine_v2 import InferenceEngineV2, LatentHandle  # noqa: F401
from .ragged import DSStateManagerConfig  # noqa: F401
from .scheduling_utils import SchedulingError, SchedulingResult  # noqa: F401
