from .async_swapper import AsyncTensorSwapper  # noqa: F401
from .partitioned_param_swapper import AsyncPartitionedParameterSwapper, PartitionedParamStatus  # noqa: F401
from .utils import SwapBuffer, SwapBufferManager, SwapBufferPool  # noqa: F401
