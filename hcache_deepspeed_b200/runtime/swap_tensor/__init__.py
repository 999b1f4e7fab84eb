from .async_swapper import AsyncTensorSwapper  # noqa: F401
from .partitioned_param_swapper import AsyncPartitionedParameterSwapper, PartitionedParamStatus  # noqa: F401
from .utils import SwapBuffer, SwapBufferManager, SwapBufferPool  # noqa: F401
from .optimizer_utils import OptimizerStateSwapInfo, OptimizerSwapper  # noqa: F401
from .partitioned_optimizer_swapper import PartitionedOptimizerSwapper  # noqa: F401
from .pipelined_optimizer_swapper import PipelinedOptimizerSwapper  # noqa: F401
