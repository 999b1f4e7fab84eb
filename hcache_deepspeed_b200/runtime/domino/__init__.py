from .transformer import DominoTransformerLayer, DominoAsyncLinear  # noqa: F401
from .async_linear import DominoAsyncColumnParallelLinear, DominoAsyncColumnParallelLinearImpl, RowParallelLinearNoComm  # noqa: F401
