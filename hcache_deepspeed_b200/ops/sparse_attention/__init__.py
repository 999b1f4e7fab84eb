from .sparse_self_attention import SparseSelfAttention  # noqa: F401
from .sparsity_config import (BigBirdSparsityConfig, BSLongformerSparsityConfig, DenseSparsityConfig, FixedSparsityConfig,  # noqa: F401
                              LocalSlidingWindowSparsityConfig, SparsityConfig, VariableSparsityConfig)
