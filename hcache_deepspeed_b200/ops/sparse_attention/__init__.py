from .bert_sparse_self_attention import BertSparseSelfAttention  # noqa: F401
from .matmul import MatMul  # noqa: F401
from .softmax import Softmax  # noqa: F401
from .sparse_attention_utils import SparseAttentionUtils  # noqa: F401
from .sparse_self_attention import SparseSelfAttention  # noqa: F401
from .sparsity_config import (BigBirdSparsityConfig, BSLongformerSparsityConfig, DenseSparsityConfig, FixedSparsityConfig,  # noqa: F401
                              LocalSlidingWindowSparsityConfig, SparsityConfig, VariableSparsityConfig)
