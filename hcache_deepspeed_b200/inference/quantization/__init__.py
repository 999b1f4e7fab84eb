from .layers import QuantizedEmbedding, QuantizedLinear  # noqa: F401
from .quantization import _init_group_wise_weight_quantization  # noqa: F401
