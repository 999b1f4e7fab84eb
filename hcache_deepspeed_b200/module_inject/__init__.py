"""Module injection: automatic tensor parallelism for arbitrary HF-style modules (AutoTP), the tensor-parallel layer
zoo, and the policy table that maps model families onto the sm_100a executor (``deepspeed.module_inject``)."""
from .auto_tp import AutoTP, ReplaceWithTensorSlicing  # noqa: F401
from .layers import (EmbeddingLayer, LinearAllreduce, LinearLayer, LmHeadLinearAllreduce, Normalize, OPTEmbedding,  # noqa: F401
                     RMSNormalize)
from .module_quantize import quantize_transformer_layer  # noqa: F401
from .replace_module import GroupQuantizer, generic_injection, replace_transformer_layer, revert_transformer_layer  # noqa: F401
from .tp_shard import get_num_kv_heads, get_shard_size, get_shard_size_list, set_num_kv_heads  # noqa: F401
