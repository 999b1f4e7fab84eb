"""Tensor-parallel sharding helpers under their ``deepspeed/inference/v2/model_implementations/sharding`` names
(``attn.py``, ``attn_out.py``, ``embedding.py``, ``mlp.py``, ``qkv.py``, ``unembed.py``, ``types.py``, ``utils.py``).
The implementation is :mod:`hcache_deepspeed_b200.parallel.sharding`."""
from enum import Enum

from .....parallel.sharding import (attn_out_in_features, get_local_heads, get_shard_endpoints, qkv_out_features,  # noqa: F401
                                    shard_attn_out_param, shard_embedding_param, shard_mlp_1_param, shard_mlp_2_param,
                                    shard_qkv_param, shard_unembed_param, sharded_intermediate_dim, sharded_unembed_dim)


class ShardingType(Enum):
    OUTER_DIMENSION = 0     # split the output features (column parallel: QKV, MLP-1, unembed)
    INNER_DIMENSION = 1     # split the input features (row parallel: attention out, MLP-2)


DEFAULT_SHARD_GRANULARITY = 32
