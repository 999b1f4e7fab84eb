"""Module registry of the v2 engine: named, configurable building blocks (attention, linear, norms, embedding,
unembedding, MoE) that can be instantiated standalone — the role of ``deepspeed/inference/v2/modules``
(``ds_module.py:19-62``, ``module_registry.py:14-58``, ``configs/*.py``, ``heuristics.py:36-195``).

Inside :class:`hcache_deepspeed_b200.models.transformer.RaggedTransformer` these operations are fused further
(norm + residual fold, QKV GEMM + RoPE + KV append, gated epilogues), so the model executor calls the op back end
directly; the modules here wrap the same back-end calls one operation at a time for users who compose their own model.
"""
from .configs import (DSEmbeddingsConfig, DSLinearConfig, DSMoEConfig, DSNormConfig, DSSelfAttentionConfig,  # noqa: F401
                      DSUnembedConfig, MaskingType, NormTypeEnum, PositionalEmbeddingType, RotateHalfConfig)
from .ds_module import ConfigBundle, DSModuleBase, DSModuleRegistryBase  # noqa: F401
from .heuristics import (instantiate_attention, instantiate_embed, instantiate_linear, instantiate_moe,  # noqa: F401
                         instantiate_post_norm, instantiate_pre_norm, instantiate_unembed)
from .implementations import (DSAttentionRegistry, DSEmbeddingRegistry, DSLinearRegistry, DSMoERegistry,  # noqa: F401
                              DSPostNormRegistry, DSPreNormRegistry, DSUnembedRegistry)
