"""Pick an implementation by name for a config (``deepspeed/inference/v2/modules/heuristics.py:36-195``). One
implementation exists per interface here (the sm_100a back end), except the pre-norm which depends on the norm type."""
from __future__ import annotations

from .configs import (DSEmbeddingsConfig, DSLinearConfig, DSMoEConfig, DSNormConfig, DSSelfAttentionConfig, DSUnembedConfig,
                      NormTypeEnum)
from .ds_module import ConfigBundle
from .implementations import (DSAttentionRegistry, DSEmbeddingRegistry, DSLinearRegistry, DSMoERegistry, DSPostNormRegistry,
                              DSPreNormRegistry, DSUnembedRegistry)


def instantiate_attention(attention_config: DSSelfAttentionConfig, engine_config=None):
    return DSAttentionRegistry.instantiate_config(ConfigBundle("dense_blocked_attention", attention_config))


def instantiate_embed(embed_config: DSEmbeddingsConfig, engine_config=None):
    return DSEmbeddingRegistry.instantiate_config(ConfigBundle("ragged_embedding", embed_config))


def instantiate_linear(linear_config: DSLinearConfig, engine_config=None):
    mode = getattr(getattr(engine_config, "quantization", None), "quantization_mode", None)
    if mode == "wf6af16":
        return DSLinearRegistry.instantiate_config(ConfigBundle("fp6_tcgen05_linear", linear_config))
    if mode == "wf8af8":
        return DSLinearRegistry.instantiate_config(ConfigBundle("fp8_tcgen05_linear", linear_config))
    return DSLinearRegistry.instantiate_config(ConfigBundle("blas_fp_linear", linear_config))


def instantiate_moe(moe_config: DSMoEConfig, engine_config=None):
    return DSMoERegistry.instantiate_config(ConfigBundle("cutlass_multi_gemm_moe", moe_config))


def instantiate_post_norm(norm_config: DSNormConfig, engine_config=None):
    return DSPostNormRegistry.instantiate_config(ConfigBundle("cuda_post_ln", norm_config))


def instantiate_pre_norm(norm_config: DSNormConfig, engine_config=None):
    name = "cuda_pre_rms" if NormTypeEnum(norm_config.type) == NormTypeEnum.RMSNorm else "cuda_pre_ln"
    return DSPreNormRegistry.instantiate_config(ConfigBundle(name, norm_config))


def instantiate_unembed(unembed_config: DSUnembedConfig, engine_config=None):
    return DSUnembedRegistry.instantiate_config(ConfigBundle("ragged_unembed", unembed_config))
