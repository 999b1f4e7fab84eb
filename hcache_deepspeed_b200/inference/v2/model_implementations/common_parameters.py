"""The parameter types that cover the supported model families: fused / unfused / Megatron-layout QKV, plain / gated /
pre-fused-gated MLP, attention output, embedding, unembedding, norm, rotary inverse frequencies and the MoE variants.
Each ``finalize`` puts the source tensors into this framework's canonical layout ([q; k; v] head-major QKV rows,
INTERLEAVED (gate, up) rows for gated MLP-1 — what the gated GEMM epilogue consumes) and hands the result to the model's
``transform_*_param`` (shard for tensor parallelism, cast, move to the device).

Reference: ``deepspeed/inference/v2/model_implementations/common_parameters/*.py`` (``qkv_parameters.py:14-115``,
``mlp_parameters.py:14-99``, ``moe_parameters.py:16-78``, ``attn_output/embedding/unembed/norm/invfreq_parameters.py``).
"""
from __future__ import annotations

import torch

from .parameter_base import ParameterBase, ParamList


def interleave_rows(gate: torch.Tensor, up: torch.Tensor) -> torch.Tensor:
    """[g0, g1, ...], [u0, u1, ...] -> [g0, u0, g1, u1, ...] along dim 0 (works for weights and biases)."""
    assert gate.shape == up.shape, "Gated MLP parameters must have the same number of neurons."
    return torch.stack([gate, up], dim=1).reshape(2 * gate.shape[0], *gate.shape[1:])


def megatron_qkv_reshape(param: torch.Tensor, head_size: int, n_heads: int) -> torch.Tensor:
    """Megatron stores [n_heads, 3, head_size, in]; canonical is [3, n_heads, head_size, in]."""
    assert param.shape[0] == 3 * n_heads * head_size
    v = param.reshape(n_heads, 3, head_size, *param.shape[1:])
    return v.transpose(0, 1).reshape(param.shape)


def transform_gqa_megatron(src_param: torch.Tensor, head_size: int, n_q_heads: int, n_kv_heads: int) -> torch.Tensor:
    """Grouped Megatron layout: per KV group, its ``n_q/n_kv`` query heads, then one K head, then one V head."""
    assert src_param.shape[0] == (2 * n_kv_heads + n_q_heads) * head_size
    ratio = n_q_heads // n_kv_heads
    g = src_param.reshape(n_kv_heads, ratio + 2, head_size, *src_param.shape[1:])
    tail = src_param.shape[1:]
    q = g[:, :ratio].reshape(n_q_heads * head_size, *tail)
    k = g[:, ratio].reshape(n_kv_heads * head_size, *tail)
    v = g[:, ratio + 1].reshape(n_kv_heads * head_size, *tail)
    return torch.cat([q, k, v], dim=0)


# ---- attention ---------------------------------------------------------------------------------------------------
class FusedQKVParameter(ParameterBase):
    params: torch.Tensor

    def finalize(self) -> torch.Tensor:
        return self.inference_model.transform_qkv_param(self.params)


class UnfusedQKVParameter(ParameterBase):
    q_params: torch.Tensor
    k_params: torch.Tensor
    v_params: torch.Tensor

    def finalize(self) -> torch.Tensor:
        return self.inference_model.transform_qkv_param(torch.cat([self.q_params, self.k_params, self.v_params], dim=0))


class MegatronQKVParameter(ParameterBase):
    params: torch.Tensor

    def finalize(self) -> torch.Tensor:
        m = self.inference_model
        return m.transform_qkv_param(megatron_qkv_reshape(self.params, m.head_size, m.n_heads))


class GQAMegatronQKVParameter(ParameterBase):
    params: torch.Tensor

    def finalize(self) -> torch.Tensor:
        m = self.inference_model
        return m.transform_qkv_param(transform_gqa_megatron(self.params, m.head_size, m.n_heads_q, m.n_heads_kv))


class AttentionOutputParameter(ParameterBase):
    params: torch.Tensor

    def finalize(self) -> torch.Tensor:
        return self.inference_model.transform_attn_out_param(self.params)


class InvFreqParameter(ParameterBase):
    params: torch.Tensor

    def finalize(self) -> torch.Tensor:
        dt = self.inference_model.activation_dtype
        return self.params.to(getattr(dt, "value", dt))


# ---- MLP ---------------------------------------------------------------------------------------------------------
class MLP1Parameter(ParameterBase):
    params: torch.Tensor

    def finalize(self) -> torch.Tensor:
        return self.inference_model.transform_mlp_1_param(self.params)


class GatedMLPParameter(ParameterBase):
    gate_params: torch.Tensor
    up_params: torch.Tensor

    def finalize(self) -> torch.Tensor:
        return self.inference_model.transform_mlp_1_param(interleave_rows(self.gate_params, self.up_params))


class FusedGatedMLPParameter(ParameterBase):
    """Checkpoint stores [gate; up] stacked in one tensor."""
    params: torch.Tensor

    def finalize(self) -> torch.Tensor:
        half = self.params.shape[0] // 2
        return self.inference_model.transform_mlp_1_param(interleave_rows(self.params[:half], self.params[half:]))


class MLP2Parameter(ParameterBase):
    params: torch.Tensor

    def finalize(self) -> torch.Tensor:
        return self.inference_model.transform_mlp_2_param(self.params)


# ---- embedding / unembedding / norm ------------------------------------------------------------------------------
class EmbeddingParameter(ParameterBase):
    params: torch.Tensor

    def finalize(self) -> torch.Tensor:
        return self.inference_model.transform_embedding_param(self.params)


class UnembedParameter(ParameterBase):
    params: torch.Tensor

    def finalize(self) -> torch.Tensor:
        return self.inference_model.transform_unembed_param(self.params)


class NormParameter(ParameterBase):
    params: torch.Tensor

    def finalize(self) -> torch.Tensor:
        return self.inference_model.transform_norm_param(self.params)


# ---- mixture of experts ------------------------------------------------------------------------------------------
class MoEGatingWeightParameter(ParameterBase):
    params: torch.Tensor

    def finalize(self) -> torch.Tensor:
        return self.inference_model.transform_moe_gate_param(self.params)


class UnfusedMoEMLP1Parameter(ParameterBase):
    experts: ParamList("n_experts")  # noqa: F821

    def finalize(self) -> torch.Tensor:
        return self.inference_model.transform_moe_mlp_1_param(torch.stack(tuple(self.experts), dim=0))


class UnfusedMoEMLP2Parameter(ParameterBase):
    experts: ParamList("n_experts")  # noqa: F821

    def finalize(self) -> torch.Tensor:
        return self.inference_model.transform_moe_mlp_2_param(torch.stack(tuple(self.experts), dim=0))


class UnfusedMoEGatedMLPParameter(ParameterBase):
    gating_experts: ParamList("n_experts")  # noqa: F821
    up_experts: ParamList("n_experts")  # noqa: F821

    def finalize(self) -> torch.Tensor:
        stacked = torch.stack([interleave_rows(g, u) for g, u in zip(self.gating_experts, self.up_experts)], dim=0)
        return self.inference_model.transform_moe_mlp_1_param(stacked)
