"""Op objects of the v1 inference path (``deepspeed/ops/transformer/inference/op_binding/*.py``: one ``BaseOp`` subclass
per fused op, each picking a compiled kernel by dtype with a torch fallback). Here every op object forwards to the one
formulation in :class:`~..ops.InferenceOps`; the classes keep the reference's names and ``op(config)(...)`` call shape."""
from __future__ import annotations

import torch

from ..config import DeepSpeedInferenceConfig
from ..ops import InferenceOps

_shared_ops = None


def _ops() -> InferenceOps:
    global _shared_ops
    if _shared_ops is None:
        _shared_ops = InferenceOps()
    return _shared_ops


class BaseOp(torch.nn.Module):
    """``forward`` of a subclass is the op; ``self.config`` carries dtype / eps / layout flags."""
    inference_module = None

    def __init__(self, config: DeepSpeedInferenceConfig = None):
        super().__init__()
        self.config = config if config is not None else DeepSpeedInferenceConfig()
        if BaseOp.inference_module is None:
            BaseOp.inference_module = _ops()


def _simple(name: str, fn_name: str, doc: str):
    fn = getattr(InferenceOps, fn_name)

    def forward(self, *args, **kwargs):
        return fn(*args, **kwargs)

    return type(name, (BaseOp,), {"forward": forward, "__doc__": doc})


BiasAddOp = _simple("BiasAddOp", "bias_add", "x + bias")
BiasGeluOp = _simple("BiasGeluOp", "bias_gelu", "gelu(x + bias)")
BiasReluOp = _simple("BiasReluOp", "bias_relu", "relu(x + bias)")
BiasResidualOp = _simple("BiasResidualOp", "bias_residual", "x + residual + bias")
GatedActivationOp = _simple("GatedActivationOp", "gated_activation", "GEGLU / SiGLU over the two halves of the last dim")
MoEResMatmulOp = _simple("MoEResMatmulOp", "moe_res_matmul", "residual-MoE mixing of the expert and dense outputs")
VectorMatMulOp = _simple("VectorMatMulOp", "vector_matmul", "x @ W")
ResidualAddOp = _simple("ResidualAddOp", "residual_add_bias", "the sequential / parallel residual wiring")
PreRMSNormOp = _simple("PreRMSNormOp", "pre_rms_norm", "(residual + x, rms_norm(residual + x))")


class VectorAddOp(BaseOp):
    def forward(self, a, b, gamma):
        return a + gamma * b


class LayerNormOp(BaseOp):
    def forward(self, vals, gamma, beta, epsilon):
        return InferenceOps.layer_norm(vals, gamma, beta, epsilon)

    @staticmethod
    def layer_norm_residual(vals, bias, res, gamma, beta, epsilon):
        return InferenceOps.layer_norm_residual(vals, bias, res, gamma, beta, epsilon)

    @staticmethod
    def layer_norm_residual_store_pre_ln_res(vals, bias, res, gamma, beta, epsilon):
        return InferenceOps.layer_norm_residual_store_pre_ln_res(vals, bias, res, gamma, beta, epsilon)


class RMSNormOp(BaseOp):
    def forward(self, vals, gamma, epsilon):
        return InferenceOps.rms_norm(vals, gamma, epsilon)


class LinearOp(BaseOp):
    def forward(self, input, weight, bias=None, add_bias: bool = False, do_flash_attn: bool = False, num_heads: int = 1,
                external_cache=None, num_layers=None):
        out = InferenceOps.vector_matmul(input, weight)
        return out + bias if (add_bias and bias is not None) else out


class QKVGemmOp(BaseOp):
    def forward(self, input, weight, bias, gamma, beta):
        c = self.config
        return InferenceOps.qkv_gemm(input, weight, None, bias, gamma, beta, c.layer_norm_eps, bias is not None,
                                     pre_layer_norm=c.pre_layer_norm, rms_norm=c.norm_type == "rms_norm")


class MLPGemmOp(BaseOp):
    def forward(self, input, residual, weight_interm, weight_out, input_bias, bias, gamma, beta):
        c = self.config
        return InferenceOps.mlp_gemm(input, residual, input_bias, weight_interm, weight_out, bias, gamma, beta, c.layer_norm_eps,
                                     c.pre_layer_norm, c.mlp_after_attn, act=c.mlp_act_func_type,
                                     rms_norm=c.norm_type == "rms_norm")


class GELUGemmOp(BaseOp):
    """``gelu(x @ W1 + b1) @ W2`` (the fused MLP of post-LN encoders)."""

    def forward(self, input, weight, bias, weight_out):
        return InferenceOps.vector_matmul(InferenceOps.bias_gelu(InferenceOps.vector_matmul(input, weight), bias), weight_out)


class SoftmaxOp(BaseOp):
    def forward(self, attn_scores, attn_mask, alibi, triangular: bool, recompute: bool, local_attention: bool, window_size: int,
                async_op: bool, layer_scale: float, head_offset: int):
        s = attn_scores.float() * layer_scale
        if alibi is not None:
            s = s + alibi.float().reshape(s.shape[0], s.shape[1], -1, s.shape[-1])[..., -s.shape[-2]:, :] if alibi.dim() != s.dim() else s + alibi.float()
        if triangular:
            q, k = s.shape[-2], s.shape[-1]
            keep = torch.ones(q, k, dtype=torch.bool, device=s.device).tril(diagonal=k - q)
            if local_attention:
                keep &= torch.ones(q, k, dtype=torch.bool, device=s.device).triu(diagonal=k - q - window_size + 1)
            s = s.masked_fill(~keep, float("-inf"))
        if attn_mask is not None:
            s = s + attn_mask.float()
        return torch.softmax(s, dim=-1).to(attn_scores.dtype)


class SoftmaxContextOp(BaseOp):
    """Attention over the layer's KV cache: rotary -> append -> softmax(QK^T) V."""

    def forward(self, query_key_value, attn_mask, heads, num_kv, norm_factor, no_masking, layer_id, num_layers, alibi,
                is_prompt: bool = None, max_out_tokens: int = None):
        c = self.config
        if is_prompt is None:
            is_prompt = query_key_value.shape[1] > 1
        return _ops().softmax_context(query_key_value, attn_mask, c.rotary_dim, c.rotate_half, c.rotate_every_two, heads, num_kv,
                                      norm_factor, c.triangular_masking, getattr(c, "local_attention", False),
                                      getattr(c, "window_size", 1), no_masking, layer_id, num_layers, alibi,
                                      getattr(c, "rope_theta", 10000.0), is_prompt, max_out_tokens or c.max_out_tokens)


class EinsumSecSmEcmOp(BaseOp):
    def forward(self, Q, W):
        return torch.einsum("sec,sm->ecm", Q, W)


class PadTransformOp(BaseOp):
    """Pad the head dim of q/k/v to a multiple of 8 for flash kernels (diffusers attention)."""

    def forward(self, query, key, value, add_padding: bool, num_heads: int, head_dim: int = None):
        def split(x):
            b, s, hd = x.shape
            x = x.view(b, s, num_heads, hd // num_heads).transpose(1, 2)
            pad = (-x.shape[-1]) % 8 if add_padding else 0
            return torch.nn.functional.pad(x, (0, pad)) if pad else x
        return split(query), split(key), split(value)


class WorkspaceOp(BaseOp):
    def allocate_workspace(self, hidden_dim, num_heads, prompt_length, batch_size, num_layers, mp_size, external_cache, rank,
                           max_out_tokens, min_out_tokens):
        return None            # caches are allocated lazily per layer on first use

    def release_workspace(self):
        return _ops().release_workspace()

    def retake_workspace(self):
        return True

    def reset_cache(self):
        return _ops().reset_cache()

    def is_allocated(self):
        return True
