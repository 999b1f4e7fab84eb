"""Interfaces, registries and the implementations that sit on the op back ends (``ops/backend.py``).
Names follow the reference's implementations so configs carry over: ``blas_fp_linear``
(``implementations/linear/blas_fp_linear.py``), ``cuda_pre_rms`` / ``cuda_pre_ln`` / ``cuda_post_ln``,
``ragged_embedding``, ``ragged_unembed``, ``dense_blocked_attention``, ``cutlass_multi_gemm_moe``."""
from __future__ import annotations

from typing import Optional, Tuple, Type

import torch

from ....ops import reference as ref
from ....ops.backend import get_ops
from ....ops.kernels import ACT_CODES
from .configs import (DSEmbeddingsConfig, DSLinearConfig, DSMoEConfig, DSNormConfig, DSSelfAttentionConfig, DSUnembedConfig,
                      NormTypeEnum, PositionalEmbeddingType)
from .ds_module import DSModuleBase, DSModuleRegistryBase

_GATED = {"geglu": "gelu", "reglu": "relu", "siglu": "silu"}


def _interface(cfg_cls):
    class _Base(DSModuleBase):
        @staticmethod
        def config_class():
            return cfg_cls

    return _Base


DSLinearBase = _interface(DSLinearConfig)
DSPreNormBase = _interface(DSNormConfig)
DSPostNormBase = _interface(DSNormConfig)
DSEmbeddingBase = _interface(DSEmbeddingsConfig)
DSUnembedBase = _interface(DSUnembedConfig)
DSSelfAttentionBase = _interface(DSSelfAttentionConfig)
DSMoEBase = _interface(DSMoEConfig)


def _registry(base):
    class _Registry(DSModuleRegistryBase):
        @staticmethod
        def associated_class() -> Type[DSModuleBase]:
            return base

    return _Registry


DSLinearRegistry, DSPreNormRegistry, DSPostNormRegistry = _registry(DSLinearBase), _registry(DSPreNormBase), _registry(DSPostNormBase)
DSEmbeddingRegistry, DSUnembedRegistry = _registry(DSEmbeddingBase), _registry(DSUnembedBase)
DSAttentionRegistry, DSMoERegistry = _registry(DSSelfAttentionBase), _registry(DSMoEBase)


@DSLinearRegistry.register_module
class BlasFPLinear(DSLinearBase):
    """``y = act(x W^T + b)``; gated activations expect interleaved (gate, up) weight rows (``transform_param``)."""

    @staticmethod
    def name():
        return "blas_fp_linear"

    @staticmethod
    def supports_config(config: DSLinearConfig) -> bool:
        return config.input_dtype == config.output_dtype and (config.activation in _GATED or config.activation in ACT_CODES)

    def __init__(self, config, implementation_config=None):
        super().__init__(config, implementation_config)
        self.gated = config.activation in _GATED
        self.act = ACT_CODES[_GATED.get(config.activation, config.activation)]

    def transform_param(self, param: torch.Tensor) -> torch.Tensor:
        """[gate; up] stacked halves -> interleaved rows (what the gated epilogue consumes)."""
        if not self.gated or param.shape[0] != self._config.out_channels * 2:
            return param
        half = param.shape[0] // 2
        return torch.stack([param[:half], param[half:]], dim=1).reshape(param.shape).contiguous()

    def forward(self, hidden_states: torch.Tensor, w: torch.Tensor, b: Optional[torch.Tensor] = None) -> torch.Tensor:
        ops = get_ops(hidden_states.device)
        n_out = w.shape[0] // 2 if self.gated else w.shape[0]
        out = torch.empty((hidden_states.shape[0], n_out), dtype=hidden_states.dtype, device=hidden_states.device)
        return ops.gemm(hidden_states, w, out, bias=b, act=self.act, gated=self.gated)


@DSLinearRegistry.register_module
class Fp8TensorCoreLinear(BlasFPLinear):
    """W8A8 linear on the fp8 tensor cores (``quantization_mode = "wf8af8"``): weights are quantised to E4M3 per output
    channel once (``transform_param``), activations per token on every call; the GEMM is ``tcgen05.mma.kind::f8f6f4``
    with both scales applied in the epilogue (``ops/csrc/gemm_fp8.cu``). Takes the place of the reference's
    ``quantized_wf6af16_linear`` (``implementations/linear/quantized_linear.py:24-205``), whose kernel is sm_80-only."""

    @staticmethod
    def name():
        return "fp8_tcgen05_linear"

    @staticmethod
    def supports_config(config: DSLinearConfig) -> bool:
        return BlasFPLinear.supports_config(config) and config.in_channels % 16 == 0

    def transform_param(self, param: torch.Tensor) -> torch.Tensor:
        param = super().transform_param(param)
        if param.dim() != 2:
            return param  # biases stay in the activation dtype
        if param.is_cuda:
            from ....ops import kernels as K
            q, scales = K.quantize_fp8_rows(param.contiguous())
        else:
            amax = param.float().abs().amax(dim=1)
            scales = torch.where(amax > 0, amax / 448.0, torch.ones_like(amax))
            q = (param.float() / scales[:, None]).to(torch.float8_e4m3fn).view(torch.uint8)
        q.scales = scales
        return q

    def forward(self, hidden_states: torch.Tensor, w: torch.Tensor, b: Optional[torch.Tensor] = None) -> torch.Tensor:
        n_out = w.shape[0] // 2 if self.gated else w.shape[0]
        if hidden_states.is_cuda:
            from ....ops import kernels as K
            out = torch.empty((hidden_states.shape[0], n_out), dtype=hidden_states.dtype, device=hidden_states.device)
            aq, a_s = K.quantize_fp8_rows(hidden_states)
            return K.gemm_fp8(aq, a_s, w, w.scales, out, bias=b, act=self.act, gated=self.gated)
        amax = hidden_states.float().abs().amax(dim=1, keepdim=True)
        a_s = torch.where(amax > 0, amax / 448.0, torch.ones_like(amax))
        ad = (hidden_states.float() / a_s).to(torch.float8_e4m3fn).float() * a_s
        wd = w.view(torch.float8_e4m3fn).float() * w.scales[:, None]
        return ref.gemm(ad, wd, bias=b, act=self.act, gated=self.gated, out_dtype=hidden_states.dtype)


@DSLinearRegistry.register_module
class Fp6TensorCoreLinear(BlasFPLinear):
    """``quantization_mode = "wf6af16"``: FP6 (E3M2) weights, 0.75 byte per weight in HBM, per-output-channel scales.
    The reference dequantises FP6 to fp16 inside an sm_80 ``mma.sync`` kernel
    (``implementations/linear/quantized_linear.py:96-205``, ``core_ops/cuda_linear/include/kernel_matmul.cuh:21-272``);
    on Blackwell the tensor cores take E3M2 operands natively (``tcgen05.mma.kind::f8f6f4``): the TMA unpacks the 6-bit
    stream into shared memory (tensor-map type 16U6_ALIGN16B) and the activations are quantised to E4M3 per token, so the
    GEMM is the fp8 kernel with a different weight format (``ops/csrc/gemm_fp8.cu``). ``in_channels`` must be a multiple
    of 128."""

    @staticmethod
    def name():
        return "fp6_tcgen05_linear"

    @staticmethod
    def supports_config(config: DSLinearConfig) -> bool:
        return BlasFPLinear.supports_config(config) and config.in_channels % 128 == 0

    def transform_param(self, param: torch.Tensor):
        param = super().transform_param(param)
        if param.dim() != 2:
            return param  # biases stay in the activation dtype
        from ....ops.fp8 import Fp8Weight
        return Fp8Weight.from_tensor(param, "e3m2")

    def forward(self, hidden_states: torch.Tensor, w, b: Optional[torch.Tensor] = None) -> torch.Tensor:
        n_out = w.shape[0] // 2 if self.gated else w.shape[0]
        if hidden_states.is_cuda:
            from ....ops import kernels as K
            out = torch.empty((hidden_states.shape[0], n_out), dtype=hidden_states.dtype, device=hidden_states.device)
            aq, a_s = K.quantize_fp8_rows(hidden_states)
            return K.gemm_fp8(aq, a_s, w.q, w.scale, out, bias=b, act=self.act, gated=self.gated, w_fmt=1)
        amax = hidden_states.float().abs().amax(dim=1, keepdim=True)
        a_s = torch.where(amax > 0, amax / 448.0, torch.ones_like(amax))
        ad = (hidden_states.float() / a_s).to(torch.float8_e4m3fn).float() * a_s
        return ref.gemm(ad, w.dequantize(), bias=b, act=self.act, gated=self.gated, out_dtype=hidden_states.dtype)


class _PreNorm(DSPreNormBase):
    layer_norm = False

    def forward(self, residual: torch.Tensor, hidden_in: Optional[torch.Tensor], gamma: torch.Tensor,
                beta: Optional[torch.Tensor] = None) -> Tuple[torch.Tensor, torch.Tensor]:
        """``residual += hidden_in`` (if given), returns ``(residual, norm(residual))``."""
        ops = get_ops(residual.device)
        out = torch.empty_like(residual)
        ops.norm(residual, gamma, out, beta=beta, delta=hidden_in, res_out=residual if hidden_in is not None else None,
                 eps=self._config.eps, layer_norm=self.layer_norm)
        return residual, out


@DSPreNormRegistry.register_module
class CudaPreRMS(_PreNorm):
    @staticmethod
    def name():
        return "cuda_pre_rms"

    @staticmethod
    def supports_config(config):
        return NormTypeEnum(config.type) == NormTypeEnum.RMSNorm and config.channels % 8 == 0


@DSPreNormRegistry.register_module
class CudaPreLN(_PreNorm):
    layer_norm = True

    @staticmethod
    def name():
        return "cuda_pre_ln"

    @staticmethod
    def supports_config(config):
        return NormTypeEnum(config.type) == NormTypeEnum.LayerNorm and config.channels % 8 == 0


@DSPostNormRegistry.register_module
class CudaPostLN(DSPostNormBase):
    @staticmethod
    def name():
        return "cuda_post_ln"

    @staticmethod
    def supports_config(config):
        return NormTypeEnum(config.type) == NormTypeEnum.LayerNorm and config.channels % 8 == 0

    def forward(self, residual, hidden_in, gamma, beta=None):
        """``norm(residual + hidden_in)`` written over ``residual`` (post-LN)."""
        ops = get_ops(residual.device)
        summed = residual + hidden_in
        ops.norm(summed, gamma, residual, beta=beta, eps=self._config.eps, layer_norm=True)
        return residual


@DSEmbeddingRegistry.register_module
class RaggedEmbedding(DSEmbeddingBase):
    @staticmethod
    def name():
        return "ragged_embedding"

    @staticmethod
    def supports_config(config: DSEmbeddingsConfig) -> bool:
        return not config.use_token_type and config.output_normalization is None and config.embedding_dim % 8 == 0

    def forward(self, ids: torch.Tensor, word_embeddings: torch.Tensor, position_embeddings: Optional[torch.Tensor] = None,
                tok_pos: Optional[torch.Tensor] = None) -> torch.Tensor:
        ops = get_ops(word_embeddings.device)
        out = torch.empty((ids.shape[0], word_embeddings.shape[1]), dtype=word_embeddings.dtype, device=word_embeddings.device)
        return ops.embed(ids, word_embeddings, out, pos_weight=position_embeddings, tok_pos=tok_pos,
                         pos_offset=self._config.positional_offset)


@DSUnembedRegistry.register_module
class RaggedUnembed(DSUnembedBase):
    @staticmethod
    def name():
        return "ragged_unembed"

    @staticmethod
    def supports_config(config: DSUnembedConfig) -> bool:
        return True

    def forward(self, hidden_states: torch.Tensor, vocab_embedding: torch.Tensor, last_rows: torch.Tensor,
                gamma: Optional[torch.Tensor] = None, beta: Optional[torch.Tensor] = None, bias: Optional[torch.Tensor] = None):
        """Gather each sequence's last token, optional final norm, vocabulary GEMM -> ``[n_seqs, vocab]``."""
        ops = get_ops(hidden_states.device)
        S = last_rows.shape[0]
        picked = torch.empty((S, hidden_states.shape[1]), dtype=hidden_states.dtype, device=hidden_states.device)
        ops.gather_rows(hidden_states, last_rows, picked)
        if self._config.norm_type is not None:
            normed = torch.empty_like(picked)
            ops.norm(picked, gamma, normed, beta=beta,
                     layer_norm=NormTypeEnum(self._config.norm_type) == NormTypeEnum.LayerNorm)
            picked = normed
        out = torch.empty((S, vocab_embedding.shape[0]), dtype=hidden_states.dtype, device=hidden_states.device)
        return ops.gemm(picked, vocab_embedding, out, bias=bias)


@DSAttentionRegistry.register_module
class DenseBlockedAttention(DSSelfAttentionBase):
    """KV append (+ rotate-half RoPE) into the paged cache, then ragged paged attention."""

    @staticmethod
    def name():
        return "dense_blocked_attention"

    @staticmethod
    def supports_config(config: DSSelfAttentionConfig) -> bool:
        pe = PositionalEmbeddingType(config.positional_embedding_type)
        return pe in (PositionalEmbeddingType.none, PositionalEmbeddingType.rotate_half) and \
            config.n_heads_q % config.n_heads_kv == 0

    def __init__(self, config, implementation_config=None):
        super().__init__(config, implementation_config)
        self.inv_freq, self.rot_dim = None, config.head_size
        if PositionalEmbeddingType(config.positional_embedding_type) == PositionalEmbeddingType.rotate_half:
            pc = config.positional_embedding_config
            self.rot_dim = pc.rotate_dim if (pc is not None and pc.rotate_dim) else config.head_size
            self.inv_freq = ref.default_inv_freq(self.rot_dim, pc.theta_base if pc is not None else 10000.0)

    def forward(self, q_k_v: torch.Tensor, kv_cache_layer: torch.Tensor, batch, inv_freqs: Optional[torch.Tensor] = None):
        c = self._config
        T = q_k_v.shape[0]
        nq, nkv, d = c.n_heads_q, c.n_heads_kv, c.head_size
        inv = inv_freqs if inv_freqs is not None else (self.inv_freq.to(q_k_v.device) if self.inv_freq is not None else None)
        q_out = torch.empty((T, nq * d), dtype=q_k_v.dtype, device=q_k_v.device)
        if q_k_v.is_cuda:
            from ....ops import kernels
            kernels.rope_kv_append(q_k_v, q_out, kv_cache_layer, batch.tok_pos(), batch.tok_slot(), inv, nq, nkv, d,
                                   self.rot_dim, batch.block_size)
        else:
            q, k, v = q_k_v.split([nq * d, nkv * d, nkv * d], dim=1)
            q, k, v = q.reshape(T, nq, d), k.reshape(T, nkv, d), v.reshape(T, nkv, d)
            if inv is not None:
                q = ref.rope_rotate_half(q, batch.tok_pos(), inv, self.rot_dim)
                k = ref.rope_rotate_half(k, batch.tok_pos(), inv, self.rot_dim)
            q_out.copy_(q.reshape(T, nq * d))
            ref.kv_append(kv_cache_layer, k, v, batch.tok_slot())
        out = torch.empty_like(q_out)
        scale = ref.attention_scale(d) * c.scale_factor
        get_ops(q_k_v.device).attention(q_out, out, kv_cache_layer, batch, nq, nkv, d, scale)
        return out


@DSMoERegistry.register_module
class MultiGemmMoE(DSMoEBase):
    """gate -> top-k -> experts -> weighted sum. Expert weights: ``mlp_1 [E, F(*2 if gated), H]``, ``mlp_2 [E, H, F]``."""

    @staticmethod
    def name():
        return "cutlass_multi_gemm_moe"

    @staticmethod
    def supports_config(config: DSMoEConfig) -> bool:
        return config.top_k in (1, 2, 4, 8) and config.top_k <= config.n_experts

    def forward(self, hidden_states, gate_w, mlp_1_w, mlp_2_w):
        c = self._config
        gated = c.activation in _GATED
        act = ACT_CODES[_GATED.get(c.activation, c.activation)]
        T, H = hidden_states.shape
        E, k = c.n_experts, c.top_k
        out = torch.zeros((T, H), dtype=torch.float32, device=hidden_states.device)
        if hidden_states.is_cuda:
            from ....ops import kernels as K
            dev = hidden_states.device
            logits = K.gemm(hidden_states, gate_w)
            scores = torch.empty(T, k, dtype=torch.float32, device=dev)
            ids, offs, slots = (torch.empty(T, k, dtype=torch.int32, device=dev) for _ in range(3))
            counts = torch.zeros(E, dtype=torch.int32, device=dev)
            pad_off = torch.zeros(E + 1, dtype=torch.int32, device=dev)
            rows = K.moe_padded_rows(T, E, k)
            xs = torch.zeros(rows, H, dtype=hidden_states.dtype, device=dev)
            hs = torch.zeros(rows, c.intermediate_features, dtype=hidden_states.dtype, device=dev)
            ys = torch.zeros(rows, H, dtype=hidden_states.dtype, device=dev)
            K.moe_topk_gating(logits, k, c.normalize_scores, scores, ids, offs, counts)
            K.moe_scatter(hidden_states, xs, ids, offs, counts, pad_off, slots, k)
            K.moe_gemm(xs, mlp_1_w.reshape(-1, H), pad_off, E, hs, act=act, gated=gated)
            K.moe_gemm(hs, mlp_2_w.reshape(-1, c.intermediate_features), pad_off, E, ys)
            K.moe_gather(ys, slots, scores, out)
        else:
            probs = torch.softmax(ref.gemm(hidden_states, gate_w, out_dtype=torch.float32), dim=-1)
            topv, topi = torch.topk(probs, k, dim=-1)
            if c.normalize_scores:
                topv = topv / topv.sum(-1, keepdim=True)
            for e in range(E):
                sel = topi == e
                rows = sel.any(-1).nonzero().flatten()
                if rows.numel() == 0:
                    continue
                h = ref.gemm(hidden_states[rows], mlp_1_w[e], act=act, gated=gated)
                y = ref.gemm(h, mlp_2_w[e], out_dtype=torch.float32)
                out.index_add_(0, rows, y * (topv * sel).sum(-1)[rows, None])
        return out.to(hidden_states.dtype)
