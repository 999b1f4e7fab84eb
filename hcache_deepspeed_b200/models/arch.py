"""Declarative description of a decoder-only transformer family.

The reference spends one ``container.py`` + ``model.py`` + ``policy.py`` per family on parameter
metaprogramming (``deepspeed/inference/v2/model_implementations/{llama_v2,opt,mistral,...}/``,
``layer_container_base.py:42-355``, ``parameter_base.py:54-255``). Here a family is an
:class:`ArchSpec` (what the network looks like) plus a weight-name map (where its tensors live in a
HuggingFace checkpoint); ONE executor (:mod:`hcache_deepspeed_b200.models.transformer`) runs every
spec on the sm_100a kernels.
"""
from __future__ import annotations

from dataclasses import dataclass, field, replace
from typing import Any, Callable, Dict, List, Optional


@dataclass
class ArchSpec:
    model_type: str
    n_layers: int
    hidden: int
    n_heads: int
    n_kv_heads: int
    head_dim: int
    intermediate: int
    vocab: int
    # normalisation
    norm: str = "rms"  # "rms" | "layer"
    norm_eps: float = 1e-5
    # MLP
    act: str = "silu"
    gated_mlp: bool = True
    # positions
    pos_emb: str = "rope"  # "rope" | "learned" | "alibi" | "none"
    alibi_scale: float = 1.0  # multiplies the ALiBi slopes (Falcon adds the bias BEFORE its 1/sqrt(d): scale = 1/sqrt(d))
    embed_norm: bool = False  # LayerNorm on the token embeddings (Bloom's word_embeddings_layernorm)
    sliding_window: int = 0  # > 0: a query at position p attends to the keys (p - window, p] (Mistral, Phi-3)
    layer_windows: Optional[List[int]] = None  # per-layer override of sliding_window (GPT-Neo: global / local alternate)
    rope_theta: float = 10000.0
    rotary_dim: Optional[int] = None  # None = head_dim
    rope_scaling: Optional[Dict[str, Any]] = None
    learned_pos_offset: int = 0  # OPT: 2
    max_positions: int = 4096
    # biases
    qkv_bias: bool = False
    attn_out_bias: bool = False
    mlp_bias: bool = False
    lm_head_bias: bool = False
    norm_bias: bool = False  # LayerNorm beta present
    # topology
    parallel_residual: bool = False  # attention and MLP both read the same pre-norm (Falcon, Phi)
    parallel_two_norms: bool = False  # Falcon new architecture: separate norms for the two branches
    tie_word_embeddings: bool = False
    final_norm: bool = True
    attn_scale: Optional[float] = None
    # mixture of experts (0 experts = dense)
    n_experts: int = 0
    top_k: int = 0
    moe_intermediate: int = 0
    shared_expert_intermediate: int = 0  # Qwen2-MoE shared expert (with sigmoid gate)
    norm_topk_prob: bool = True
    extras: Dict[str, Any] = field(default_factory=dict)

    # ---- derived -------------------------------------------------------------------------------
    @property
    def rot_dim(self) -> int:
        return self.rotary_dim if self.rotary_dim is not None else self.head_dim

    @property
    def q_dim(self) -> int:
        return self.n_heads * self.head_dim

    @property
    def kv_dim(self) -> int:
        return self.n_kv_heads * self.head_dim

    @property
    def is_moe(self) -> bool:
        return self.n_experts > 0

    def hidden_state_bytes_per_token_layer(self, elem: int = 2) -> int:
        """Bytes HCache stores per token per layer (the QKV GEMM input)."""
        return self.hidden * elem

    def kv_bytes_per_token_layer(self, elem: int = 2) -> int:
        """Bytes plain KV offload stores per token per layer."""
        return 2 * self.kv_dim * elem

    def window_of(self, layer: int) -> int:
        return int(self.layer_windows[layer]) if self.layer_windows is not None else int(self.sliding_window)

    def scaled(self, **kw) -> "ArchSpec":
        return replace(self, **kw)

    def param_count(self) -> int:
        h, i = self.hidden, self.intermediate
        attn = h * (self.q_dim + 2 * self.kv_dim) + self.q_dim * h
        if self.is_moe:
            mlp = self.n_experts * (3 if self.gated_mlp else 2) * h * self.moe_intermediate + h * self.n_experts
            if self.shared_expert_intermediate:
                mlp += 3 * h * self.shared_expert_intermediate + h
        else:
            mlp = (3 if self.gated_mlp else 2) * h * i
        emb = self.vocab * h * (1 if self.tie_word_embeddings else 2)
        return self.n_layers * (attn + mlp + 2 * h) + emb + h


#: HF ``model_type`` -> function(hf_config) -> ArchSpec
ARCH_BUILDERS: Dict[str, Callable[[Any], ArchSpec]] = {}


def register_arch(*model_types: str):

    def deco(fn):
        for mt in model_types:
            ARCH_BUILDERS[mt] = fn
        return fn

    return deco


def _get(cfg: Any, *names: str, default=None):
    for n in names:
        if isinstance(cfg, dict):
            if n in cfg and cfg[n] is not None:
                return cfg[n]
        elif getattr(cfg, n, None) is not None:
            return getattr(cfg, n)
    return default


def arch_from_hf_config(cfg: Any) -> ArchSpec:
    mt = _get(cfg, "model_type")
    if mt not in ARCH_BUILDERS:
        raise ValueError(f"Unsupported model type {mt}; supported: {sorted(ARCH_BUILDERS)}")
    return ARCH_BUILDERS[mt](cfg)


def supported_model_types() -> List[str]:
    return sorted(ARCH_BUILDERS)


# ---------------------------------------------------------------------------------------------------
# family builders (ref: each family's model.py property block, e.g. llama_v2/model.py:36-116)
# ---------------------------------------------------------------------------------------------------
def _rope_theta(cfg, default=10000.0) -> float:
    theta = _get(cfg, "rope_theta")
    if theta is None:
        rp = _get(cfg, "rope_parameters", "rope_scaling")
        if isinstance(rp, dict) and rp.get("rope_theta") is not None:
            theta = rp["rope_theta"]
    return float(theta if theta is not None else default)


def _window_fields(cfg, mt) -> Dict[str, Any]:
    """Sliding-window settings of the llama-like families. Mistral / Phi-3: ``sliding_window`` on every layer. Qwen2:
    only with ``use_sliding_window``, and then on the layers HF marks ``sliding_attention`` (``layer_types``) or, for
    older configs, on layers >= ``max_window_layers``."""
    w = _get(cfg, "sliding_window")
    if not w or mt not in ("mistral", "phi3", "qwen2", "mixtral", "qwen2_moe"):
        return {}
    n = _get(cfg, "num_hidden_layers")
    if mt in ("qwen2", "qwen2_moe"):
        if not _get(cfg, "use_sliding_window", default=False):
            return {}
        types = _get(cfg, "layer_types")
        if types:
            return {"layer_windows": [int(w) if t == "sliding_attention" else 0 for t in types]}
        first = _get(cfg, "max_window_layers", default=n)
        return {"layer_windows": [int(w) if i >= first else 0 for i in range(n)]}
    types = _get(cfg, "layer_types")
    if types and any(t != "sliding_attention" for t in types):
        return {"layer_windows": [int(w) if t == "sliding_attention" else 0 for t in types]}
    return {"sliding_window": int(w)}


@register_arch("llama", "mistral", "qwen2", "phi3", "internlm")
def _llama_like(cfg) -> ArchSpec:
    mt = _get(cfg, "model_type")
    h = _get(cfg, "hidden_size")
    nh = _get(cfg, "num_attention_heads")
    return ArchSpec(
        model_type=mt, n_layers=_get(cfg, "num_hidden_layers"), hidden=h, n_heads=nh,
        n_kv_heads=_get(cfg, "num_key_value_heads", default=nh), head_dim=_get(cfg, "head_dim", default=h // nh),
        intermediate=_get(cfg, "intermediate_size"), vocab=_get(cfg, "vocab_size"), norm="rms",
        norm_eps=_get(cfg, "rms_norm_eps", default=1e-5), act=_get(cfg, "hidden_act", default="silu"), gated_mlp=True,
        pos_emb="rope", rope_theta=_rope_theta(cfg), max_positions=_get(cfg, "max_position_embeddings", default=4096),
        qkv_bias=(mt == "qwen2") or bool(_get(cfg, "attention_bias", default=False)),
        attn_out_bias=bool(_get(cfg, "attention_bias", default=False)) and mt != "qwen2",
        mlp_bias=bool(_get(cfg, "mlp_bias", default=False)),
        tie_word_embeddings=bool(_get(cfg, "tie_word_embeddings", default=False)), **_window_fields(cfg, mt))


@register_arch("opt")
def _opt(cfg) -> ArchSpec:
    h = _get(cfg, "hidden_size")
    nh = _get(cfg, "num_attention_heads")
    if _get(cfg, "word_embed_proj_dim", default=h) != h:
        raise ValueError("OPT variants with word_embed_proj_dim != hidden_size (opt-350m) are not supported")
    if not _get(cfg, "do_layer_norm_before", default=True):
        raise ValueError("post-LayerNorm OPT variants are not supported")
    return ArchSpec(
        model_type="opt", n_layers=_get(cfg, "num_hidden_layers"), hidden=h, n_heads=nh, n_kv_heads=nh,
        head_dim=h // nh, intermediate=_get(cfg, "ffn_dim"), vocab=_get(cfg, "vocab_size"), norm="layer", norm_eps=1e-5,
        act=_get(cfg, "activation_function", default="relu"), gated_mlp=False, pos_emb="learned", learned_pos_offset=2,
        max_positions=_get(cfg, "max_position_embeddings", default=2048), qkv_bias=bool(_get(cfg, "enable_bias", default=True)),
        attn_out_bias=bool(_get(cfg, "enable_bias", default=True)), mlp_bias=bool(_get(cfg, "enable_bias", default=True)),
        norm_bias=True, tie_word_embeddings=True, final_norm=True)


@register_arch("falcon")
def _falcon(cfg) -> ArchSpec:
    h = _get(cfg, "hidden_size")
    nh = _get(cfg, "num_attention_heads")
    new_arch = bool(_get(cfg, "new_decoder_architecture", default=False))
    if new_arch:
        nkv = _get(cfg, "num_kv_heads", default=nh)
    else:
        nkv = 1 if _get(cfg, "multi_query", default=True) else nh
    alibi = bool(_get(cfg, "alibi", default=False))
    return ArchSpec(
        model_type="falcon", n_layers=_get(cfg, "num_hidden_layers"), hidden=h, n_heads=nh, n_kv_heads=nkv,
        head_dim=h // nh, intermediate=_get(cfg, "ffn_hidden_size", default=4 * h), vocab=_get(cfg, "vocab_size"),
        norm="layer", norm_eps=_get(cfg, "layer_norm_epsilon", default=1e-5), act="gelu", gated_mlp=False,
        pos_emb="alibi" if alibi else "rope", alibi_scale=(h // nh) ** -0.5 if alibi else 1.0, rope_theta=_rope_theta(cfg), max_positions=_get(cfg, "max_position_embeddings", default=2048),
        qkv_bias=bool(_get(cfg, "bias", default=False)), attn_out_bias=bool(_get(cfg, "bias", default=False)),
        mlp_bias=bool(_get(cfg, "bias", default=False)), norm_bias=True,
        parallel_residual=bool(_get(cfg, "parallel_attn", default=True)), parallel_two_norms=new_arch,
        tie_word_embeddings=True, extras={"new_decoder_architecture": new_arch})


@register_arch("phi")
def _phi(cfg) -> ArchSpec:
    h = _get(cfg, "hidden_size")
    nh = _get(cfg, "num_attention_heads")
    hd = h // nh
    return ArchSpec(
        model_type="phi", n_layers=_get(cfg, "num_hidden_layers"), hidden=h, n_heads=nh,
        n_kv_heads=_get(cfg, "num_key_value_heads", default=nh), head_dim=hd, intermediate=_get(cfg, "intermediate_size"),
        vocab=_get(cfg, "vocab_size"), norm="layer", norm_eps=_get(cfg, "layer_norm_eps", default=1e-5),
        act=_get(cfg, "hidden_act", default="gelu_new"), gated_mlp=False, pos_emb="rope", rope_theta=_rope_theta(cfg),
        rotary_dim=int(hd * _get(cfg, "partial_rotary_factor", default=0.5)),
        max_positions=_get(cfg, "max_position_embeddings", default=2048), qkv_bias=True, attn_out_bias=True, mlp_bias=True,
        lm_head_bias=True, norm_bias=True, parallel_residual=True)


@register_arch("qwen")
def _qwen(cfg) -> ArchSpec:
    h = _get(cfg, "hidden_size")
    nh = _get(cfg, "num_attention_heads")
    return ArchSpec(
        model_type="qwen", n_layers=_get(cfg, "num_hidden_layers"), hidden=h, n_heads=nh, n_kv_heads=nh,
        head_dim=_get(cfg, "kv_channels", default=h // nh), intermediate=_get(cfg, "intermediate_size") // 2,
        vocab=_get(cfg, "vocab_size"), norm="rms", norm_eps=_get(cfg, "layer_norm_epsilon", default=1e-6), act="silu",
        gated_mlp=True, pos_emb="rope", rope_theta=float(_get(cfg, "rotary_emb_base", default=10000.0)),
        max_positions=_get(cfg, "seq_length", "max_position_embeddings", default=8192), qkv_bias=True)


# ---- families the reference serves only through v1 kernel injection (module_inject/containers/{gpt2,gptj,gptneox}.py).
# Here they are three more specs for the same executor, so init_inference(replace_with_kernel_inject=True) runs them
# on the sm_100a kernels with paged KV, and HCache restore works for them too.
@register_arch("gpt2")
def _gpt2(cfg) -> ArchSpec:
    h = _get(cfg, "n_embd", "hidden_size")
    nh = _get(cfg, "n_head", "num_attention_heads")
    if _get(cfg, "scale_attn_by_inverse_layer_idx", default=False) or not _get(cfg, "scale_attn_weights", default=True):
        raise ValueError("GPT-2 variants with non-standard attention scaling are not supported")
    return ArchSpec(
        model_type="gpt2", n_layers=_get(cfg, "n_layer", "num_hidden_layers"), hidden=h, n_heads=nh, n_kv_heads=nh,
        head_dim=h // nh, intermediate=_get(cfg, "n_inner", default=4 * h), vocab=_get(cfg, "vocab_size"), norm="layer",
        norm_eps=_get(cfg, "layer_norm_epsilon", default=1e-5), act=_get(cfg, "activation_function", default="gelu_new"),
        gated_mlp=False, pos_emb="learned", learned_pos_offset=0, max_positions=_get(cfg, "n_positions", default=1024),
        qkv_bias=True, attn_out_bias=True, mlp_bias=True, norm_bias=True, tie_word_embeddings=True)


@register_arch("gpt_neo")
def _gpt_neo(cfg) -> ArchSpec:
    """GPT-Neo: learned positions, UNSCALED attention scores, layers alternate between global and local (sliding-window)
    attention as listed in ``attention_layers``."""
    h = _get(cfg, "hidden_size")
    nh = _get(cfg, "num_heads", "num_attention_heads")
    n = _get(cfg, "num_layers", "num_hidden_layers")
    kinds = list(_get(cfg, "attention_layers", default=["global"] * n))
    win = int(_get(cfg, "window_size", default=256))
    return ArchSpec(
        model_type="gpt_neo", n_layers=n, hidden=h, n_heads=nh, n_kv_heads=nh, head_dim=h // nh,
        intermediate=_get(cfg, "intermediate_size", default=None) or 4 * h, vocab=_get(cfg, "vocab_size"), norm="layer",
        norm_eps=_get(cfg, "layer_norm_epsilon", default=1e-5), act=_get(cfg, "activation_function", default="gelu_new"),
        gated_mlp=False, pos_emb="learned", learned_pos_offset=0, max_positions=_get(cfg, "max_position_embeddings", default=2048),
        qkv_bias=False, attn_out_bias=True, mlp_bias=True, norm_bias=True, tie_word_embeddings=True, attn_scale=1.0,
        layer_windows=[win if k == "local" else 0 for k in kinds])


@register_arch("gptj")
def _gptj(cfg) -> ArchSpec:
    h = _get(cfg, "n_embd", "hidden_size")
    nh = _get(cfg, "n_head", "num_attention_heads")
    hd = h // nh
    return ArchSpec(
        model_type="gptj", n_layers=_get(cfg, "n_layer", "num_hidden_layers"), hidden=h, n_heads=nh, n_kv_heads=nh,
        head_dim=hd, intermediate=_get(cfg, "n_inner", default=4 * h), vocab=_get(cfg, "vocab_size"), norm="layer",
        norm_eps=_get(cfg, "layer_norm_epsilon", default=1e-5), act=_get(cfg, "activation_function", default="gelu_new"),
        gated_mlp=False, pos_emb="rope", rope_theta=10000.0, rotary_dim=_get(cfg, "rotary_dim", default=hd),
        max_positions=_get(cfg, "n_positions", default=2048), qkv_bias=False, attn_out_bias=False, mlp_bias=True,
        lm_head_bias=True, norm_bias=True, parallel_residual=True,
        tie_word_embeddings=bool(_get(cfg, "tie_word_embeddings", default=False)),
        extras={"rope_interleaved_checkpoint": True})     # rows are re-ordered once at load (families.map_gptj)


@register_arch("gpt_neox")
def _gpt_neox(cfg) -> ArchSpec:
    h = _get(cfg, "hidden_size")
    nh = _get(cfg, "num_attention_heads")
    hd = h // nh
    pct = _get(cfg, "rotary_pct", "partial_rotary_factor")
    if pct is None:
        rp = _get(cfg, "rope_parameters", "rope_scaling")
        pct = rp.get("partial_rotary_factor") if isinstance(rp, dict) else None
    pct = 0.25 if pct is None else float(pct)
    theta = _get(cfg, "rotary_emb_base")
    par = bool(_get(cfg, "use_parallel_residual", default=True))
    return ArchSpec(
        model_type="gpt_neox", n_layers=_get(cfg, "num_hidden_layers"), hidden=h, n_heads=nh, n_kv_heads=nh, head_dim=hd,
        intermediate=_get(cfg, "intermediate_size", default=4 * h), vocab=_get(cfg, "vocab_size"), norm="layer",
        norm_eps=_get(cfg, "layer_norm_eps", default=1e-5), act=_get(cfg, "hidden_act", default="gelu"), gated_mlp=False,
        pos_emb="rope", rope_theta=float(theta) if theta is not None else _rope_theta(cfg), rotary_dim=int(hd * pct),
        max_positions=_get(cfg, "max_position_embeddings", default=2048),
        qkv_bias=bool(_get(cfg, "attention_bias", default=True)), attn_out_bias=bool(_get(cfg, "attention_bias", default=True)),
        mlp_bias=True, norm_bias=True, parallel_residual=par, parallel_two_norms=par,
        tie_word_embeddings=bool(_get(cfg, "tie_word_embeddings", default=False)))


@register_arch("bloom")
def _bloom(cfg) -> ArchSpec:
    h = _get(cfg, "hidden_size", "n_embed")
    nh = _get(cfg, "n_head", "num_attention_heads")
    if _get(cfg, "apply_residual_connection_post_layernorm", default=False):
        raise ValueError("Bloom variants with the residual taken after the LayerNorm are not supported")
    return ArchSpec(
        model_type="bloom", n_layers=_get(cfg, "n_layer", "num_hidden_layers"), hidden=h, n_heads=nh, n_kv_heads=nh,
        head_dim=h // nh, intermediate=4 * h, vocab=_get(cfg, "vocab_size"), norm="layer",
        norm_eps=_get(cfg, "layer_norm_epsilon", default=1e-5), act="gelu_tanh", gated_mlp=False, pos_emb="alibi",
        max_positions=_get(cfg, "seq_length", "max_position_embeddings", default=2048), qkv_bias=True, attn_out_bias=True,
        mlp_bias=True, norm_bias=True, embed_norm=True, tie_word_embeddings=True)


@register_arch("mixtral")
def _mixtral(cfg) -> ArchSpec:
    base = _llama_like(cfg)
    return base.scaled(model_type="mixtral", n_experts=_get(cfg, "num_local_experts"), top_k=_get(cfg, "num_experts_per_tok"),
                       moe_intermediate=_get(cfg, "intermediate_size"), norm_topk_prob=True)


@register_arch("qwen2_moe")
def _qwen2_moe(cfg) -> ArchSpec:
    base = _llama_like(cfg)
    return base.scaled(model_type="qwen2_moe", qkv_bias=True, attn_out_bias=False, n_experts=_get(cfg, "num_experts"),
                       top_k=_get(cfg, "num_experts_per_tok"), moe_intermediate=_get(cfg, "moe_intermediate_size"),
                       shared_expert_intermediate=_get(cfg, "shared_expert_intermediate_size", default=0),
                       norm_topk_prob=bool(_get(cfg, "norm_topk_prob", default=False)))


# ---------------------------------------------------------------------------------------------------
# named configurations used by benchmarks and tests (random-init weights; shapes of the public models)
# ---------------------------------------------------------------------------------------------------
def named_arch(name: str) -> ArchSpec:
    n = name.lower().replace("_", "-")
    llama = dict(model_type="llama", norm="rms", norm_eps=1e-5, act="silu", gated_mlp=True, pos_emb="rope",
                 max_positions=4096, vocab=32000, head_dim=128)
    opt = dict(model_type="opt", norm="layer", norm_eps=1e-5, act="relu", gated_mlp=False, pos_emb="learned",
               learned_pos_offset=2, max_positions=2048, vocab=50272, qkv_bias=True, attn_out_bias=True, mlp_bias=True,
               norm_bias=True, tie_word_embeddings=True)
    table = {
        "llama-2-7b": ArchSpec(n_layers=32, hidden=4096, n_heads=32, n_kv_heads=32, intermediate=11008, **llama),
        "llama-2-13b": ArchSpec(n_layers=40, hidden=5120, n_heads=40, n_kv_heads=40, intermediate=13824, **llama),
        "llama-2-70b": ArchSpec(n_layers=80, hidden=8192, n_heads=64, n_kv_heads=8, intermediate=28672, **llama),
        "opt-125m": ArchSpec(n_layers=12, hidden=768, n_heads=12, n_kv_heads=12, head_dim=64, intermediate=3072, **opt),
        "opt-30b": ArchSpec(n_layers=48, hidden=7168, n_heads=56, n_kv_heads=56, head_dim=128, intermediate=28672, **opt),
        "tiny-llama-test": ArchSpec(n_layers=2, hidden=256, n_heads=4, n_kv_heads=2, intermediate=512,
                                    **{**llama, "vocab": 512, "head_dim": 64}),
        "tiny-mixtral-test": ArchSpec(n_layers=2, hidden=256, n_heads=4, n_kv_heads=2, intermediate=512,
                                      **{**llama, "vocab": 512, "head_dim": 64, "model_type": "mixtral"}, n_experts=8,
                                      top_k=2, moe_intermediate=384, norm_topk_prob=True),
        "tiny-opt-test": ArchSpec(n_layers=2, hidden=256, n_heads=4, n_kv_heads=4, intermediate=512,
                                  **{**opt, "vocab": 512, "head_dim": 64, "max_positions": 512}),
    }
    if n not in table:
        raise KeyError(f"unknown named architecture {name}; known: {sorted(table)}")
    return table[n]
