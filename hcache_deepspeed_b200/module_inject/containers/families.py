"""Per-family policies for v1 kernel injection: HF block -> :class:`LayerTensors`.
Reference: ``deepspeed/module_inject/containers/{gpt2,gptneox,gptj,gptneo,bert,distil_bert,opt,bloom}.py`` (each a
container class plus an ``HF*LayerPolicy`` with ``attention()``, ``mlp()``, ``layernorm()``)."""
from __future__ import annotations

import torch

from .base import LayerTensors, register_policy, t


def _cat(*xs):
    return None if any(x is None for x in xs) else torch.cat([x.detach() for x in xs], dim=-1)


def _b(mod):
    return None if getattr(mod, "bias", None) is None else mod.bias.detach()


def _per_head_to_qkv(w: torch.Tensor, heads: int) -> torch.Tensor:
    """Fused QKV stored per head as [heads, 3, d] along the OUTPUT dim -> [q | k | v] (works for [.., out] and [out])."""
    d = w.shape[-1] // (3 * heads)
    v = w.reshape(*w.shape[:-1], heads, 3, d)
    return v.transpose(-3, -2).reshape(w.shape)


def _rope_theta(cfg, default=10000.0):
    rp = getattr(cfg, "rope_parameters", None)
    if isinstance(rp, dict) and "rope_theta" in rp:
        return float(rp["rope_theta"])
    return float(getattr(cfg, "rope_theta", getattr(cfg, "rotary_emb_base", default)))


_ACT = {"gelu_new": "gelu_new", "gelu_pytorch_tanh": "gelu_new", "gelu_fast": "gelu_new", "gelu": "gelu", "relu": "relu",
        "silu": "silu", "swish": "silu"}


@register_policy("GPT2Block", "gpt2")
def gpt2(block, cfg) -> LayerTensors:
    a, m = block.attn, block.mlp           # Conv1D modules already store [in, out]
    return LayerTensors(
        qkvw=a.c_attn.weight.detach(), qkvb=_b(a.c_attn), ow=a.c_proj.weight.detach(), ob=_b(a.c_proj),
        inter_w=m.c_fc.weight.detach(), inter_b=_b(m.c_fc), out_w=m.c_proj.weight.detach(), out_b=_b(m.c_proj),
        attn_nw=block.ln_2.weight.detach(), attn_nb=_b(block.ln_2), input_nw=block.ln_1.weight.detach(), input_nb=_b(block.ln_1),
        hidden_size=cfg.n_embd, heads=cfg.n_head, eps=cfg.layer_norm_epsilon, act=_ACT.get(cfg.activation_function, "gelu"),
        scale_attn_by_inverse_layer_idx=bool(getattr(cfg, "scale_attn_by_inverse_layer_idx", False)))


@register_policy("GPTNeoXLayer", "gptneox")
def gptneox(block, cfg) -> LayerTensors:
    a, m, h = block.attention, block.mlp, cfg.num_attention_heads
    d = cfg.hidden_size // h
    qkvw = _per_head_to_qkv(t(a.query_key_value.weight), h)
    qkvb = None if a.query_key_value.bias is None else _per_head_to_qkv(a.query_key_value.bias.detach(), h)
    rp = getattr(cfg, "rope_parameters", None) or {}
    pct = rp.get("partial_rotary_factor", getattr(cfg, "rotary_pct", 1.0)) if isinstance(rp, dict) else getattr(cfg, "rotary_pct", 1.0)
    return LayerTensors(
        qkvw=qkvw, qkvb=qkvb, ow=t(a.dense.weight), ob=_b(a.dense),
        inter_w=t(m.dense_h_to_4h.weight), inter_b=_b(m.dense_h_to_4h), out_w=t(m.dense_4h_to_h.weight), out_b=_b(m.dense_4h_to_h),
        attn_nw=block.post_attention_layernorm.weight.detach(), attn_nb=_b(block.post_attention_layernorm),
        input_nw=block.input_layernorm.weight.detach(), input_nb=_b(block.input_layernorm),
        hidden_size=cfg.hidden_size, heads=h, eps=cfg.layer_norm_eps, act=_ACT.get(cfg.hidden_act, "gelu"),
        mlp_after_attn=not getattr(cfg, "use_parallel_residual", True), rotary_dim=int(d * pct), rotate_half=True,
        rope_theta=_rope_theta(cfg))


@register_policy("GPTJBlock", "gptj")
def gptj(block, cfg) -> LayerTensors:
    a, m = block.attn, block.mlp
    return LayerTensors(
        qkvw=_cat(t(a.q_proj.weight), t(a.k_proj.weight), t(a.v_proj.weight)), qkvb=None, ow=t(a.out_proj.weight), ob=None,
        inter_w=t(m.fc_in.weight), inter_b=_b(m.fc_in), out_w=t(m.fc_out.weight), out_b=_b(m.fc_out),
        attn_nw=None, attn_nb=None, input_nw=block.ln_1.weight.detach(), input_nb=_b(block.ln_1),
        hidden_size=cfg.n_embd, heads=cfg.n_head, eps=cfg.layer_norm_epsilon, act=_ACT.get(cfg.activation_function, "gelu_new"),
        mlp_after_attn=False, rotary_dim=cfg.rotary_dim or cfg.n_embd // cfg.n_head, rotate_every_two=True,
        rope_theta=_rope_theta(cfg))


@register_policy("GPTNeoBlock", "gptneo")
def gptneo(block, cfg) -> LayerTensors:
    a, m = block.attn.attention, block.mlp
    lt = LayerTensors(
        qkvw=_cat(t(a.q_proj.weight), t(a.k_proj.weight), t(a.v_proj.weight)), qkvb=None, ow=t(a.out_proj.weight), ob=_b(a.out_proj),
        inter_w=t(m.c_fc.weight), inter_b=_b(m.c_fc), out_w=t(m.c_proj.weight), out_b=_b(m.c_proj),
        attn_nw=block.ln_2.weight.detach(), attn_nb=_b(block.ln_2), input_nw=block.ln_1.weight.detach(), input_nb=_b(block.ln_1),
        hidden_size=cfg.hidden_size, heads=cfg.num_heads, eps=cfg.layer_norm_epsilon, act=_ACT.get(cfg.activation_function, "gelu_new"))
    lt.scale_attention = False             # GPT-Neo does not divide the scores by sqrt(d)
    return lt


@register_policy("OPTDecoderLayer", "opt")
def opt(block, cfg) -> LayerTensors:
    a = block.self_attn
    return LayerTensors(
        qkvw=_cat(t(a.q_proj.weight), t(a.k_proj.weight), t(a.v_proj.weight)), qkvb=_cat(_b(a.q_proj), _b(a.k_proj), _b(a.v_proj)),
        ow=t(a.out_proj.weight), ob=_b(a.out_proj), inter_w=t(block.fc1.weight), inter_b=_b(block.fc1),
        out_w=t(block.fc2.weight), out_b=_b(block.fc2),
        attn_nw=block.final_layer_norm.weight.detach(), attn_nb=_b(block.final_layer_norm),
        input_nw=block.self_attn_layer_norm.weight.detach(), input_nb=_b(block.self_attn_layer_norm),
        hidden_size=cfg.hidden_size, heads=cfg.num_attention_heads, eps=1e-5, act=_ACT.get(cfg.activation_function, "relu"),
        pre_layer_norm=bool(cfg.do_layer_norm_before))


@register_policy("BloomBlock", "bloom")
def bloom(block, cfg) -> LayerTensors:
    a, m, h = block.self_attention, block.mlp, cfg.n_head
    qkvb = None if a.query_key_value.bias is None else _per_head_to_qkv(a.query_key_value.bias.detach(), h)
    return LayerTensors(
        qkvw=_per_head_to_qkv(t(a.query_key_value.weight), h), qkvb=qkvb, ow=t(a.dense.weight), ob=_b(a.dense),
        inter_w=t(m.dense_h_to_4h.weight), inter_b=_b(m.dense_h_to_4h), out_w=t(m.dense_4h_to_h.weight), out_b=_b(m.dense_4h_to_h),
        attn_nw=block.post_attention_layernorm.weight.detach(), attn_nb=_b(block.post_attention_layernorm),
        input_nw=block.input_layernorm.weight.detach(), input_nb=_b(block.input_layernorm),
        hidden_size=cfg.hidden_size, heads=h, eps=cfg.layer_norm_epsilon, act="gelu_new", use_alibi=True)


@register_policy("BertLayer", "bert")
def bert(block, cfg) -> LayerTensors:
    s, o = block.attention.self, block.attention.output
    return LayerTensors(
        qkvw=_cat(t(s.query.weight), t(s.key.weight), t(s.value.weight)), qkvb=_cat(_b(s.query), _b(s.key), _b(s.value)),
        ow=t(o.dense.weight), ob=_b(o.dense), inter_w=t(block.intermediate.dense.weight), inter_b=_b(block.intermediate.dense),
        out_w=t(block.output.dense.weight), out_b=_b(block.output.dense),
        attn_nw=o.LayerNorm.weight.detach(), attn_nb=_b(o.LayerNorm),
        input_nw=block.output.LayerNorm.weight.detach(), input_nb=_b(block.output.LayerNorm),
        hidden_size=cfg.hidden_size, heads=cfg.num_attention_heads, eps=cfg.layer_norm_eps, act=_ACT.get(cfg.hidden_act, "gelu"),
        pre_layer_norm=False, triangular_masking=False)


@register_policy("TransformerBlock", "distil_bert")
def distil_bert(block, cfg) -> LayerTensors:
    a, f = block.attention, block.ffn
    return LayerTensors(
        qkvw=_cat(t(a.q_lin.weight), t(a.k_lin.weight), t(a.v_lin.weight)), qkvb=_cat(_b(a.q_lin), _b(a.k_lin), _b(a.v_lin)),
        ow=t(a.out_lin.weight), ob=_b(a.out_lin), inter_w=t(f.lin1.weight), inter_b=_b(f.lin1), out_w=t(f.lin2.weight),
        out_b=_b(f.lin2), attn_nw=block.sa_layer_norm.weight.detach(), attn_nb=_b(block.sa_layer_norm),
        input_nw=block.output_layer_norm.weight.detach(), input_nb=_b(block.output_layer_norm),
        hidden_size=cfg.dim, heads=cfg.n_heads, eps=1e-12, act=_ACT.get(cfg.activation, "gelu"), pre_layer_norm=False,
        triangular_masking=False)


def _llama_like(block, cfg) -> LayerTensors:
    """Llama / Mistral / Qwen2 / InternLM decoder layers: RMSNorm, rotate-half RoPE over the whole head, grouped KV
    heads, SiLU-gated MLP (the fused layer wants ``[up | gate]`` columns: ``up * silu(gate)``)."""
    a, m = block.self_attn, block.mlp
    h = cfg.num_attention_heads
    kv = getattr(cfg, "num_key_value_heads", None) or h
    d = getattr(cfg, "head_dim", None) or cfg.hidden_size // h
    bias = [x.bias for x in (a.q_proj, a.k_proj, a.v_proj)]
    qkvb = None if all(b is None for b in bias) else _cat(*[
        (b.detach() if b is not None else torch.zeros(x.weight.shape[0], dtype=x.weight.dtype, device=x.weight.device))
        for b, x in zip(bias, (a.q_proj, a.k_proj, a.v_proj))])
    return LayerTensors(
        qkvw=_cat(t(a.q_proj.weight), t(a.k_proj.weight), t(a.v_proj.weight)), qkvb=qkvb, ow=t(a.o_proj.weight), ob=_b(a.o_proj),
        inter_w=_cat(t(m.up_proj.weight), t(m.gate_proj.weight)), inter_b=None, out_w=t(m.down_proj.weight), out_b=_b(m.down_proj),
        attn_nw=block.post_attention_layernorm.weight.detach(), attn_nb=None,
        input_nw=block.input_layernorm.weight.detach(), input_nb=None,
        hidden_size=cfg.hidden_size, heads=h, num_kv=kv, eps=cfg.rms_norm_eps, act="gated_silu", norm_type="rms_norm",
        rotary_dim=d, rotate_half=True, rope_theta=_rope_theta(cfg))


for _cls, _family in (("LlamaDecoderLayer", "llama"), ("MistralDecoderLayer", "mistral"), ("Qwen2DecoderLayer", "qwen2"),
                      ("InternLMDecoderLayer", "internlm")):
    register_policy(_cls, _family)(_llama_like)


@register_policy("ParallelTransformerLayer", "megatron_gpt")
def megatron_gpt(block, cfg) -> LayerTensors:
    """Megatron-LM GPT layer (``containers/megatron_gpt.py`` of the reference): fused ``query_key_value`` stored per head
    as ``[heads, 3, d]`` (checkpoint version >= 2; ``cfg.megatron_v2=False`` means the older ``[3, heads, d]`` = already
    ``[q | k | v]``), pre-LN, GELU MLP, learned positions (no rotary in the layer)."""
    a = getattr(block, "attention", None) or block.self_attention
    m = block.mlp
    h = getattr(cfg, "num_attention_heads", None) or cfg.n_head
    hidden = getattr(cfg, "hidden_size", None) or cfg.n_embd
    qkvw, qkvb = t(a.query_key_value.weight), _b(a.query_key_value)
    if getattr(cfg, "megatron_v2", True):
        qkvw = _per_head_to_qkv(qkvw, h)
        qkvb = None if qkvb is None else _per_head_to_qkv(qkvb, h)
    return LayerTensors(
        qkvw=qkvw, qkvb=qkvb, ow=t(a.dense.weight), ob=_b(a.dense),
        inter_w=t(m.dense_h_to_4h.weight), inter_b=_b(m.dense_h_to_4h), out_w=t(m.dense_4h_to_h.weight), out_b=_b(m.dense_4h_to_h),
        attn_nw=block.post_attention_layernorm.weight.detach(), attn_nb=_b(block.post_attention_layernorm),
        input_nw=block.input_layernorm.weight.detach(), input_nb=_b(block.input_layernorm),
        hidden_size=hidden, heads=h, eps=getattr(cfg, "layernorm_epsilon", 1e-5), act="gelu")
