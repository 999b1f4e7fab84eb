"""Logit parity with HuggingFace reference implementations on CPU (fp32), one tiny random model per family.
Covers the family weight mappers + ArchSpec builders + executor semantics (the role of the reference's
per-model container tests and of tests/unit/inference/test_inference.py's before/after comparison)."""
import pytest
import torch

transformers = pytest.importorskip("transformers")

from hcache_deepspeed_b200.inference.v2 import RaggedInferenceEngineConfig, build_engine_from_module  # noqa: E402


def _cfg():
    return RaggedInferenceEngineConfig(state_manager=dict(
        max_tracked_sequences=16, max_ragged_batch_size=256, max_ragged_sequence_count=8, max_context=512,
        memory_config=dict(mode="allocate", size=32 * 1024 * 1024)))


def _tiny(family):
    T = transformers
    common = dict(hidden_size=128, num_hidden_layers=2, num_attention_heads=4, vocab_size=300, max_position_embeddings=512)
    if family == "llama":
        return T.LlamaForCausalLM(T.LlamaConfig(intermediate_size=256, num_key_value_heads=2, **common))
    if family == "llama_mha_bias":
        return T.LlamaForCausalLM(T.LlamaConfig(intermediate_size=256, num_key_value_heads=4, attention_bias=True, mlp_bias=True, **common))
    if family == "mistral":
        return T.MistralForCausalLM(T.MistralConfig(intermediate_size=256, num_key_value_heads=2, sliding_window=None, **common))
    if family == "qwen2":
        return T.Qwen2ForCausalLM(T.Qwen2Config(intermediate_size=256, num_key_value_heads=2, **common))
    if family == "opt":
        return T.OPTForCausalLM(T.OPTConfig(ffn_dim=256, word_embed_proj_dim=128, **common))
    if family == "falcon_mqa":
        return T.FalconForCausalLM(T.FalconConfig(hidden_size=128, num_hidden_layers=2, num_attention_heads=4, vocab_size=300,
                                                  multi_query=True, parallel_attn=True, new_decoder_architecture=False,
                                                  alibi=False, bias=False))
    if family == "falcon_new":
        return T.FalconForCausalLM(T.FalconConfig(hidden_size=128, num_hidden_layers=2, num_attention_heads=4, num_kv_heads=2,
                                                  vocab_size=300, new_decoder_architecture=True, alibi=False, bias=False))
    if family == "phi":
        return T.PhiForCausalLM(T.PhiConfig(intermediate_size=256, partial_rotary_factor=0.5, **common))
    if family == "phi3":
        return T.Phi3ForCausalLM(T.Phi3Config(intermediate_size=256, num_key_value_heads=2, pad_token_id=0, **common))
    if family == "mixtral":
        return T.MixtralForCausalLM(T.MixtralConfig(intermediate_size=128, num_key_value_heads=2, num_local_experts=4,
                                                    num_experts_per_tok=2, sliding_window=None, **common))
    if family == "qwen2_moe":
        return T.Qwen2MoeForCausalLM(T.Qwen2MoeConfig(intermediate_size=256, moe_intermediate_size=64, shared_expert_intermediate_size=128,
                                                      num_key_value_heads=2, num_experts=4, num_experts_per_tok=2,
                                                      decoder_sparse_step=1, mlp_only_layers=[], **common))
    if family == "gpt2":
        return T.GPT2LMHeadModel(T.GPT2Config(n_embd=128, n_layer=2, n_head=4, vocab_size=300, n_positions=512))
    if family == "gptj":
        return T.GPTJForCausalLM(T.GPTJConfig(n_embd=128, n_layer=2, n_head=4, vocab_size=300, n_positions=512, rotary_dim=16))
    if family == "gpt_neox":
        return T.GPTNeoXForCausalLM(T.GPTNeoXConfig(intermediate_size=256, rotary_pct=0.5, use_parallel_residual=True, **common))
    if family == "gpt_neox_sequential":
        return T.GPTNeoXForCausalLM(T.GPTNeoXConfig(intermediate_size=256, rotary_pct=1.0, use_parallel_residual=False, **common))
    if family == "bloom":
        return T.BloomForCausalLM(T.BloomConfig(hidden_size=128, n_layer=2, n_head=4, vocab_size=300))
    if family == "bloom_6heads":      # a head count that is not a power of two: the interleaved slope sequence
        return T.BloomForCausalLM(T.BloomConfig(hidden_size=192, n_layer=2, n_head=6, vocab_size=300))
    if family == "falcon_alibi":
        return T.FalconForCausalLM(T.FalconConfig(hidden_size=128, num_hidden_layers=2, num_attention_heads=4, vocab_size=300,
                                                  multi_query=False, parallel_attn=False, new_decoder_architecture=False,
                                                  alibi=True, bias=True))
    if family == "gpt_neo":       # local layers with a window shorter than the test prompts
        return T.GPTNeoForCausalLM(T.GPTNeoConfig(hidden_size=128, num_layers=2, num_heads=4, vocab_size=300, window_size=8,
                                                  attention_types=[[["global", "local"], 1]], max_position_embeddings=512))
    if family == "mistral_window":
        return T.MistralForCausalLM(T.MistralConfig(intermediate_size=256, num_key_value_heads=2, sliding_window=10, **common))
    if family == "phi3_window":
        return T.Phi3ForCausalLM(T.Phi3Config(intermediate_size=256, num_key_value_heads=2, pad_token_id=0, sliding_window=12,
                                              **common))
    raise KeyError(family)


FAMILIES = ["llama", "llama_mha_bias", "mistral", "qwen2", "opt", "falcon_mqa", "falcon_new", "phi", "phi3", "mixtral", "qwen2_moe",
            "gpt2", "gptj", "gpt_neox", "gpt_neox_sequential",
            "bloom", "bloom_6heads", "falcon_alibi", "gpt_neo", "mistral_window", "phi3_window"]


@pytest.mark.parametrize("family", FAMILIES)
def test_family_matches_huggingface(family):
    torch.manual_seed(0)
    try:
        model = _tiny(family).eval().float()
    except Exception as e:  # config signature drift between transformers versions
        pytest.skip(f"cannot build HF {family}: {type(e).__name__}: {e}")
    eng = build_engine_from_module(model, _cfg(), dtype=torch.float32, device=torch.device("cpu"))
    g = torch.Generator().manual_seed(1)
    a = torch.randint(0, 300, (31, ), generator=g)
    b = torch.randint(0, 300, (9, ), generator=g)
    logits, latents = eng.put([1, 2], [a, b])
    with torch.no_grad():
        ra = model(a[None]).logits[0, -1]
        rb = model(b[None]).logits[0, -1]
    torch.testing.assert_close(logits[0], ra, rtol=2e-4, atol=2e-4)
    torch.testing.assert_close(logits[1], rb, rtol=2e-4, atol=2e-4)
    # one decode step on top of the cached context
    nxt = ra.argmax()[None]
    l2, lat2 = eng.put([1], [nxt])
    with torch.no_grad():
        r2 = model(torch.cat([a, nxt])[None]).logits[0, -1]
    torch.testing.assert_close(l2[0], r2, rtol=2e-4, atol=2e-4)
    # HCache: restore a fresh sequence from the captured hidden states, next logits must agree
    saved = torch.cat([latents[0], lat2[0]], dim=1)
    eng.restore_kv([7], [torch.cat([a, nxt])], [saved])
    probe = r2.argmax()[None]
    la, _ = eng.put([1], [probe])
    lb, _ = eng.put([7], [probe])
    torch.testing.assert_close(la, lb, rtol=1e-5, atol=1e-5)
