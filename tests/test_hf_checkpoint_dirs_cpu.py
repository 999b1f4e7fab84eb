"""``build_hf_engine`` on real checkpoint directories (the user-facing entry point, ref engine_factory.py:68-135 +
checkpoint/huggingface_engine.py): single safetensors file, sharded safetensors with an index, a ``pytorch_model.bin``, and
a directory written by ``InferenceEngineV2.serialize``."""
import json
import os

import pytest
import torch

transformers = pytest.importorskip("transformers")

from hcache_deepspeed_b200.inference.v2 import build_hf_engine  # noqa: E402

CFG = {"state_manager": dict(max_tracked_sequences=8, max_ragged_batch_size=128, max_ragged_sequence_count=4, max_context=256,
                             memory_config=dict(mode="allocate", size=16 << 20))}


def _models():
    T = transformers
    common = dict(hidden_size=64, num_hidden_layers=2, num_attention_heads=4, vocab_size=120, max_position_embeddings=128)
    return {"llama": lambda: T.LlamaForCausalLM(T.LlamaConfig(intermediate_size=128, num_key_value_heads=2, **common)),
            "bloom": lambda: T.BloomForCausalLM(T.BloomConfig(hidden_size=64, n_layer=2, n_head=4, vocab_size=120)),
            "gpt2": lambda: T.GPT2LMHeadModel(T.GPT2Config(n_embd=64, n_layer=2, n_head=4, vocab_size=120, n_positions=128))}


@pytest.mark.parametrize("family", ["llama", "bloom", "gpt2"])
@pytest.mark.parametrize("layout", ["safetensors", "sharded_index", "bin"])
def test_build_hf_engine_from_directory(family, layout, tmp_path):
    torch.manual_seed(0)
    model = _models()[family]().eval().float()
    ids = torch.randint(1, 120, (13, ))
    with torch.no_grad():
        want = model(ids[None]).logits[0, -1]
    d = str(tmp_path / "ckpt")
    if layout == "safetensors":
        model.save_pretrained(d)
        assert os.path.exists(os.path.join(d, "model.safetensors"))
    elif layout == "sharded_index":
        model.save_pretrained(d, max_shard_size="40KB")
        assert os.path.exists(os.path.join(d, "model.safetensors.index.json"))
    else:
        model.save_pretrained(d)
        os.remove(os.path.join(d, "model.safetensors"))
        sd = {k: v for k, v in model.state_dict().items()}
        torch.save(sd, os.path.join(d, "pytorch_model.bin"))
    eng = build_hf_engine(d, CFG, dtype=torch.float32)
    got, _ = eng.put([1], [ids])
    torch.testing.assert_close(got[0], want, rtol=2e-4, atol=2e-4)
    if layout == "safetensors":
        # serialize -> a directory build_hf_engine recognises by itself
        out = str(tmp_path / "ds")
        eng.serialize(out)
        eng2 = build_hf_engine(out, CFG)
        got2, _ = eng2.put([1], [ids])
        torch.testing.assert_close(got2[0], got[0], rtol=0, atol=0)


def test_build_hf_engine_rejects_unknown_model_type(tmp_path):
    d = tmp_path / "x"
    d.mkdir()
    (d / "config.json").write_text(json.dumps({"model_type": "t5", "architectures": ["T5ForConditionalGeneration"]}))
    with pytest.raises(FileNotFoundError, match="no \\*.safetensors or \\*.bin weights"):
        build_hf_engine(str(d), CFG)
    torch.save({"w": torch.zeros(1)}, d / "pytorch_model.bin")
    with pytest.raises(ValueError, match="Unsupported model type t5"):
        build_hf_engine(str(d), CFG)
