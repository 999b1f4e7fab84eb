"""AutoTP, TP layers and shard arithmetic (reference: tests/unit/inference/test_inference.py AutoTP cases,
tests/unit/model_parallelism/test_autotp_training.py)."""
import pytest
import torch

from tests.autotp_helper import autotp_worker
from tests.dist_utils import run_distributed

pytestmark = pytest.mark.timeout(600)


def test_shard_sizes_follow_heads():
    from hcache_deepspeed_b200.module_inject import tp_shard
    tp_shard.set_num_kv_heads(6)
    assert tp_shard.get_shard_size_list(6 * 64, 4) == [128, 128, 64, 64]
    assert sum(tp_shard.get_shard_size_list(6 * 64, 4)) == 384
    assert tp_shard.get_shard_size_list(1000, 3, "lm_head") == [334, 333, 333]
    tp_shard.set_num_kv_heads(None)
    assert tp_shard.get_shard_size_list(10, 4) == [3, 3, 2, 2]


def test_fused_qkv_split():
    from hcache_deepspeed_b200.module_inject.fusedqkv_utils import prepare_tp_fused_qkvw, shard_chunk_mlp
    from hcache_deepspeed_b200.module_inject import tp_shard
    tp_shard.set_num_kv_heads(None)
    h = 8
    w = torch.arange(3 * h * 2, dtype=torch.float32).reshape(3 * h, 2)
    parts = [prepare_tp_fused_qkvw("GLMBlock", w, 2, r) for r in range(2)]
    q, k, v = torch.chunk(w, 3, dim=0)
    assert torch.equal(parts[0], torch.cat([q[:4], k[:4], v[:4]])) and torch.equal(parts[1], torch.cat([q[4:], k[4:], v[4:]]))
    wb = torch.arange(4 * 3 * 2 * 2, dtype=torch.float32).reshape(4 * 3 * 2, 2)   # 4 heads x (q,k,v) x d=2
    p0 = prepare_tp_fused_qkvw("BloomBlock", wb, 2, 0, num_heads=4)
    assert torch.equal(p0, wb[:12])
    gu = torch.arange(8 * 2, dtype=torch.float32).reshape(8, 2)
    w1, _ = shard_chunk_mlp(gu, None, 1, 2)
    assert torch.equal(w1, torch.cat([gu[2:4], gu[6:8]]))


def test_group_quantizer_roundtrip():
    from hcache_deepspeed_b200.module_inject.replace_module import GroupQuantizer
    torch.manual_seed(0)
    w = torch.randn(64, 32)
    q = GroupQuantizer(q_int8=True, group_size=8).quantize(w)
    assert q.dtype == torch.int8 and q.scale.numel() == 8
    back = GroupQuantizer.dequantize(q.data, q.scale, torch.float32)
    assert float((back - w).abs().max()) < float(w.abs().max()) / 100


def test_policy_table_lists_reference_containers():
    from hcache_deepspeed_b200.module_inject.policy import policy_for, replace_policies
    names = {p.name for p in replace_policies()}
    for n in ("bert", "bloom", "clip", "distil_bert", "gpt2", "gptj", "gptneo", "gptneox", "internlm", "llama", "llama2",
              "megatron_gpt", "megatron_gpt_moe", "opt", "unet", "vae"):
        assert n in names
    assert policy_for("llama").supported and policy_for("opt").supported and not policy_for("bloom").supported


@pytest.mark.parametrize("world", [2, 4])
def test_autotp_matches_unsharded(world):
    for r in run_distributed(autotp_worker, world):
        assert r["err"] < 1e-4
        assert r["gems"] == ["mlp.down_proj", "self_attn.o_proj"]
        assert r["types"] == ["LinearLayer", "LinearAllreduce", "LinearAllreduce", "LmHeadLinearAllreduce"]
        assert r["heads"] == 4 // world and r["q_rows"] == 32 // world and r["o_cols"] == 32 // world


def test_autotp_on_huggingface_architectures():
    """AutoTP (no kernel injection) over two gloo ranks on fourteen real HuggingFace architectures: Conv1D projections
    (GPT-2), every fused-QKV layout (per-head interleaved: BLOOM / GPT-NeoX / Falcon; [q | k | v]: GPT-2; multi-query and
    grouped Falcon; Phi-3's GQA fused QKV + fused gate/up), ALiBi slopes per rank (BLOOM), attention modules that read head
    counts from their config (Phi-3), MoE experts (Mixtral). Sharded logits == unsharded logits."""
    pytest.importorskip("transformers")
    from tests.autotp_helper import HF_AUTOTP_CASES, hf_autotp_worker
    res = run_distributed(hf_autotp_worker, 2, list(HF_AUTOTP_CASES), timeout=600)
    for r in res:
        bad = {k: v for k, v in r.items() if not (isinstance(v, float) and v < 1e-4) and not str(v).startswith("skip")}
        assert not bad, bad
        assert sum(isinstance(v, float) for v in r.values()) >= 10


def test_autotp_four_ranks_mha_mqa_and_gqa_limit():
    pytest.importorskip("transformers")
    from tests.autotp_helper import hf_autotp_worker
    res = run_distributed(hf_autotp_worker, 4, ["bloom", "gpt2", "falcon_mqa", "llama_gqa"], timeout=600)
    for r in res:
        assert r["bloom"] < 1e-4 and r["gpt2"] < 1e-4 and r["falcon_mqa"] < 1e-4
        assert "cannot be split over 4 ranks" in str(r["llama_gqa"])      # 2 KV heads: a clear error on every rank, no hang


def test_v1_kernel_inject_under_tensor_parallelism_hf_models():
    pytest.importorskip("transformers")
    from tests.autotp_helper import v1_inject_tp_worker
    kinds = ["llama_gqa", "bloom", "gpt2", "falcon_mqa", "mixtral", "gpt_neo", "phi3"]
    for r in run_distributed(v1_inject_tp_worker, 2, kinds, timeout=600):
        for k in kinds:
            assert r[k]["injected"] and r[k]["err"] < 1e-4 and r[k]["gen_equal"], (k, r[k])


def test_init_inference_injection_policy_names_row_parallel_linears():
    from tests.autotp_helper import injection_policy_worker
    for r in run_distributed(injection_policy_worker, 2):
        assert r["types"] == ["LinearLayer", "LinearAllreduce"] and r["err"] < 1e-5, r
        assert "NoSuchBlock" in r["bad"]


@pytest.mark.parametrize("world,tp", [(2, 2), (4, 2)])
def test_training_tensor_parallelism_matches_unsharded_model(world, tp):
    pytest.importorskip("transformers")
    from tests.autotp_helper import tp_training_worker
    for r in run_distributed(tp_training_worker, world, tp, timeout=600):
        assert r["dp"] == world // tp and r["q_rows"] == 64 // tp
        assert r["got"] == pytest.approx(r["want"], rel=2e-4, abs=2e-4), r


def test_self_partitioning_tp_layers_2ranks():
    from tests.autotp_helper import tp_layers_worker
    from tests.dist_utils import run_distributed
    for r in run_distributed(tp_layers_worker, 2):
        assert r["col_err"] < 1e-5 and r["row_err"] < 1e-5 and r["conv_err"] < 1e-5 and r["convrow_err"] < 1e-5, r
        assert r["col_shape"] == (3, 6) and r["row_shape"] == (4, 3) and r["frozen"], r
        assert r["gathered"] and r["train_grad_err"] < 1e-5 and r["trainable"], r
        assert r["conv2d_err"] < 1e-5 and r["gateup_err"] < 1e-5 and r["mode_roundtrip"], r
