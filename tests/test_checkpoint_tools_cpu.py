"""zero_to_fp32 consolidation, universal checkpoint conversion + resume under another dp degree, checkpoint inspection
(reference: tests/unit/checkpoint/{test_zero_optimizer,test_universal_checkpoint}.py)."""
import os

import pytest
import torch

from tests.dist_utils import run_distributed
from tests.runtime_helper import ckpt_tools_worker

pytestmark = pytest.mark.timeout(600)


@pytest.mark.parametrize("stage", [1, 2, 3])
def test_zero_to_fp32_consolidates_shards(stage, tmp_path):
    from hcache_deepspeed_b200.checkpoint import (DeepSpeedCheckpoint, convert_zero_checkpoint_to_fp32_state_dict,
                                                  get_fp32_state_dict_from_zero_checkpoint)
    d = str(tmp_path / "ck")
    states = run_distributed(ckpt_tools_worker, 2, stage, d, "save")
    sd = get_fp32_state_dict_from_zero_checkpoint(d)          # uses the 'latest' file
    for k, v in states[0].items():
        torch.testing.assert_close(sd[k], v, rtol=1e-6, atol=1e-7)
    out = convert_zero_checkpoint_to_fp32_state_dict(d, str(tmp_path / "fp32"), tag="t3")
    assert os.path.exists(out) and set(torch.load(out, weights_only=False)) >= set(states[0])
    ck = DeepSpeedCheckpoint(os.path.join(d, "t3"))
    assert ck.original_dp_degree == 2 and ck.original_tp_degree == 1 and len(ck.zero_files) == 2
    # lazy mode hands out descriptions that materialise one by one; a small shard limit gives numbered files + an index
    from hcache_deepspeed_b200.checkpoint.zero_to_fp32 import (GatheredTensor, get_model_state_files, get_optim_files,
                                                                natural_keys, parse_model_states, parse_optim_states,
                                                                to_torch_tensor, zero3_partitioned_param_info)
    lazy = get_fp32_state_dict_from_zero_checkpoint(d, tag="t3", lazy_mode=True)
    assert any(isinstance(v, GatheredTensor) for v in lazy.values())
    full = to_torch_tensor(lazy)
    for k, v in states[0].items():
        torch.testing.assert_close(full[k], v, rtol=1e-6, atol=1e-7)
    assert all(v.numel() == 0 or v.shape == full[k].shape for k, v in to_torch_tensor(lazy, return_empty_tensor=True).items())
    import json
    for safe in (False, True):
        idx = convert_zero_checkpoint_to_fp32_state_dict(d, str(tmp_path / f"sharded{int(safe)}"), max_shard_size="3KB", tag="t3",
                                                         safe_serialization=safe)
        meta = json.load(open(idx))
        files = sorted(set(meta["weight_map"].values()))
        assert idx.endswith(".index.json") and len(files) > 1 and set(meta["weight_map"]) >= set(states[0])
        merged = {}
        for f in files:
            path = os.path.join(os.path.dirname(idx), f)
            if safe:
                from safetensors.torch import load_file
                merged.update(load_file(path))
            else:
                merged.update(torch.load(path, weights_only=False))
        for k, v in states[0].items():
            torch.testing.assert_close(merged[k], v, rtol=1e-6, atol=1e-7)
    tag_dir = os.path.join(d, "t3")
    got_stage, world, flats = parse_optim_states(get_optim_files(tag_dir), tag_dir)
    assert (got_stage, world, len(flats)) == (stage, 2, 2)
    ms = parse_model_states(get_model_state_files(tag_dir))
    assert ms[0].param_shapes and isinstance(ms[0].shared_params, dict)
    assert zero3_partitioned_param_info(10, 4) == (3, 2) and sorted(["r10", "r2"], key=natural_keys) == ["r2", "r10"]


def test_universal_checkpoint_resumes_with_other_dp_degree(tmp_path):
    from hcache_deepspeed_b200.checkpoint import convert_to_universal
    d = str(tmp_path / "ck")
    states = run_distributed(ckpt_tools_worker, 2, 2, d, "save")
    uni = convert_to_universal(d, str(tmp_path / "ck" / "t3_universal"), tag="t3")
    assert os.path.isdir(os.path.join(uni, "zero", "head.weight")) and os.path.exists(os.path.join(uni, "zero", "head.weight", "exp_avg.pt"))
    res = run_distributed(ckpt_tools_worker, 1, 1, uni, "universal")   # dp 2 / stage 2  ->  dp 1 / stage 1
    assert res[0]["restored"] == len(states[0]) and res[0]["exp_avg_norm"] > 0
    for k, v in states[0].items():
        torch.testing.assert_close(res[0]["state"][k], v, rtol=1e-6, atol=1e-7)
    # the config switch: "checkpoint": {"load_universal": true} + engine.load_checkpoint(dir) under dp 3 / stage 1
    res = run_distributed(ckpt_tools_worker, 3, 1, d, "universal_config")
    for r in res:
        assert r["path"].endswith("t3_universal") and r["steps"] == 3 and r["exp_avg_norm"] > 0
        for k, v in states[0].items():
            torch.testing.assert_close(r["state"][k], v, rtol=1e-6, atol=1e-7)


def test_reshape_3d_maps_and_zero_checkpoint(tmp_path):
    """(pp, tp, dp) contraction maps and ZeRO optimizer-file merging (reference: checkpoint/reshape_*.py, zero_checkpoint.py)."""
    import torch
    from hcache_deepspeed_b200.checkpoint import reshape as R
    m = R.reshape_meg_2d_parallel(4, 4, 2, 2)
    assert m.get_data(0, 0) == [0, 1, 4, 5] and m.get_data(1, 1) == [10, 11, 14, 15]
    maps = R.model_3d_desc(2, 2, 2).reshape(R.model_3d_desc(1, 1, 1))
    assert len(maps) == 1 and maps[0].get_data(0, 0) == [0, 4, 1, 5, 2, 6, 3, 7]
    maps = R.model_3d_desc(1, 2, 4).reshape(R.model_3d_desc(1, 2, 2))
    assert maps[0].get_data(0, 1) == [1, 3] and maps[1].get_data(0, 1) == [5, 7]
    ok, errs = R.model_3d_desc(1, 2, 2).can_reshape(R.model_3d_desc(1, 4, 2))
    assert not ok and "TP" in errs[0]
    tp, pp, dp = R.get_mpu_ranks(2, 4, 2)
    assert tp[0] == [0, 1] and pp[0] == [0, 4, 8, 12] and dp[:2] == [[0, 2], [1, 3]]

    # a tp=1, pp=1, dp=4 ZeRO checkpoint: flat fp32 partitions of 6 elements, the last rank carries 2 of padding
    for r in range(4):
        sd = {"optimizer_state_dict": {
            "base_optimizer_state": {"state": {0: {"exp_avg": torch.full((6, ), float(r)), "step": 3}}},
            "group_paddings": [2 if r == 3 else 0], "partition_count": [4],
            "single_partition_of_fp32_groups": [torch.arange(6.0) + 10 * r]}}
        torch.save(sd, R.get_zero_ckpt_name_for_rank(str(tmp_path), r, 0))
    torch.save({}, R.get_model_ckpt_name_for_rank(str(tmp_path), "00"))
    ck = R.ZeROCheckpoint(str(tmp_path))
    assert (ck.get_src_pp_degree(), ck.get_src_tp_degree(), ck.get_src_dp_degree()) == (1, 1, 4)
    ck.reshape(R.model_3d_desc(1, 1, 2))
    assert ck.get_file_indices_for_rank(0, 0, 1) == [2, 3]
    sd = ck.get_state_for_rank(0, 0, 1)["optimizer_state_dict"]
    assert sd["partition_count"] == [2] and sd["group_paddings"] == [0]
    ea = sd["base_optimizer_state"]["state"][0]["exp_avg"]
    assert ea.tolist() == [2.0] * 6 + [3.0] * 4            # rank 3's padding stripped before the merge
    assert sd["single_partition_of_fp32_groups"][0].numel() == 12
    c = R.clone_tensors_for_torch_save({"a": [torch.ones(4)[:2]]})
    assert c["a"][0].untyped_storage().size() == 8


def test_deepspeed_checkpoint_2d_views_of_a_pipeline_checkpoint(tmp_path):
    """Megatron-DeepSpeed layout (one file per layer and tensor-parallel slice + one model-states file per (pipe, tensor)
    rank): the layer / stage maps and the merged states when opened under a smaller tp x pp (ref:
    checkpoint/deepspeed_checkpoint.py:96-260)."""
    import torch
    from hcache_deepspeed_b200.checkpoint import DeepSpeedCheckpoint
    tp, pp, blocks, h = 2, 2, 4, 4
    g = torch.Generator().manual_seed(0)
    full = {}
    for layer in range(blocks + 2):
        if layer == 0:
            sd = {"word_embeddings.weight": torch.randn(8, h, generator=g), "position_embeddings.weight": torch.randn(6, h, generator=g)}
        elif layer == blocks + 1:
            sd = {"weight": torch.randn(h, generator=g), "bias": torch.randn(h, generator=g)}
        else:
            sd = {"input_layernorm.weight": torch.randn(h, generator=g),
                  "self_attention.query_key_value.weight": torch.randn(3 * h, h, generator=g),
                  "self_attention.dense.weight": torch.randn(h, h, generator=g),
                  "self_attention.dense.bias": torch.randn(h, generator=g),
                  "mlp.dense_h_to_4h.weight": torch.randn(4 * h, h, generator=g),
                  "mlp.dense_4h_to_h.weight": torch.randn(h, 4 * h, generator=g)}
        full[layer] = sd
        for r in range(tp):
            part = {}
            for k, v in sd.items():
                if k in ("input_layernorm.weight", "self_attention.dense.bias", "position_embeddings.weight") or layer == blocks + 1:
                    part[k] = v.clone()
                elif k in ("self_attention.dense.weight", "mlp.dense_4h_to_h.weight"):
                    part[k] = v.chunk(tp, dim=1)[r].clone()
                else:
                    part[k] = v.chunk(tp, dim=0)[r].clone()
            torch.save(part, tmp_path / f"layer_{layer:02d}-model_{r:02d}-model_states.pt")
    for r in range(tp * pp):
        torch.save({"iteration": 17, "args": {"hidden": h}, "checkpoint_info": {"v": 1}, "module": None}, tmp_path / f"mp_rank_{r:02d}_model_states.pt")

    ck = DeepSpeedCheckpoint(str(tmp_path))
    assert (ck.original_tp_degree, ck.original_pp_degree) == (2, 2) and ck.get_iteration() == 17 and ck.get_checkpoint_info() == {"v": 1}
    assert ck.get_embedding_layer_id() == "layer_00" and ck.get_final_norm_layer_id() == "layer_05"
    assert ck.get_pp_transformer_map(0) == ["layer_01", "layer_02"] and ck.get_pp_transformer_map(1) == ["layer_03", "layer_04"]
    assert [len(f) for f in ck.transformer_file_map[(1, 0)]] == [1, 1]
    st = ck.get_transformer_state(tp_index=1, pp_index=1)      # unchanged degrees: the slice files as they are
    torch.testing.assert_close(st[0]["mlp.dense_4h_to_h.weight"], full[3]["mlp.dense_4h_to_h.weight"].chunk(2, dim=1)[1])

    one = DeepSpeedCheckpoint(str(tmp_path), tp_degree=1, pp_degree=1)
    assert one.is_change_tp_degree() and one.is_change_pp_degree() and not one.is_change_dp_degree()
    assert len(one.get_embedding_files(0)) == 2 and len(one.get_final_norm_files(0)) == 2
    emb = one.get_embedding_state(0)
    torch.testing.assert_close(emb["word_embeddings.weight"], full[0]["word_embeddings.weight"])
    torch.testing.assert_close(emb["position_embeddings.weight"], full[0]["position_embeddings.weight"])
    torch.testing.assert_close(one.get_final_norm_state(0)["weight"], full[blocks + 1]["weight"])
    layers = one.get_transformer_state(0, 0)
    assert len(layers) == blocks and one.get_pp_transformer_map(0) == [f"layer_{i:02d}" for i in range(1, blocks + 1)]
    for i, sd in enumerate(layers, start=1):
        for k, v in full[i].items():
            torch.testing.assert_close(sd[k], v)
    assert [os.path.basename(f) for f in one.get_2d_parallel_files(0, 0)] == [f"mp_rank_{r:02d}_model_states.pt" for r in range(4)]
    assert one.get_2d_parallel_state(0, 0)["iteration"] == 17
    one.show_2d_mapping(); one.show_tp_embedding_map(); one.show_tp_final_norm_map(); one.show_pp_transformer_map(); one.show_transformer_file_map()


def test_universal_checkpoint_from_and_into_zero3(tmp_path):
    """ZeRO-3 (dp 2) -> universal -> ZeRO-3 under dp 3 and ZeRO-1 under dp 1: weights and Adam moments survive
    (ref: ds_to_universal.py extract_zero_shards_stage3 / merge_zero3_slices, stage3.py load_hp_checkpoint_state)."""
    from hcache_deepspeed_b200.checkpoint import convert_to_universal
    d = str(tmp_path / "ck")
    states = run_distributed(ckpt_tools_worker, 2, 3, d, "save")
    uni = convert_to_universal(d, str(tmp_path / "ck" / "t3_universal"), tag="t3")
    norms = []
    for world, stage in ((3, 3), (1, 1)):
        res = run_distributed(ckpt_tools_worker, world, stage, uni, "universal")
        for r in res:
            assert r["restored"] == len(states[0]) and r["exp_avg_norm"] > 0
            for k, v in states[0].items():
                torch.testing.assert_close(r["state"][k], v, rtol=1e-6, atol=1e-7)
        norms.append(res[0]["exp_avg_norm"])
    assert norms[0] == pytest.approx(norms[1], rel=1e-5)       # the same moments, whatever layout they were cut into


def test_universal_tp_merge_rules_and_reslicing(tmp_path):
    """Joining tensor-parallel slices by the checkpoint's pattern lists, and cutting them again for another tp degree
    (ref: ds_to_universal.py merge_tp_slices :232, universal_checkpoint.py load_hp_checkpoint_state)."""
    from hcache_deepspeed_b200.checkpoint.ds_to_universal import dump_param_fragment, merge_tp_slices
    from hcache_deepspeed_b200.checkpoint.universal_checkpoint import SubparamShape, load_hp_checkpoint_state
    info = {"tp_replicated_parameter_patterns": [r".*norm.weight"], "parameter_to_average_patterns": [r".*router.bias"],
            "parameter_with_row_parallelism_patterns": [r".*dense.weight"], "vocabulary_parameter_patterns": [r".*embed.weight"],
            "parameter_with_2_sub_params_cat_dim_0": [r".*gate_up.weight"], "original_vocab_size": 10,
            "parameter_with_sub_params": [dict(patterns=[r".*qkv.weight"], shape=((8, 4, 4), 6), partition_dim=0)]}
    g = torch.Generator().manual_seed(0)
    full = {"l.norm.weight": torch.randn(6, generator=g), "l.router.bias": torch.randn(4, generator=g),
            "l.dense.weight": torch.randn(6, 8, generator=g), "l.col.weight": torch.randn(8, 6, generator=g),
            "l.gate_up.weight": torch.randn(16, 6, generator=g), "l.qkv.weight": torch.randn(16, 6, generator=g),
            "l.embed.weight": torch.randn(10, 6, generator=g)}

    def cut(name, t, r, tp=2):
        if "norm" in name:
            return t.clone()
        if "router" in name:
            return t + (0.5 if r == 0 else -0.5)
        if "dense" in name:
            return t.chunk(tp, dim=1)[r]
        if "gate_up" in name:
            return torch.cat([h.chunk(tp, dim=0)[r] for h in t.chunk(2, dim=0)], dim=0)
        if "qkv" in name:
            return torch.cat([p.chunk(tp, dim=0)[r] for p in t.split([8, 4, 4], dim=0)], dim=0)
        if "embed" in name:
            return torch.cat([t, torch.zeros(2, 6)], dim=0).chunk(tp, dim=0)[r]      # vocabulary padded to 12 rows
        return t.chunk(tp, dim=0)[r]

    for name, t in full.items():
        merged, meta = merge_tp_slices(name, [cut(name, t, r) for r in range(2)], info)
        torch.testing.assert_close(merged, t)
        dump_param_fragment(str(tmp_path / "zero"), name, {"fp32": merged}, meta)
        for tp in (1, 2):
            for r in range(tp):
                got = load_hp_checkpoint_state(str(tmp_path), name, "fp32", tp_rank=r, tp_world_size=tp, padded_vocab_size=12)
                want = cut(name, t, r, tp) if tp > 1 else (torch.cat([t, torch.zeros(2, 6)]) if "embed" in name else t)
                if "router" in name:
                    want = t
                torch.testing.assert_close(got, want)
    assert isinstance(SubparamShape(**info["parameter_with_sub_params"][0]).shape, tuple)
