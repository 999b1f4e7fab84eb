"""Tensor parallelism end to end on CPU ranks (gloo): TP=2 / TP=3 logits == TP=1 logits, incl. GQA with replicated
KV heads, uneven head splits, OPT biases, and HCache restore under TP."""
import pytest
import torch

from tests.dist_utils import run_distributed


def _run_rank(rank, world, arch_kwargs, seed):
    import torch
    import torch.distributed as dist
    from hcache_deepspeed_b200.inference.v2 import RaggedInferenceEngineConfig
    from hcache_deepspeed_b200.inference.v2.engine_v2 import InferenceEngineV2
    from hcache_deepspeed_b200.inference.v2.model_implementations import InferenceV2Policy
    from hcache_deepspeed_b200.models.arch import ArchSpec
    from hcache_deepspeed_b200.models.weights import load_weights, random_weights
    from hcache_deepspeed_b200.models.transformer import RaggedTransformer, TPComm

    torch.manual_seed(0)
    spec = ArchSpec(**arch_kwargs)
    cfg = RaggedInferenceEngineConfig(tensor_parallel={"tp_size": world}, state_manager=dict(
        max_tracked_sequences=8, max_ragged_batch_size=128, max_ragged_sequence_count=4, max_context=256,
        memory_config=dict(mode="allocate", size=16 * 1024 * 1024)))

    # every rank derives its shard from the SAME full (tp=1) random model
    full = random_weights(spec, 0, 1, torch.float32, torch.device("cpu"), seed=seed)

    class FromFull(InferenceV2Policy):
        def build_model(self, engine_config, mp_group):
            from hcache_deepspeed_b200.models.weights import shard_info, interleave_gate_up
            from tests.tp_shard_helper import shard_full_weights
            w = shard_full_weights(spec, full, rank, world)
            comm = TPComm(group=mp_group, size=world, rank=rank)
            return RaggedTransformer(spec, w, engine_config.state_manager, dtype=torch.float32, device=torch.device("cpu"),
                                     tp_rank=rank, tp_size=world, comm=comm)

    eng = InferenceEngineV2(FromFull(arch=spec, random_seed=seed), cfg)
    g = torch.Generator().manual_seed(3)
    a, b = torch.randint(0, spec.vocab, (23, ), generator=g), torch.randint(0, spec.vocab, (5, ), generator=g)
    logits, lat = eng.put([1, 2], [a, b])
    nxt = logits.argmax(-1)
    l2, lat2 = eng.put([1, 2], [nxt[0:1], nxt[1:2]])
    # restore under TP: rank 0 holds the latents (they are replicated), broadcast to the others like a scheduler would
    if world > 1:
        objs = [[lat[0], lat2[0]]] if rank == 0 else [None]
        dist.broadcast_object_list(objs, src=0)
        lat_a = torch.cat(objs[0], dim=1)
    else:
        lat_a = torch.cat([lat[0], lat2[0]], dim=1)
    eng.restore_kv([7], [torch.cat([a, nxt[0:1]])], [lat_a])
    probe = l2.argmax(-1)[0:1]
    la, _ = eng.put([1], [probe])
    lb, _ = eng.put([7], [probe])
    return logits, l2, la, lb


BASE = dict(model_type="llama", n_layers=2, hidden=64, head_dim=16, intermediate=96, vocab=101, norm="rms", act="silu",
            gated_mlp=True, pos_emb="rope", max_positions=256)
CASES = {
    "mha_tp2": (dict(BASE, n_heads=4, n_kv_heads=4), 2),
    "gqa_split_groups_tp2": (dict(BASE, n_heads=4, n_kv_heads=2), 2),
    "gqa_replicated_kv_tp2": (dict(BASE, n_heads=4, n_kv_heads=1), 2),
    "mha_uneven_tp3": (dict(BASE, n_heads=4, n_kv_heads=4, hidden=64), 3),
    "opt_bias_tp2": (dict(model_type="opt", n_layers=2, hidden=64, n_heads=4, n_kv_heads=4, head_dim=16, intermediate=128,
                          vocab=101, norm="layer", act="relu", gated_mlp=False, pos_emb="learned", learned_pos_offset=2,
                          max_positions=256, qkv_bias=True, attn_out_bias=True, mlp_bias=True, norm_bias=True,
                          tie_word_embeddings=True), 2),
    # ALiBi: each rank adds the slopes of ITS heads; 6 heads = the interleaved (non power of two) slope sequence
    "bloom_alibi_tp3": (dict(model_type="bloom", n_layers=2, hidden=96, n_heads=6, n_kv_heads=6, head_dim=16, intermediate=384,
                             vocab=101, norm="layer", act="gelu_tanh", gated_mlp=False, pos_emb="alibi", embed_norm=True,
                             max_positions=256, qkv_bias=True, attn_out_bias=True, mlp_bias=True, norm_bias=True,
                             tie_word_embeddings=True), 3),
}


@pytest.mark.parametrize("case", sorted(CASES))
def test_tp_matches_single_rank(case):
    arch, world = CASES[case]
    single = run_distributed(_run_rank, 1, arch, 11)[0]
    multi = run_distributed(_run_rank, world, arch, 11)
    for r in range(world):
        for want, got in zip(single, multi[r]):
            torch.testing.assert_close(got, want, rtol=2e-4, atol=2e-4)
    # restore == live on every rank
    for r in range(world):
        torch.testing.assert_close(multi[r][2], multi[r][3], rtol=1e-5, atol=1e-5)
