"""Tensor parallelism on real GPUs (needs >= 2): the NCCL baseline and the flag-protocol collectives (GEMM tail signal,
all-reduce / reduce-scatter + all-gather inside the norm kernel over peer or NVLS multicast addresses, all-gather -> GEMM
waits, logits gather) must both reproduce the single-GPU logits; HCache restore (replicated and PCIe-sharded) under TP.
The same script runs for emulated ranks on one GPU in tests/test_gpu_tp_loopback.py."""
import pytest
import torch

from tests.dist_utils import run_distributed

pytestmark = [pytest.mark.gpu, pytest.mark.timeout(900)]

ARCH = dict(model_type="llama", n_layers=3, hidden=1024, n_heads=8, n_kv_heads=8, head_dim=128, intermediate=2816, vocab=2048,
            norm="rms", act="silu", gated_mlp=True, pos_emb="rope", max_positions=2048)


def _rank_fn(rank, world, arch_kwargs, fused, seed, nvls=True):
    import os
    os.environ["HDS_TP_NVLS"] = "1" if nvls else "0"
    os.environ["HDS_TP_STRICT"] = "1"     # a missing symmetric-memory set-up must fail the test, not fall back to NCCL
    import torch
    import torch.distributed as dist
    from hcache_deepspeed_b200.inference.v2 import RaggedInferenceEngineConfig
    from hcache_deepspeed_b200.inference.v2.engine_v2 import InferenceEngineV2
    from hcache_deepspeed_b200.inference.v2.model_implementations import InferenceV2Policy
    from hcache_deepspeed_b200.models.arch import ArchSpec
    from hcache_deepspeed_b200.models.transformer import RaggedTransformer, TPComm
    from hcache_deepspeed_b200.models.weights import random_weights
    from tests.tp_shard_helper import shard_full_weights

    torch.cuda.set_device(rank)
    dev = torch.device("cuda", rank)
    spec = ArchSpec(**arch_kwargs)
    cfg = RaggedInferenceEngineConfig(tensor_parallel={"tp_size": world, "fused_collectives": fused}, state_manager=dict(
        max_tracked_sequences=8, max_ragged_batch_size=768, max_ragged_sequence_count=4, max_context=2048,
        memory_config=dict(mode="allocate", size=256 * 1024 * 1024)))
    full = random_weights(spec, 0, 1, torch.bfloat16, torch.device("cpu"), seed=seed)

    class FromFull(InferenceV2Policy):
        def build_model(self, engine_config, mp_group):
            w = shard_full_weights(spec, full, rank, world)
            for lw in w.layers:
                for k, v in list(lw.__dict__.items()):
                    if isinstance(v, torch.Tensor):
                        setattr(lw, k, v.to(dev).contiguous())
            for k in ("embed", "pos_embed", "final_norm_g", "final_norm_b", "lm_head", "lm_head_b"):
                if getattr(w, k) is not None:
                    setattr(w, k, getattr(w, k).to(dev).contiguous())
            comm = TPComm(group=mp_group, size=world, rank=rank)
            if world > 1 and fused:
                from hcache_deepspeed_b200.parallel.tp_comm import make_tp_comm
                comm = make_tp_comm(mp_group, world, rank, dev)
            return RaggedTransformer(spec, w, engine_config.state_manager, dtype=torch.bfloat16, device=dev, tp_rank=rank,
                                     tp_size=world, comm=comm)

    eng = InferenceEngineV2(FromFull(arch=spec, random_seed=seed), cfg)
    used_fused = bool(getattr(eng.model().comm, "fused", False))
    g = torch.Generator().manual_seed(3)
    a, b = torch.randint(0, spec.vocab, (300, ), generator=g), torch.randint(0, spec.vocab, (40, ), generator=g)
    logits, lat = eng.put([1, 2], [a, b])
    outs = [logits.float().cpu()]
    # decode inputs are fixed tokens, not each run's own arg-max: with random weights the top logits are nearly tied
    # and a rounding-level difference would otherwise send the runs down different token paths
    for step in range(5):
        cur = torch.randint(0, spec.vocab, (2, ), generator=g)                     # the last steps replay the decode CUDA graph (flag waits inside the graph)
        lg, _ = eng.put([1, 2], [cur[0:1], cur[1:2]], capture_latents=False)
        outs.append(lg.float().cpu())
    mid = torch.randint(0, spec.vocab, (40, ), generator=g)   # fp32 partial sums + reduce-scatter / all-gather
    lg, _ = eng.put([1], [mid], capture_latents=False)
    outs.append(lg.float().cpu())
    if world > 1:
        objs = [lat[0]] if rank == 0 else [None]
        dist.broadcast_object_list(objs, src=0)
        lat_a = objs[0]
    else:
        lat_a = lat[0]
    eng.restore_kv([7], [a], [lat_a.pin_memory()])
    probe = torch.randint(0, spec.vocab, (9, ), generator=g)
    lb, _ = eng.put([7], [probe], capture_latents=False)
    outs.append(lb.float().cpu())
    if used_fused:
        # sharded restore: every rank passes only its row slice; the projection kernels gather over NVLink
        lo, hi = eng.latent_row_range(int(a.numel()))
        mine = lat_a[:, lo:hi].contiguous().pin_memory()
        for defer in (False, True):
            uid = 20 + int(defer)
            eng.restore_kv([uid], [a], [mine], sharded=True, defer=defer)
            lc, _ = eng.put([uid], [probe], capture_latents=False)
            # fp32 reductions land in arrival order: equal up to rounding, not bit for bit
            torch.testing.assert_close(lc.float().cpu(), outs[-1], rtol=3e-2, atol=3e-2)
    torch.cuda.synchronize()
    nvls_used = bool(getattr(eng.model().comm, "use_nvls", False))
    if world > 1:
        dist.barrier()
    return outs, used_fused, nvls_used


@pytest.mark.parametrize("world,fused,nvls", [(2, False, True), (2, True, True), (2, True, False), (4, True, True), (8, True, True),
                                              (8, True, False)])
def test_tp_matches_single_gpu(world, fused, nvls):
    if torch.cuda.device_count() < world:
        pytest.skip(f"needs {world} GPUs")
    single = run_distributed(_rank_fn, 1, ARCH, False, 5)[0][0]
    multi = run_distributed(_rank_fn, world, ARCH, fused, 5, nvls, timeout=600.0)
    for r in range(world):
        outs, used_fused, nvls_used = multi[r]
        if fused:
            assert used_fused, "fused collectives silently fell back to NCCL"
        for want, got in zip(single, outs):
            torch.testing.assert_close(got, want, rtol=6e-2, atol=6e-2)
    print(f"tp{world} fused={fused} nvls requested={nvls} used={multi[0][2]}: logits match the single-GPU engine")
    torch.testing.assert_close(multi[0][0][0], multi[world - 1][0][0], rtol=2e-2, atol=2e-2)  # ranks agree up to reduction order
