"""Tensor parallelism on ONE GPU: ``LoopbackWorld`` runs the flag-protocol collectives (GEMM tail signals, the
all-reduce / reduce-scatter + all-gather norm kernel, all-gather -> GEMM waits, logits gather, sharded HCache restore)
for 2 / 4 / 8 emulated ranks — one host thread per rank, one stream, a host rendezvous before every waiting launch, so
no kernel ever spins on a flag that has not been produced. The logits must match the single-rank engine built from the
same full weights. (Real multi-GPU runs of the same script: tests/test_gpu_tp.py.)"""
import threading

import pytest
import torch

pytestmark = [pytest.mark.gpu, pytest.mark.timeout(600)]

LLAMA = dict(model_type="llama", n_layers=3, hidden=1024, n_heads=8, n_kv_heads=8, head_dim=128, intermediate=2816, vocab=2048,
             norm="rms", act="silu", gated_mlp=True, pos_emb="rope", max_positions=2048)
GQA = dict(LLAMA, n_kv_heads=2)       # fewer KV heads than ranks at tp=4/8: KV heads are replicated
OPT = dict(model_type="opt", n_layers=2, hidden=1024, n_heads=16, n_kv_heads=16, head_dim=64, intermediate=4096, vocab=2048,
           norm="layer", act="relu", gated_mlp=False, pos_emb="learned", max_positions=2048, learned_pos_offset=2,
           qkv_bias=True, mlp_bias=True, attn_out_bias=True, norm_bias=True)


def _spec(kwargs):
    from hcache_deepspeed_b200.models.arch import ArchSpec
    fields = ArchSpec.__dataclass_fields__
    return ArchSpec(**{k: v for k, v in kwargs.items() if k in fields})


def _build_engine(spec, full, rank, world, comm, quant=None):
    from hcache_deepspeed_b200.inference.v2 import RaggedInferenceEngineConfig
    from hcache_deepspeed_b200.inference.v2.engine_v2 import InferenceEngineV2
    from hcache_deepspeed_b200.inference.v2.model_implementations import InferenceV2Policy
    from hcache_deepspeed_b200.models.transformer import RaggedTransformer, TPComm
    from tests.tp_shard_helper import shard_full_weights
    dev = torch.device("cuda", torch.cuda.current_device())
    cfg = RaggedInferenceEngineConfig(state_manager=dict(
        max_tracked_sequences=8, max_ragged_batch_size=768, max_ragged_sequence_count=4, max_context=2048,
        memory_config=dict(mode="allocate", size=192 * 1024 * 1024)), use_cuda_graph_decode=False)

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
            model = RaggedTransformer(spec, w, engine_config.state_manager, dtype=torch.bfloat16, device=dev, tp_rank=rank,
                                      tp_size=world, comm=comm if comm is not None else TPComm(size=1, rank=0))
            if quant:
                model.quantize_weights_fp8()
            return model

    return InferenceEngineV2(FromFull(arch=spec, random_seed=0), cfg)


def _script(eng, spec, rank, world, comm, shared):
    """prefill (340 tokens: 16-bit partial sums, reduce-scatter + all-gather) -> 3 decode steps (one-shot all-reduce) ->
    a 40-token continuation (fp32 partial sums, reduce-scatter + all-gather) -> HCache restore (replicated and sharded)."""
    g = torch.Generator().manual_seed(3)
    a, b = torch.randint(0, spec.vocab, (300, ), generator=g), torch.randint(0, spec.vocab, (40, ), generator=g)
    logits, lat = eng.put([1, 2], [a, b])
    outs = [logits.float().cpu()]
    # decode inputs are fixed tokens, not each run's own arg-max: with random weights the top logits are nearly tied
    # and a rounding-level difference would otherwise send the runs down different token paths
    for step in range(3):
        cur = torch.randint(0, spec.vocab, (2, ), generator=g)
        lg, _ = eng.put([1, 2], [cur[0:1], cur[1:2]], capture_latents=False)
        outs.append(lg.float().cpu())
    mid = torch.randint(0, spec.vocab, (40, ), generator=g)
    lg, _ = eng.put([1], [mid], capture_latents=False)
    outs.append(lg.float().cpu())
    if rank == 0:
        shared["lat"] = lat[0]
    if comm is not None:
        comm.pre_wait()
    lat_a = shared["lat"]
    eng.restore_kv([7], [a], [lat_a.pin_memory()])
    probe = torch.randint(0, spec.vocab, (9, ), generator=g)
    lb, _ = eng.put([7], [probe], capture_latents=False)
    outs.append(lb.float().cpu())
    if world > 1:
        lo, hi = eng.latent_row_range(int(a.numel()))
        mine = lat_a[:, lo:hi].contiguous().pin_memory()
        for defer in (False, True):
            uid = 20 + int(defer)
            eng.restore_kv([uid], [a], [mine], sharded=True, defer=defer)
            lc, _ = eng.put([uid], [probe], capture_latents=False)
            torch.testing.assert_close(lc.float().cpu(), outs[-1], rtol=3e-2, atol=3e-2)
    torch.cuda.synchronize()
    return outs


def _run_world(arch, tp, quant=None, seed=5):
    from hcache_deepspeed_b200.models.weights import random_weights
    from hcache_deepspeed_b200.ops import native
    from hcache_deepspeed_b200.parallel.tp_comm import LoopbackWorld
    native.cuda_lib()
    spec = _spec(arch)
    full = random_weights(spec, 0, 1, torch.bfloat16, torch.device("cpu"), seed=seed)
    dev = torch.device("cuda", torch.cuda.current_device())
    single = _script(_build_engine(spec, full, 0, 1, None, quant), spec, 0, 1, None, {})
    world = LoopbackWorld(tp, dev)
    shared = {}
    multi = world.run(lambda r, comm: _script(_build_engine(spec, full, r, tp, comm, quant), spec, r, tp, comm, shared))
    torch.cuda.synchronize()
    for c in world.comms:
        c.close()
    return single, multi


@pytest.mark.parametrize("tp", [2, 4, 8])
def test_loopback_llama(tp):
    single, multi = _run_world(LLAMA, tp)
    for r in range(tp):
        for want, got in zip(single, multi[r]):
            torch.testing.assert_close(got, want, rtol=6e-2, atol=6e-2)
    for r in range(1, tp):   # ranks agree with one another far more tightly than with the differently-sharded single run
        torch.testing.assert_close(multi[r][0], multi[0][0], rtol=2e-2, atol=2e-2)


def test_loopback_gqa_replicated_kv_heads():
    single, multi = _run_world(GQA, 4)
    for r in range(4):
        for want, got in zip(single, multi[r]):
            torch.testing.assert_close(got, want, rtol=6e-2, atol=6e-2)


def test_loopback_opt_layernorm_bias():
    single, multi = _run_world(OPT, 2)
    for r in range(2):
        for want, got in zip(single, multi[r]):
            torch.testing.assert_close(got, want, rtol=6e-2, atol=6e-2)


def test_loopback_fp8_tp():
    """wf8af8 under tensor parallelism: the fp8 GEMMs signal, the activation quantiser waits; logits track the bf16 run."""
    bf16, _ = _run_world(LLAMA, 1 + 1)
    single8, multi8 = _run_world(LLAMA, 2, quant=True)
    for r in range(2):
        for want, got in zip(single8, multi8[r]):
            cos = torch.nn.functional.cosine_similarity(got.flatten(), want.flatten(), dim=0)
            assert cos > 0.97, float(cos)
    cos = torch.nn.functional.cosine_similarity(multi8[0][0].flatten(), bf16[0].flatten(), dim=0)
    assert cos > 0.95, float(cos)


def test_flag_counters_and_gather_cols():
    """Protocol bookkeeping: n signals leave n in every peer's arrive slot and in `sent`; the column gather assembles
    uneven shards."""
    from hcache_deepspeed_b200.ops import kernels
    from hcache_deepspeed_b200.parallel.tp_comm import LoopbackWorld
    dev = torch.device("cuda", torch.cuda.current_device())
    world = LoopbackWorld(3, dev)
    widths = [40, 24, 33]
    col0 = [0, 40, 64]
    rows = 5

    def body(r, comm):
        loc = comm.alloc_symmetric((rows, 40), torch.bfloat16)
        loc[:, :widths[r]] = torch.arange(rows * widths[r], device=dev).reshape(rows, -1).to(torch.bfloat16) + 1000 * r
        for _ in range(4):
            comm.signal()
        out = torch.zeros(rows, sum(widths), dtype=torch.bfloat16, device=dev)
        comm.gather_logits(loc[:, :widths[r]], out, [(c, c + w) for c, w in zip(col0, widths)])
        torch.cuda.synchronize()
        return out.float().cpu(), comm.pad.cpu()

    res = world.run(body)
    want = torch.cat([torch.arange(rows * w).reshape(rows, -1).float() + 1000 * r for r, w in enumerate(widths)], dim=1)
    want = want.to(torch.bfloat16).float()
    for r in range(3):
        out, pad = res[r]
        torch.testing.assert_close(out, want, rtol=0, atol=0)
        assert int(pad[33]) == 4                         # sent
        assert int(pad[32]) == 0                         # ticket back at zero
        for p in range(3):
            if p != r:
                assert int(pad[p]) == 4, (r, p, pad[:8])
    assert not kernels.tp_disarm()
