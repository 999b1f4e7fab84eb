"""TP timing breakdown (run under torchrun): barrier cost, sharded restore, new-turn prefill, decode iteration."""
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from hcache_deepspeed_b200.inference.v2 import RaggedInferenceEngineConfig, build_engine_from_arch  # noqa: E402
from hcache_deepspeed_b200.models.arch import named_arch  # noqa: E402
from hcache_deepspeed_b200.ops import kernels  # noqa: E402

rank, world, local = (int(os.environ.get(k, d)) for k, d in (("RANK", 0), ("WORLD_SIZE", 1), ("LOCAL_RANK", 0)))
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
if world > 1:
    dist.init_process_group("nccl", device_id=dev)
B, ctx, n_new = 8, 4096, 64
spec = named_arch("llama-2-7b")
max_ctx = ctx + n_new + 128
cfg = RaggedInferenceEngineConfig(tensor_parallel={"tp_size": world}, state_manager={
    "max_tracked_sequences": 32, "max_ragged_batch_size": ctx + 64, "max_ragged_sequence_count": 8, "max_context": max_ctx,
    "memory_config": {"mode": "allocate", "size": int((B + 1) * (max_ctx + 128) * spec.kv_bytes_per_token_layer() * spec.n_layers / world)}},
    hcache={"enabled": True, "capture": "off"}, use_cuda_graph_decode=os.environ.get("GRAPH", "1") == "1")
eng = build_engine_from_arch(spec, cfg, seed=1, device=dev)
model = eng.model()
fused = bool(getattr(model.comm, "fused", False))
lat = torch.empty((spec.n_layers, ctx, spec.hidden), dtype=torch.bfloat16, pin_memory=True).normal_(0, 0.5)
lo, hi = eng.latent_row_range(ctx)
mine = lat[:, lo:hi].contiguous().pin_memory()
ctx_tok, new_tok = torch.randint(0, spec.vocab, (ctx, )), torch.randint(0, spec.vocab, (n_new, ))
ev = lambda: torch.cuda.Event(enable_timing=True)  # noqa: E731


def timed(fn, n=1):
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    s, e = ev(), ev()
    s.record()
    for _ in range(n):
        fn()
    e.record()
    torch.cuda.synchronize()
    return s.elapsed_time(e) / n


def say(msg):
    if rank == 0:
        print(msg, flush=True)


if fused:
    say(f"symm barrier: {timed(model.comm.barrier, 50) * 1000:.1f} us")
    d32 = model._delta_bufs[0][:64]
    say(f"all_reduce_delta barrier: {timed(lambda: model.comm.all_reduce_delta(d32), 50) * 1000:.1f} us")
uid = [100]


def restore(defer, sharded):
    uid[0] += 1
    if sharded:
        eng.restore_kv([uid[0]], [ctx_tok], [mine], do_checks=False, sharded=True, defer=defer)
    else:
        eng.restore_kv([uid[0]], [ctx_tok], [lat], do_checks=False, defer=defer)


for sharded in ([False, True] if fused else [False]):
    restore(False, sharded)
    eng.flush(uid[0])
    t = timed(lambda: restore(False, sharded))
    say(f"restore (sharded={sharded}, no defer): {t:.2f} ms")
    t2 = timed(lambda: eng.put([uid[0]], [new_tok], do_checks=False))
    say(f"  new-turn prefill: {t2:.2f} ms")
    eng.flush(uid[0])

    def both():
        restore(True, sharded)
        eng.put([uid[0]], [new_tok], do_checks=False)
    both()
    eng.flush(uid[0])
    say(f"restore(defer)+prefill (sharded={sharded}): {timed(both):.2f} ms")
    eng.flush(uid[0])
uids = list(range(B))
for u in uids:
    eng.restore_kv([u], [ctx_tok], [mine if fused else lat], do_checks=False, sharded=fused)
    eng.put([u], [new_tok], do_checks=False)
cur = [torch.tensor([1]) for _ in uids]


def dec():
    lg, _ = eng.put(uids, cur, do_checks=False)
    kernels.argmax(lg).cpu()


for _ in range(4):
    dec()
say(f"decode iteration (B={B}): {timed(dec, 8):.3f} ms  graph_replays={eng.graph_replays}")
if world > 1:
    dist.barrier()
torch.cuda.synchronize()
os._exit(0)
