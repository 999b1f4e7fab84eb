"""Decode-iteration time under tensor parallelism and what the flag exchanges cost (run under torchrun).
One process per GPU builds the TP engine once, fills B sessions of ctx tokens, then times graph-replayed decode iterations
for several settings switched at run time: default, no flag waits/signals (WRONG results, timing only), one-shot vs
reduce-scatter norm shape, NVLS on/off.   usage: torchrun ... tools/tp_decode_probe.py [--model M] [--sessions-per-gpu S] [--ctx C]"""
import argparse
import json
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from hcache_deepspeed_b200.inference.v2 import RaggedInferenceEngineConfig, build_engine_from_arch  # noqa: E402
from hcache_deepspeed_b200.models import transformer as tr  # noqa: E402
from hcache_deepspeed_b200.models.arch import named_arch  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--model", default="llama-2-7b")
ap.add_argument("--sessions-per-gpu", type=int, default=8)
ap.add_argument("--ctx", type=int, default=4096)
ap.add_argument("--iters", type=int, default=24)
ap.add_argument("--strong", action="store_true", help="also time a fixed 8-session batch")
a = ap.parse_args()
rank, world, local = (int(os.environ.get(k, d)) for k, d in (("RANK", 0), ("WORLD_SIZE", 1), ("LOCAL_RANK", 0)))
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
os.environ.setdefault("HDS_TP_STRICT", "1")
if world > 1:
    dist.init_process_group("nccl", device_id=dev)
spec = named_arch(a.model)
B, ctx = a.sessions_per_gpu * world, a.ctx
max_ctx = ctx + 1024
cfg = RaggedInferenceEngineConfig(tensor_parallel={"tp_size": world}, state_manager={
    "max_tracked_sequences": 2 * B, "max_ragged_batch_size": ctx + 64, "max_ragged_sequence_count": B, "max_context": max_ctx,
    "memory_config": {"mode": "allocate", "size": int((B + 1) * (max_ctx + 128) * spec.kv_bytes_per_token_layer() * spec.n_layers / world)}},
    hcache={"enabled": False})
eng = build_engine_from_arch(spec, cfg, seed=1, device=dev)
model = eng.model()
comm = model.comm
g = torch.Generator().manual_seed(0)
for u in range(B):
    eng.put([u], [torch.randint(0, spec.vocab, (ctx, ), generator=g)], do_checks=False)
torch.cuda.synchronize()
ev = lambda: torch.cuda.Event(enable_timing=True)  # noqa: E731


def decode_ms(n_seqs, label):
    eng._decode_graphs.clear()
    eng._decode_warm.clear()
    uids = list(range(n_seqs))
    toks = [torch.randint(0, spec.vocab, (1, ), generator=g) for _ in uids]
    for _ in range(4):      # eager warm-up + graph capture
        eng.put(uids, toks, do_checks=False)
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    s, e = ev(), ev()
    s.record()
    for _ in range(a.iters):
        eng.put(uids, toks, do_checks=False)
    e.record()
    torch.cuda.synchronize()
    t = torch.tensor([s.elapsed_time(e) / a.iters], device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    if rank == 0:
        w_gb = model.w.n_bytes() / 1e9
        kv_gb = n_seqs * (ctx + 8) * spec.kv_bytes_per_token_layer() * spec.n_layers / world / 1e9
        print(json.dumps({"case": label, "tp": world, "sessions": n_seqs, "ms_per_iteration": round(t.item(), 4),
                          "tokens_per_s": round(n_seqs / t.item() * 1e3, 1), "hbm_floor_ms": round((w_gb + kv_gb) / 6.5, 3),
                          "graph_replays": eng.graph_replays}), flush=True)
    return t.item()


batches = [B] + ([8] if a.strong and B != 8 else [])
for n in batches:
    decode_ms(n, "default")
    if world > 1 and getattr(comm, "fused", False):
        comm.debug_nosync = True
        decode_ms(n, "no waits/signals (invalid results: timing only)")
        comm.debug_nosync = False
        keep = tr.TP_ONESHOT_MAX_TOKENS
        tr.TP_ONESHOT_MAX_TOKENS = 1 << 20
        decode_ms(n, "one-shot all-reduce norm")
        tr.TP_ONESHOT_MAX_TOKENS = 0
        decode_ms(n, "reduce-scatter + all-gather norm")
        tr.TP_ONESHOT_MAX_TOKENS = keep
        if comm.use_nvls:
            comm.use_nvls = False
            decode_ms(n, "default, NVLS off")
            tr.TP_ONESHOT_MAX_TOKENS = 1 << 20
            decode_ms(n, "one-shot, NVLS off")
            tr.TP_ONESHOT_MAX_TOKENS = 0
            decode_ms(n, "rs+ag, NVLS off")
            tr.TP_ONESHOT_MAX_TOKENS = keep
            comm.use_nvls = True
if world > 1:
    dist.barrier()
    torch.cuda.synchronize()
    sys.stdout.flush()
    os._exit(0)
