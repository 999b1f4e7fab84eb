"""Builds the bench workload (Llama-2-7B, B sessions restored at ctx tokens), then runs a few decode iterations and
one restore+prefill inside NVTX ranges so ncu can list per-kernel device time for exactly those phases:
  ncu --nvtx --nvtx-include "decode/" --metrics gpu__time_duration.sum --csv ...
Also prints CUDA-event timings of the same phases (valid only when NOT run under ncu)."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from hcache_deepspeed_b200.inference.v2 import RaggedInferenceEngineConfig, build_engine_from_arch  # noqa: E402
from hcache_deepspeed_b200.models.arch import named_arch  # noqa: E402
from hcache_deepspeed_b200.ops import kernels  # noqa: E402

B = int(os.environ.get("B", 8))
ctx = int(os.environ.get("CTX", 4096))
n_new, D = 64, int(os.environ.get("D", 4))
layers = int(os.environ.get("LAYERS", 32))
dev = torch.device("cuda", 0)
spec = named_arch(os.environ.get("MODEL", "llama-2-7b")).scaled(n_layers=layers)
# per-rank shapes of a tensor-parallel run on one GPU (no collectives): e.g. HEADS=4 KV_HEADS=4 INTER=1376 VOCAB=4000 B=64
# is one rank of Llama-2-7B at TP=8 with 64 sessions
over = {k: int(os.environ[e]) for k, e in (("n_heads", "HEADS"), ("n_kv_heads", "KV_HEADS"), ("intermediate", "INTER"),
                                           ("vocab", "VOCAB")) if e in os.environ}
if over:
    spec = spec.scaled(**over)
max_ctx = ctx + n_new + 64 + 64
cfg = RaggedInferenceEngineConfig(state_manager={
    "max_tracked_sequences": max(32, 2 * B), "max_ragged_batch_size": max(ctx, B * n_new) + 64, "max_ragged_sequence_count": max(B, 8),
    "max_context": max_ctx, "memory_config": {"mode": "allocate", "size": int((B + 1) * (max_ctx + 128) * spec.kv_bytes_per_token_layer() * layers)}},
    hcache={"enabled": True, "capture": "off"}, use_cuda_graph_decode=os.environ.get("GRAPH", "1") == "1")
eng = build_engine_from_arch(spec, cfg, seed=1, device=dev)
lat = torch.empty((layers, ctx, spec.hidden), dtype=torch.bfloat16, pin_memory=True)
lat.copy_((torch.randn(lat.shape) * 0.5).to(torch.bfloat16))
ctx_tok = torch.randint(0, spec.vocab, (ctx, ))
new_tok = torch.randint(0, spec.vocab, (n_new, ))
ev = lambda: torch.cuda.Event(enable_timing=True)  # noqa: E731


def restore_prefill(uid):
    eng.restore_kv([uid], [ctx_tok], [lat], do_checks=False)
    logits, _ = eng.put([uid], [new_tok], do_checks=False)
    return kernels.argmax(logits).cpu()


uids = list(range(B))
cur = [restore_prefill(u) for u in uids]
for _ in range(2):  # warm
    logits, _ = eng.put(uids, cur, do_checks=False)
    nxt = kernels.argmax(logits).cpu()
    cur = [nxt[i:i + 1] for i in range(B)]
torch.cuda.synchronize()
s, e = ev(), ev()
torch.cuda.nvtx.range_push("decode")
s.record()
for _ in range(D):
    logits, _ = eng.put(uids, cur, do_checks=False)
    nxt = kernels.argmax(logits).cpu()
    cur = [nxt[i:i + 1] for i in range(B)]
e.record()
torch.cuda.synchronize()
torch.cuda.nvtx.range_pop()
print(f"decode: {s.elapsed_time(e) / D:.3f} ms/iter (B={B}, ctx={ctx}, layers={layers})")
restore_prefill(100)
eng.flush(100)
torch.cuda.synchronize()
s, e = ev(), ev()
torch.cuda.nvtx.range_push("ttft")
s.record()
restore_prefill(101)
e.record()
torch.cuda.synchronize()
torch.cuda.nvtx.range_pop()
print(f"restore+prefill TTFT: {s.elapsed_time(e):.3f} ms")
# restore only / prefill only
eng.flush(101)
s, m, e = ev(), ev(), ev()
s.record()
eng.restore_kv([102], [ctx_tok], [lat], do_checks=False)
m.record()
logits, _ = eng.put([102], [new_tok], do_checks=False)
e.record()
torch.cuda.synchronize()
print(f"restore: {s.elapsed_time(m):.3f} ms  prefill(new turn): {m.elapsed_time(e):.3f} ms  "
      f"H2D bytes {lat.numel() * 2 / 1e9:.3f} GB -> {lat.numel() * 2 / 1e6 / s.elapsed_time(m):.1f} GB/s")
