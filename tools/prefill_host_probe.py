"""Where the non-kernel time of a fresh prefill goes: host enqueue time of put() vs device time, cProfile of put()."""
import cProfile
import os
import pstats
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from hcache_deepspeed_b200.inference.v2 import RaggedInferenceEngineConfig, build_engine_from_arch  # noqa: E402
from hcache_deepspeed_b200.models.arch import named_arch  # noqa: E402

dev = torch.device("cuda", 0)
spec = named_arch("llama-2-7b")
n_tok = 4160
cfg = RaggedInferenceEngineConfig(
    state_manager={"max_tracked_sequences": 16, "max_ragged_batch_size": n_tok + 64, "max_ragged_sequence_count": 8,
                   "max_context": n_tok + 256,
                   "memory_config": {"mode": "allocate", "size": int(3 * (n_tok + 256) * spec.kv_bytes_per_token_layer() * spec.n_layers)}},
    hcache={"enabled": True, "capture": "off"})
eng = build_engine_from_arch(spec, cfg, seed=1234, dtype=torch.bfloat16, device=dev)
toks = torch.randint(0, spec.vocab, (n_tok, )).pin_memory()
for i in range(4):
    eng.put([i], [toks], do_checks=False)
    torch.cuda.synchronize()
    eng.flush(i)
for i in range(4, 8):
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    logits, _ = eng.put([i], [toks], do_checks=False)
    t1 = time.perf_counter()
    torch.cuda.synchronize()
    t2 = time.perf_counter()
    eng.flush(i)
    print(f"put() returned after {1e3 * (t1 - t0):.2f} ms host time; device done at {1e3 * (t2 - t0):.2f} ms", flush=True)
pr = cProfile.Profile()
pr.enable()
logits, _ = eng.put([9], [toks], do_checks=False)
pr.disable()
torch.cuda.synchronize()
pstats.Stats(pr).sort_stats("cumulative").print_stats(22)
