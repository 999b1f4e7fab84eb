"""Fresh-prompt prefill latency of the ragged engine (no hidden-state cache): Llama-2-7B, one 4160-token prompt per put().
Complements bench.py, whose TTFT is the RESTORED-context path. usage: python tools/prefill_probe.py [HDS_GEMM_2SM=0|1]"""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from hcache_deepspeed_b200.inference.v2 import RaggedInferenceEngineConfig, build_engine_from_arch  # noqa: E402
from hcache_deepspeed_b200.models.arch import named_arch  # noqa: E402

dev = torch.device("cuda", 0)
torch.cuda.set_device(0)
spec = named_arch("llama-2-7b")
n_tok = int(os.environ.get("PROBE_TOKENS", 4160))
cfg = RaggedInferenceEngineConfig(
    state_manager={"max_tracked_sequences": 16, "max_ragged_batch_size": n_tok + 64, "max_ragged_sequence_count": 8,
                   "max_context": n_tok + 256,
                   "memory_config": {"mode": "allocate", "size": int(3 * (n_tok + 256) * spec.kv_bytes_per_token_layer() * spec.n_layers)}},
    hcache={"enabled": True, "capture": "off"},
    quantization={"quantization_mode": os.environ.get("PROBE_QUANT") or None})
eng = build_engine_from_arch(spec, cfg, seed=1234, dtype=torch.bfloat16, device=dev)
g = torch.Generator().manual_seed(3)
toks = torch.randint(0, spec.vocab, (n_tok, ), generator=g).pin_memory()
times = []
for i in range(8):
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    logits, _ = eng.put([i], [toks], do_checks=False)
    nxt = int(logits[0].argmax())           # device -> host read of the result, as a client would
    e1.record()
    torch.cuda.synchronize()
    eng.flush(i)
    times.append(e0.elapsed_time(e1))
steady = sorted(times[3:])
print(json.dumps({"case": "fresh_prefill", "model": "llama-2-7b", "tokens": n_tok, "ms_median": round(steady[len(steady) // 2], 3),
                  "ms_min": round(steady[0], 3), "tokens_per_s": round(n_tok / steady[len(steady) // 2] * 1e3, 1),
                  "gemm_2sm": os.environ.get("HDS_GEMM_2SM", "1"),
                  "quant": os.environ.get("PROBE_QUANT", "none")}))
