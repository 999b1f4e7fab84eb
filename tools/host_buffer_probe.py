"""Host time of building + uploading a 64-sequence decode batch's metadata with the three staging-memory kinds
(HDS_FAST_HOST_BUFFER unset = pinned torch memory, 'mapped', 'wc' = + write-combined). One GPU:
    python tools/host_buffer_probe.py >> gpurun_out/host_buffer_probe.jsonl"""
import json
import os
import sys
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def run(mode):
    if mode:
        os.environ["HDS_FAST_HOST_BUFFER"] = mode
    else:
        os.environ.pop("HDS_FAST_HOST_BUFFER", None)
    from hcache_deepspeed_b200.inference.v2 import RaggedInferenceEngineConfig, build_engine_from_arch
    from hcache_deepspeed_b200.models.arch import named_arch
    spec = named_arch("llama-2-7b").scaled(n_layers=1)
    B, ctx = 64, 2048
    cfg = RaggedInferenceEngineConfig(state_manager={
        "max_tracked_sequences": B, "max_ragged_batch_size": 4096, "max_ragged_sequence_count": B, "max_context": 4096,
        "memory_config": {"mode": "allocate", "size": 8 << 30}}, hcache={"enabled": False}, use_cuda_graph_decode=False)
    eng = build_engine_from_arch(spec, cfg, seed=1, dtype=torch.bfloat16, device=torch.device("cuda"))
    g = torch.Generator().manual_seed(0)
    for u in range(B):
        eng.put([u], [torch.randint(0, spec.vocab, (ctx, ), generator=g)])
    uids = list(range(B))
    toks = [torch.randint(0, spec.vocab, (1, ), generator=g) for _ in range(B)]
    batch, sm, model = eng._batch, eng._state_manager, eng._model

    def build():
        batch.clear()
        for uid, t in zip(uids, toks):
            seq = sm.get_or_create_sequence(uid)
            model.maybe_allocate_kv(seq, 1)
            seq.pre_forward(1)
            batch.insert_sequence(seq, t, do_checks=False)
        batch.finalize()
        for uid in uids:                       # undo pre_forward: the probe does not run the forward
            sm.get_or_create_sequence(uid)._in_flight_tokens = 0

    for _ in range(20):
        build()
    torch.cuda.synchronize()
    n = 300
    t0 = time.perf_counter()
    for _ in range(n):
        build()
    host = (time.perf_counter() - t0) / n
    torch.cuda.synchronize()
    total = (time.perf_counter() - t0) / n
    # a full decode step for scale
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(50):
        eng.put(uids, toks, do_checks=False)
    torch.cuda.synchronize()
    step = (time.perf_counter() - t0) / 50
    print(json.dumps({"staging": mode or "pinned", "batch_build_host_us": round(host * 1e6, 1), "batch_build_synced_us": round(total * 1e6, 1),
                      "decode_step_1layer_us": round(step * 1e6, 1), "sequences": B}), flush=True)


if __name__ == "__main__":
    run(sys.argv[1] if len(sys.argv) > 1 and sys.argv[1] != "pinned" else None)
