"""2-GPU check of the training side on real NCCL: ZeRO-3 with trace-driven async prefetch + the schedule compiler
(selective gather, prefetch plan, optimizer-state offload) must reproduce single-process AdamW.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29611 tools/zero3_gpu_check.py
"""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import hcache_deepspeed_b200 as ds  # noqa: E402
from tests.runtime_helper import TinyNet, make_batches, reference_run  # noqa: E402


def main():
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    dev = os.environ.get("HDS_CHECK_DEVICE", "cuda")         # "cpu" = dry run over gloo
    if dev == "cuda":
        torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", 0)))
    ds.init_distributed(dist_backend="nccl" if dev == "cuda" else "gloo")
    steps, micro, hidden = 8, 4, 32
    want_losses, want_state = reference_run(steps=steps, global_bs=micro * world)
    torch.manual_seed(1234)
    model = TinyNet(hidden).to(dev)
    cfg = {"train_micro_batch_size_per_gpu": micro, "gradient_accumulation_steps": 1, "steps_per_print": 1000,
           "optimizer": {"type": "AdamW", "params": {"lr": 1e-2, "weight_decay": 0.0}},
           "zero_optimization": {"stage": 3, "stage3_param_persistence_threshold": 0, "stage3_prefetch_bucket_size": 100000},
           "compile": {"deepcompile": True, "selective_gather_budget_bytes": 3000, "prefetch_bucket_size": 100000,
                       "offload_opt_states": True}}
    engine, _, _, _ = ds.initialize(model=model, config=cfg)
    engine.compile()
    losses = []
    for x, y in make_batches(steps, micro * world, hidden):
        lo = rank * micro
        loss = engine(x[lo:lo + micro].to(dev), y[lo:lo + micro].to(dev))
        engine.backward(loss)
        engine.step()
        losses.append(float(loss))
    from hcache_deepspeed_b200.runtime.zero import GatheredParameters
    worst = 0.0
    for n, p in engine.module.named_parameters():
        with GatheredParameters([p]):
            worst = max(worst, float((p.data.float().cpu() - want_state[n]).abs().max()))
    t = torch.tensor(losses, device=dev)
    torch.distributed.all_reduce(t)
    mean = (t / world).tolist()
    off = engine.optimizer.parameter_offload
    out = {"rank": rank, "max_param_diff": worst, "max_loss_diff": max(abs(a - b) for a, b in zip(mean, want_losses)),
           "prefetched": off.stats["prefetched"], "trace_invalidations": off.stats["trace_invalidations"],
           "compile_report": engine.compile_report, "backend": torch.distributed.get_backend()}
    if rank == 0:
        print(json.dumps(out))
    assert worst < 2e-4 and out["max_loss_diff"] < 5e-4 and off.stats["prefetched"] > 0, out
    torch.distributed.barrier()


if __name__ == "__main__":
    main()
