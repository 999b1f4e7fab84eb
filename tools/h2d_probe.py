"""Host->device bandwidth probe: each rank copies a pinned buffer to its GPU, alone and then all ranks at once
(torchrun). Tells whether per-GPU PCIe links add up on this box — the scaling limit of TP-sharded restore."""
import os
import time

import torch
import torch.distributed as dist

rank, world, local = (int(os.environ.get(k, d)) for k, d in (("RANK", 0), ("WORLD_SIZE", 1), ("LOCAL_RANK", 0)))
torch.cuda.set_device(local)
if world > 1:
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
n = 1 << 30
host = torch.empty(n, dtype=torch.uint8, pin_memory=True)
devb = torch.empty(n, dtype=torch.uint8, device="cuda")


def run(iters=4):
    torch.cuda.synchronize()
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record()
    for _ in range(iters):
        devb.copy_(host, non_blocking=True)
    e.record()
    torch.cuda.synchronize()
    return iters * n / (s.elapsed_time(e) / 1e3) / 1e9


run(1)
for r in range(world):  # one rank at a time
    if world > 1:
        dist.barrier()
    if r == rank:
        print(f"rank {rank} alone: {run():.1f} GB/s", flush=True)
if world > 1:
    dist.barrier()
    bw = run()
    t = torch.tensor([bw], device="cuda")
    dist.all_reduce(t)
    print(f"rank {rank} concurrent: {bw:.1f} GB/s (sum over ranks {t.item():.1f} GB/s)", flush=True)
    dist.barrier()
    torch.cuda.synchronize()
os._exit(0)
