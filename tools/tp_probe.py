"""Kernel-level check of hds_norm_tp on real GPUs: every combination of {fp32, 16-bit partial sums} x {all-reduce,
reduce-scatter + all-gather} x {peer loads/stores, multimem.ld_reduce, multimem.st} against a torch reference built
from all ranks' partial sums (NCCL all_gather), plus timings of one reduction point per shape.
usage: python tools/tp_probe.py N   (spawns N ranks)"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def rank_fn(rank, world):
    import torch
    import torch.distributed as dist
    from hcache_deepspeed_b200.ops import kernels
    from hcache_deepspeed_b200.parallel.tp_comm import FusedTPComm
    torch.cuda.set_device(rank)
    dev = torch.device("cuda", rank)
    dist.init_process_group("nccl", device_id=dev)
    comm = FusedTPComm(dist.group.WORLD, world, rank, dev)
    H, Tmax = 4096, 4160
    p32 = comm.alloc_symmetric((Tmax, H), torch.float32)
    p16 = comm.alloc_symmetric((Tmax, H), torch.bfloat16)
    out_sym = comm.alloc_symmetric((Tmax, H), torch.bfloat16)
    resid = torch.randn(Tmax, H, device=dev, generator=torch.Generator(device=dev).manual_seed(1)).to(torch.bfloat16)
    gamma = torch.ones(H, dtype=torch.bfloat16, device=dev)
    g = torch.Generator(device=dev).manual_seed(100 + rank)
    lines = []
    for T in (8, 64, 340, 4160):
        part = torch.randn(T, H, device=dev, generator=g) * 0.5
        all_parts = [torch.empty_like(part) for _ in range(world)]
        dist.all_gather(all_parts, part)
        for p16bit in (False, True):
            src = p16 if p16bit else p32
            parts = [p.to(torch.bfloat16).float() if p16bit else p for p in all_parts]
            total = sum(parts)
            x = (resid[:T].float() + total).to(torch.bfloat16).float()
            want = (x * torch.rsqrt((x * x).mean(-1, keepdim=True) + 1e-5)).to(torch.bfloat16).float()
            per = -(-T // world)
            r0 = min(T, rank * per)
            n = min(T, r0 + per) - r0
            for scatter in (False, True):
                for ld in (False, True):
                    for st in ((False, True) if scatter else (False, )):
                        src[:T].copy_(part)
                        out_sym[:T].zero_()
                        res = torch.empty(T, H, dtype=torch.bfloat16, device=dev)
                        torch.cuda.synchronize()
                        dist.barrier()
                        comm.signal()
                        kw = {}
                        if scatter:
                            kw = dict(row0=r0, n_rows=n, signal=True, out_mc=comm.multicast_ptr(out_sym[:T]) if st else None,
                                      out_ptrs=None if st else comm.peer_ptrs_raw(out_sym[:T]))
                        kernels.norm_tp(comm.ctx, resid[:T], gamma, out_sym[:T], comm.peer_ptrs_raw(src[:T]),
                                        src_mc=comm.multicast_ptr(src[:T]) if ld else None, partial_16bit=p16bit, res_out=res,
                                        eps=1e-5, **kw)
                        if scatter:
                            comm.wait()
                        torch.cuda.synchronize()
                        dist.barrier()
                        got = out_sym[:T].float()
                        err = (got - want).abs().max().item()
                        bad = int(((got - want).abs() > 0.05).sum())
                        # timing: 20 back-to-back reduction points (signal kernel + norm [+ wait kernel])
                        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                        dist.barrier()
                        e0.record()
                        for _ in range(20):
                            comm.signal()
                            kernels.norm_tp(comm.ctx, resid[:T], gamma, out_sym[:T], comm.peer_ptrs_raw(src[:T]),
                                            src_mc=comm.multicast_ptr(src[:T]) if ld else None, partial_16bit=p16bit,
                                            res_out=res, eps=1e-5, **kw)
                            if scatter:
                                comm.wait()
                        e1.record()
                        torch.cuda.synchronize()
                        us = e0.elapsed_time(e1) * 1000 / 20
                        lines.append(f"T={T:5d} partial={'bf16' if p16bit else 'fp32'} shape={'rs+ag' if scatter else 'allred'} "
                                     f"ld={'multimem' if ld else 'peer'} st={'multimem' if st else ('peer' if scatter else '-')} "
                                     f"max_err={err:.4f} bad={bad} us_per_point={us:.1f}")
    dist.barrier()
    return lines


if __name__ == "__main__":
    from tests.dist_utils import run_distributed
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 2
    res = run_distributed(rank_fn, n, timeout=600.0)
    for r in (0, n - 1):
        print(f"---- rank {r}")
        for ln in res[r]:
            print(ln)
