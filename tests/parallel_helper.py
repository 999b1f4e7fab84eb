"""Worker functions for the MoE / sequence-parallel multi-rank tests."""
import math

import torch


def _attn(q, k, v):
    # q, k, v: [b, s, h, d] full sequence, causal
    qh, kh, vh = (t.permute(0, 2, 1, 3) for t in (q, k, v))
    s = torch.matmul(qh, kh.transpose(-1, -2)) / math.sqrt(q.shape[-1])
    n = q.shape[1]
    s = s.masked_fill(torch.triu(torch.ones(n, n, dtype=torch.bool), 1), float("-inf"))
    return torch.matmul(torch.softmax(s, -1), vh).permute(0, 2, 1, 3)


def ulysses_worker(rank, world, heads):
    import hcache_deepspeed_b200 as ds
    from hcache_deepspeed_b200 import comm as dist
    from hcache_deepspeed_b200.sequence.layer import DistributedAttention
    ds.init_distributed(dist_backend="gloo")
    torch.manual_seed(0)
    b, s, d = 2, 8 * world, 16
    q, k, v = (torch.randn(b, s, heads, d) for _ in range(3))
    want = _attn(q, k, v)
    sl = slice(rank * s // world, (rank + 1) * s // world)
    ql, kl, vl = (t[:, sl].clone().requires_grad_(True) for t in (q, k, v))
    attn = DistributedAttention(_attn, dist.get_world_group())
    out = attn(ql, kl, vl)
    out.sum().backward()
    qf = q.clone().requires_grad_(True)
    _attn(qf, k, v).sum().backward()
    return {"err": float((out - want[:, sl]).abs().max()), "gerr": float((ql.grad - qf.grad[:, sl]).abs().max())}


def fpdt_worker(rank, world, chunks, offload):
    import hcache_deepspeed_b200 as ds
    from hcache_deepspeed_b200 import comm as dist
    from hcache_deepspeed_b200.sequence.fpdt_layer import FPDT_Attention
    ds.init_distributed(dist_backend="gloo")
    torch.manual_seed(0)
    b, s, h, d = 1, 8 * world * chunks, 4, 8
    q, k, v = (torch.randn(b, s, h, d) for _ in range(3))
    want = _attn(q, k, v)
    sl = slice(rank * s // world, (rank + 1) * s // world)
    out = FPDT_Attention(dist.get_world_group(), num_chunks=chunks, offloading=offload)(q[:, sl], k[:, sl], v[:, sl])
    return {"err": float((out - want[:, sl]).abs().max())}


class _Expert(torch.nn.Module):
    def __init__(self, d):
        super().__init__()
        self.a, self.b = torch.nn.Linear(d, 2 * d), torch.nn.Linear(2 * d, d)

    def forward(self, x):
        return self.b(torch.relu(self.a(x)))


def moe_worker(rank, world, ep_size, k):
    """Expert-parallel MoE == the same layer with all experts local (ep=1) given identical weights and tokens."""
    import hcache_deepspeed_b200 as ds
    from hcache_deepspeed_b200 import comm as dist
    from hcache_deepspeed_b200.moe.layer import MoE
    ds.init_distributed(dist_backend="gloo")
    d, E, tokens = 16, 4, 32
    torch.manual_seed(3)
    ref = MoE(d, _Expert(d), num_experts=E, ep_size=1, k=k, capacity_factor=4.0, min_capacity=4, use_rts=False,
              top2_2nd_expert_sampling=False)
    # give every expert distinct weights
    for i, e in enumerate(ref.deepspeed_moe.experts.deepspeed_experts):
        torch.manual_seed(100 + i)
        for p in e.parameters():
            p.data.normal_(0, 0.2)
    torch.manual_seed(3)
    par = MoE(d, _Expert(d), num_experts=E, ep_size=ep_size, k=k, capacity_factor=4.0, min_capacity=4, use_rts=False,
              top2_2nd_expert_sampling=False)
    par.set_deepspeed_parallelism()
    par.deepspeed_moe.gate.wg.weight.data.copy_(ref.deepspeed_moe.gate.wg.weight.data)
    ep_rank = dist.get_rank() % ep_size
    nl = E // ep_size
    for li, e in enumerate(par.deepspeed_moe.experts.deepspeed_experts):
        src = ref.deepspeed_moe.experts.deepspeed_experts[ep_rank * nl + li]
        e.load_state_dict(src.state_dict())
    # every rank of an EP group routes its own tokens; compare each rank's output with the all-local layer on its tokens
    torch.manual_seed(50 + rank)
    x = torch.randn(tokens, d, requires_grad=True)
    y, l_aux, counts = par(x)
    y.sum().backward()
    x2 = x.detach().clone().requires_grad_(True)
    y2, l_aux2, counts2 = ref(x2)
    y2.sum().backward()
    return {"err": float((y - y2).abs().max()), "gerr": float((x.grad - x2.grad).abs().max()),
            "laux": float(l_aux), "laux_ref": float(l_aux2), "counts": counts.tolist(),
            "is_moe_param": all(getattr(p, "allreduce", True) is False for p in par.deepspeed_moe.experts.parameters())}


def onebit_worker(rank, world):
    """Compressed all-reduce approximates the mean and its error feedback keeps the long-run average unbiased."""
    import hcache_deepspeed_b200 as ds
    from hcache_deepspeed_b200.runtime.comm.compressed import CompressedBackend
    ds.init_distributed(dist_backend="gloo")
    b = CompressedBackend()
    torch.manual_seed(rank)
    n = 1000
    we, se = torch.zeros(b.padded_size(n)), torch.zeros(b.padded_size(n) // world)
    acc_c, acc_t = torch.zeros(n), torch.zeros(n)
    for it in range(60):
        g = torch.randn(n) + 0.5
        exact = g.clone()
        ds.comm.all_reduce(exact)
        exact /= world
        m = g.clone()
        b.compressed_allreduce(m, we, se)
        acc_c += m
        acc_t += exact
    return {"rel": float((acc_c - acc_t).norm() / acc_t.norm()), "same": float(m.sum())}


def onebit_adam_worker(rank, world):
    import hcache_deepspeed_b200 as ds
    ds.init_distributed(dist_backend="gloo")
    from tests.runtime_helper import TinyNet, make_batches
    torch.manual_seed(1234)
    cfg = {"train_micro_batch_size_per_gpu": 4, "optimizer": {"type": "OneBitAdam", "params": {"lr": 5e-3, "freeze_step": 6}},
           "steps_per_print": 1000}
    engine, _, _, _ = ds.initialize(model=TinyNet(32), config=cfg)
    losses = []
    x, y = make_batches(1, 4 * world, 32)[0]      # one fixed batch: the loss must go down
    for _ in range(14):
        lo = rank * 4
        loss = engine(x[lo:lo + 4], y[lo:lo + 4])
        engine.backward(loss)
        engine.step()
        losses.append(float(loss.detach()))
    return {"losses": losses, "frozen": engine.basic_optimizer.adam_freeze_key, "allreduce_off": not engine.enable_backward_allreduce,
            "w": engine.module.head.weight.detach().clone()}


def shm_allreduce_worker(rank, world):
    """Shared-memory all-reduce (C++ engine) against gloo, through comm.inference_all_reduce."""
    import hcache_deepspeed_b200 as ds
    from hcache_deepspeed_b200 import comm as dist
    from hcache_deepspeed_b200.comm import shm
    ds.init_distributed(dist_backend="gloo")
    g = shm.attach(slot_bytes=4096)                  # small slots: large tensors exercise the chunked path
    out = {}
    for dtype, n in ((torch.float32, 7), (torch.float32, 5000), (torch.bfloat16, 3001)):
        torch.manual_seed(100 + rank)
        x = torch.randn(n).to(dtype)
        want = x.clone()
        torch.distributed.all_reduce(want)           # gloo
        got = x.clone()
        assert shm.usable(got)
        dist.inference_all_reduce(got)
        out[f"{dtype}_{n}"] = float((got.float() - want.float()).abs().max())
    for i in range(50):                              # back-to-back rounds: slot reuse must not race
        t = torch.full((1500, ), float(rank + i))
        dist.inference_all_reduce(t)
        assert float(t[0]) == sum(r + i for r in range(world)) and float(t[-1]) == float(t[0])
    y = torch.ones(4, dtype=torch.int32)             # unsupported dtype falls back to torch.distributed
    dist.inference_all_reduce(y)
    out["int_fallback"] = int(y[0])
    out["name"] = g.name
    shm.detach()
    return out


def ccl_backend_worker(rank, world):
    """``init_distributed(dist_backend="ccl")``: every covered collective runs through the C++ shared-memory engine and
    agrees with gloo; uncovered calls fall back; sub-groups get their own segment."""
    import os
    os.environ["HDS_SHM_SIMD"] = "off" if rank == -1 else os.environ.get("HDS_SHM_SIMD", "on")
    import hcache_deepspeed_b200 as ds
    from hcache_deepspeed_b200 import comm as dist
    from hcache_deepspeed_b200.comm import comm as C, shm
    from hcache_deepspeed_b200.comm.ccl import CCLBackend
    from hcache_deepspeed_b200.comm.reduce_op import ReduceOp
    ds.init_distributed(dist_backend="ccl")
    be = C.cdb
    assert isinstance(be, CCLBackend) and be.single_node and be.is_initialized()
    eng = be.native_engine()
    out = {"avx512": eng.info()["avx512"]}
    gloo = torch.distributed.new_group(list(range(world)), backend="gloo")     # not attached: the oracle

    def oracle_all_reduce(x, op):
        y = x.clone().float() if x.dtype in (torch.float16, torch.bfloat16) else x.clone()
        torch.distributed.all_reduce(y, op=op, group=gloo)
        return y

    errs = {}
    for dtype in (torch.float32, torch.bfloat16, torch.float16, torch.int32, torch.int64):
        for n in (1, 37, 4096 + 5, 300_001):      # one exchange, a partial vector, shard-per-rank, several rounds
            torch.manual_seed(7 * rank + n)
            x = (torch.randn(n) * 4).to(dtype) if dtype.is_floating_point else torch.randint(-50, 50, (n, ), dtype=dtype)
            for op, top in ((ReduceOp.SUM, torch.distributed.ReduceOp.SUM), (ReduceOp.MAX, torch.distributed.ReduceOp.MAX),
                            (ReduceOp.MIN, torch.distributed.ReduceOp.MIN)):
                got = x.clone()
                be.all_reduce(got, op=op)
                want = oracle_all_reduce(x, top)
                if dtype.is_floating_point:
                    want = want.to(dtype)        # one rounding of the fp32 result
                    e = float((got.float() - want.float()).abs().max())
                    tol = {torch.float32: 1e-5, torch.bfloat16: 0.26, torch.float16: 0.04}[dtype] if op == ReduceOp.SUM else 0.0
                    assert e <= tol, (dtype, n, op, e)
                else:
                    assert torch.equal(got, want), (dtype, n, op)
    info = eng.info()
    assert info["allreduce_one_exchange"] > 0 and info["allreduce_shard_per_rank"] > 0
    # results are bit-identical on every rank (fixed summation order / one owner per shard)
    x = torch.randn(70_001) * (rank + 1)
    be.all_reduce(x)
    ref = [torch.empty_like(x) for _ in range(world)]
    torch.distributed.all_gather(ref, x, group=gloo)
    assert all(torch.equal(r, ref[0]) for r in ref)

    # broadcast / all-gather / reduce-scatter
    for n in (5, 250_003):
        b = torch.arange(n, dtype=torch.float64) * (1 if rank == world - 1 else -1)
        be.broadcast(b, src=world - 1)
        assert torch.equal(b, torch.arange(n, dtype=torch.float64))
        mine = torch.full((n, ), rank, dtype=torch.int16)
        flat = torch.empty(world * n, dtype=torch.int16)
        be.all_gather_into_tensor(flat, mine)
        assert torch.equal(flat.view(world, n), torch.arange(world, dtype=torch.int16)[:, None].expand(world, n))
        lst = [torch.empty(n, dtype=torch.int16) for _ in range(world)]
        be.all_gather(lst, mine)
        assert all(int(t[0]) == r and int(t[-1]) == r for r, t in enumerate(lst))
        torch.manual_seed(rank + n)
        inp = torch.randn(world * n)
        got = torch.empty(n)
        be.reduce_scatter_tensor(got, inp)
        want = inp.clone()
        torch.distributed.all_reduce(want, group=gloo)
        torch.testing.assert_close(got, want[rank * n:(rank + 1) * n], rtol=1e-5, atol=1e-5)
        hb = (torch.randn(world * n) * 3).to(torch.bfloat16)
        gb = torch.empty(n, dtype=torch.bfloat16)
        be.reduce_scatter_tensor(gb, hb, op=ReduceOp.MAX)
        wb = hb.float()
        torch.distributed.all_reduce(wb, op=torch.distributed.ReduceOp.MAX, group=gloo)
        assert torch.equal(gb.float(), wb[rank * n:(rank + 1) * n])
    be.barrier()
    for i in range(200):                               # back-to-back small calls: the two banks must not race
        t = torch.full((33, ), float(rank + i))
        dist.all_reduce(t)                             # the functional API is routed too
        assert float(t[0]) == sum(r + i for r in range(world))
    # not covered -> torch.distributed, and the counters say so
    z = torch.ones(6, dtype=torch.float64)
    be.all_reduce(z)
    nc = torch.ones(4, 6)[:, ::2]
    be.all_reduce(nc)
    prod = torch.full((3, ), 2.0)
    be.all_reduce(prod, op=ReduceOp.PRODUCT)
    assert float(z[0]) == world and float(nc[0, 0]) == world and float(prod[0]) == 2.0**world
    out["calls"] = dict(be.calls)
    # a sub-group gets its own segment
    if world >= 3:
        sub = be.new_group([0, 2])
        if rank in (0, 2):
            assert be.native_engine(sub) is not None
            t = torch.full((10, ), float(rank))
            be.all_reduce(t, group=sub)
            assert float(t[0]) == 2.0
            b = torch.full((4, ), float(rank))
            be.broadcast(b, src=2, group=sub)          # global rank 2 = group rank 1
            assert float(b[0]) == 2.0
    names = [eng.name]
    be.close()
    assert shm.attached() is None and not shm.route_all_collectives
    out["name"] = names[0]
    return out


def ulysses_engine_worker(rank, world):
    """Ulysses sequence parallelism THROUGH the training engine (``"sequence_parallel_size": world``): every rank feeds its
    slice of the sequence, ``DistributedAttention`` all-to-alls heads <-> sequence, gradients are reduced over the
    sequence-data-parallel group; losses equal full-sequence training on one process."""
    import torch, traceback, math
    import hcache_deepspeed_b200 as ds
    from hcache_deepspeed_b200.sequence.layer import DistributedAttention
    from hcache_deepspeed_b200.utils import groups
    ds.init_distributed(dist_backend="gloo")
    H, NH, S, V = 32, 4, 16, 50
    def core(q, k, v):   # [b, s, h, d]
        q, k, v = (t.transpose(1, 2) for t in (q, k, v))
        return torch.nn.functional.scaled_dot_product_attention(q, k, v, is_causal=True).transpose(1, 2)
    class Net(torch.nn.Module):
        def __init__(self, sp_group=None):
            super().__init__()
            self.emb = torch.nn.Embedding(V, H); self.qkv = torch.nn.Linear(H, 3 * H); self.o = torch.nn.Linear(H, V)
            self.attn = DistributedAttention(core, sp_group, scatter_idx=2, gather_idx=1) if sp_group is not None else core
        def forward(self, ids, labels):
            x = self.emb(ids); b, s, _ = x.shape
            q, k, v = self.qkv(x).view(b, s, 3, NH, H // NH).unbind(2)
            logits = self.o(self.attn(q, k, v).reshape(b, s, H))
            return torch.nn.functional.cross_entropy(logits.reshape(-1, V), labels.reshape(-1), reduction="sum")
    if True:
        torch.manual_seed(0)
        ref = Net()
        g = torch.Generator().manual_seed(3)
        ids = torch.randint(0, V, (2, S), generator=g); labels = torch.randint(0, V, (2, S), generator=g)
        import copy
        sd = copy.deepcopy(ref.state_dict())
        opt = torch.optim.SGD(ref.parameters(), lr=0.05)
        want = []
        for _ in range(4):
            l = ref(ids, labels) / (2 * S); l.backward(); opt.step(); opt.zero_grad(); want.append(round(float(l.detach()), 5))
        cfg = {"train_micro_batch_size_per_gpu": 2, "optimizer": {"type": "SGD", "params": {"lr": 0.05}}, "steps_per_print": 1000,
               "sequence_parallel_size": world, "zero_optimization": {"stage": 1}}
        net = Net(sp_group=None)
        net.load_state_dict(sd)
        engine, _, _, _ = ds.initialize(model=net, config=cfg)
        net.attn = DistributedAttention(core, groups._get_sequence_parallel_group(), scatter_idx=2, gather_idx=1)
        lo, hi = rank * S // world, (rank + 1) * S // world
        got = []
        for _ in range(4):
            l = engine(ids[:, lo:hi], labels[:, lo:hi]) / (2 * S) * world    # each rank holds 1/world of the tokens; dp averaging divides by world
            engine.backward(l); engine.step()
            t = l.detach().clone() / world; torch.distributed.all_reduce(t); got.append(round(float(t), 5))
        return {"want": want, "got": got, "dp": engine.dp_world_size, "sp": engine.sequence_parallel_size}


def comm_off_worker(rank, world):
    """The switches that turn families of collectives into no-ops on a backend object (ref: comm/torch.py:60-110)."""
    import importlib
    import torch
    import hcache_deepspeed_b200 as ds
    ds.init_distributed(dist_backend="gloo")
    T = importlib.import_module("hcache_deepspeed_b200.comm.torch")
    b = T.TorchBackend(backend="gloo")
    t = torch.full((2, ), float(rank + 1))
    T.all_reduce_comm_off(True)
    off = b.all_reduce(t)
    skipped = t.tolist()
    T.all_reduce_comm_off(False)
    b.all_reduce(t)
    T.backward_comm_off(True)
    flags = (T.DS_COMM_ALL_GATHER_OFF, T.DS_COMM_REDUCE_SCATTER_OFF)
    gathered = b.all_gather_into_tensor(torch.zeros(2 * world), torch.ones(2))
    T.backward_comm_off(False)
    return {"noop": type(off).__name__, "wait": off.wait(), "skipped": skipped, "reduced": t.tolist(), "flags": flags,
            "gather_noop": type(gathered).__name__, "probes": (T.has_all_reduce_coalesced(), T.has_coalescing_manager())}


def mp_compress_worker(rank, world):
    """Column- then row-parallel compressible linears over a model-parallel group equal a dense two-layer product,
    forward and input gradient (ref: compression/basic_layer.py:680-840)."""
    import torch
    import hcache_deepspeed_b200 as ds
    from hcache_deepspeed_b200.compression import basic_layer as BL
    ds.init_distributed(dist_backend="gloo")
    group = ds.comm.new_group(list(range(world)))

    class Mpu:
        def get_model_parallel_group(self):
            return group

        def get_model_parallel_world_size(self):
            return world

        def get_model_parallel_rank(self):
            return rank

    torch.manual_seed(0)
    w1, b1, w2, b2 = torch.randn(8, 6), torch.randn(8), torch.randn(4, 8), torch.randn(4)
    col = BL.ColumnParallelLinear_Compress(Mpu(), 6, 8, gather_output=False)
    row = BL.RowParallelLinear_Compress(Mpu(), 8, 4, input_is_parallel=True)
    with torch.no_grad():
        col.weight.copy_(w1.chunk(world, 0)[rank]); col.bias.copy_(b1.chunk(world, 0)[rank])
        row.weight.copy_(w2.chunk(world, 1)[rank]); row.bias.copy_(b2)
    x = torch.randn(3, 6, requires_grad=True)
    h, _ = col(x)
    y, _ = row(torch.tanh(h))
    y.pow(2).sum().backward()
    xr = x.detach().clone().requires_grad_(True)
    yr = torch.tanh(xr @ w1.t() + b1) @ w2.t() + b2
    yr.pow(2).sum().backward()
    full = BL.ColumnParallelLinear_Compress(Mpu(), 6, 8, gather_output=True, skip_bias_add=True)
    out, bias = full(x.detach())
    sc = BL.scatter_to_model_parallel_region(torch.arange(8.).reshape(1, 8))
    return {"fwd": float((y - yr).abs().max()), "grad": float((x.grad - xr.grad).abs().max()),
            "gathered_shape": tuple(out.shape), "bias_back": bias is not None and tuple(bias.shape) == (8 // world, ),
            "scatter": sc.tolist(), "gather": BL.gather_from_model_parallel_region(sc).tolist()}
