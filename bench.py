#!/usr/bin/env python
"""Headline benchmark: restored multi-turn sessions on the ragged engine (BASELINE.json config 2..5).

One STEP = one serving round over B synthetic multi-turn sessions:
  for each session: restore_kv(4K-token context from the pinned host hidden-state cache)  -> H2D + K/V projection
                    put(new-turn prompt)                                                   -> first token (TTFT sample)
  then D decode iterations over the B-session batch (greedy), then every session is flushed.
Metric `value` = decode tokens/s (whole job: tensor parallel over --gpus ranks, so one token stream);
`ttft_p50_ms` = p50 over all (step, session) samples of restore+prefill latency. Everything is measured through the
public engine API (`restore_kv` / `put`) with inputs coming from pinned host memory and results read back to the
host every forward; `value` times the decode phase with CUDA events, `e2e.value` is the same phase by host clock
around the API calls (includes host scheduling + H2D/D2H). Random-init weights, synthetic tokens.

N > 1: launched by torchrun (one rank per GPU, NCCL); TP = N.
`--impl reference` runs the unmodified reference from baseline/_ref through deepspeed.init_inference.
"""
import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)


def parse():
    p = argparse.ArgumentParser()
    p.add_argument("--gpus", type=int, default=1)
    p.add_argument("--steps", type=int, default=5)
    p.add_argument("--warmup", type=int, default=3)
    p.add_argument("--impl", default="ours", choices=["ours", "reference"])
    p.add_argument("--model", default="llama-2-7b")
    p.add_argument("--ctx", type=int, default=4096, help="restored context tokens per session")
    p.add_argument("--new-tokens", type=int, default=64, help="tokens of the new turn's prompt")
    p.add_argument("--decode-steps", type=int, default=32)
    p.add_argument("--sessions", type=int, default=8, help="concurrent sessions PER GPU (weak scaling: global batch = sessions x gpus)")
    p.add_argument("--parallelism", default="tp", choices=["tp", "dp"],
                   help="N > 1: 'tp' (default, the headline metric: decode tokens/s at TP = N) = one engine tensor-parallel over "
                        "all GPUs (flag-protocol NVLink collectives, PCIe-sharded restore); 'dp' = one full replica per GPU "
                        "serving its own sessions (no collective on the path; secondary)")
    p.add_argument("--source", default="dram", choices=["dram", "nvme"],
                   help="where a session's hidden states live before the restore: pinned host DRAM (headline) or a file read "
                        "back through the O_DIRECT aio engine inside the timed region (NVMe tier)")
    p.add_argument("--nvme-dir", default="/tmp/hds_bench_nvme")
    p.add_argument("--quant", default="none", choices=["none", "wf8af8"],
                   help="none (default, the bf16 headline configuration) or wf8af8: E4M3 weights + per-token fp8 activations; "
                        "reported with dtype fp8-w8a8 and NOT comparable with the bf16 baseline")
    p.add_argument("--force-kv-layers", type=int, default=0,
                   help="restore the LAST N layers from offloaded KV pages and the rest from hidden states (a fixed layer split, "
                        "e.g. the OPT-30B TP=4 NVMe configuration), instead of what the planner would choose")
    p.add_argument("--plan", default="hidden", choices=["auto", "hidden", "planner", "adaptive"],
                   help="hidden: every layer from hidden states (fastest measured on 1xB200: 20.4 ms vs 21.6 ms); auto: planner mix with a recomputed prefix")
    return p.parse_args()


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled during the timed region (B200_PROFILING.md recipe)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index: int):
        self.idx, self.proc, self.lines = gpu_index, None, []

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "200",
                                          "-i", str(self.idx)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        sm, mx, reasons = [], [], set()
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1]))
                mx.append(float(f[2]))
            except ValueError:
                continue
            for name, val in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[5:9]):
                if val.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


def run_ours(a):
    import torch
    import torch.distributed as dist
    from hcache_deepspeed_b200.inference.v2 import RaggedInferenceEngineConfig, build_engine_from_arch
    from hcache_deepspeed_b200.models.arch import named_arch
    from hcache_deepspeed_b200.ops import kernels, native

    rank = int(os.environ.get("RANK", 0))
    world = int(os.environ.get("WORLD_SIZE", 1))
    local = int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    from hcache_deepspeed_b200.utils.numa import bind_to_gpu_numa
    numa_cpus = bind_to_gpu_numa(local)   # pinned hidden-state buffers land on the GPU's own NUMA node
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)
    spec = named_arch(a.model)
    tp = world if a.parallelism == "tp" else 1          # ranks that share one engine
    replicas = world // tp                              # independent engines (data parallel)
    B, ctx, n_new, D = a.sessions * tp, a.ctx, a.new_tokens, a.decode_steps   # sessions handled by THIS engine
    max_ctx = ctx + n_new + D + 64
    cfg = RaggedInferenceEngineConfig(
        tensor_parallel={"tp_size": tp},
        state_manager={"max_tracked_sequences": max(2 * B, 16), "max_ragged_batch_size": max(ctx, B * n_new) + 64,
                       "max_ragged_sequence_count": max(B, 8), "max_context": max_ctx,
                       "memory_config": {"mode": "allocate",
                                         "size": int((B + 1) * (max_ctx + 128) * spec.kv_bytes_per_token_layer() * spec.n_layers / tp)}},
        hcache={"enabled": True, "capture": "off", "shard_capture": tp > 1},
        quantization={"quantization_mode": None if a.quant == "none" else a.quant})
    if tp > 1:
        os.environ.setdefault("HDS_TP_STRICT", "1")   # the headline run must not silently fall back to NCCL collectives
    eng = build_engine_from_arch(spec, cfg, seed=1234, dtype=torch.bfloat16, device=dev)
    model = eng.model()
    assert native.cuda_lib_loaded(), "sm_100a kernel library was not loaded"
    if tp > 1 and os.environ.get("HDS_TP_STRICT") == "1":
        assert getattr(model.comm, "fused", False), "tensor-parallel run without the fused flag-protocol collectives"

    # synthetic host-DRAM hidden-state cache: one pinned [L, ctx, H] tensor (every session H2Ds its own copy). Tensor
    # parallel: every rank keeps (and transfers) only its 1/tp row slice of the hidden states; the restore kernels gather
    # the slices from peer HBM over NVLink — so only that slice is ever allocated (a 70B / 16K-token cache is 21 GB)
    g = torch.Generator().manual_seed(7)
    sharded = tp > 1 and bool(getattr(model.comm, "fused", False))
    lo, hi = eng.latent_row_range(ctx) if sharded else (0, ctx)
    latents = torch.empty((spec.n_layers, hi - lo, spec.hidden), dtype=torch.bfloat16, pin_memory=True)
    for l in range(spec.n_layers):   # layer by layer: the fp32 staging of a whole 70B cache would not fit comfortably
        latents[l].copy_((torch.randn((hi - lo, spec.hidden), generator=g) * 0.5).to(torch.bfloat16))
    lat_slice = latents
    ctx_tokens = torch.randint(0, spec.vocab, (ctx, ), generator=g).pin_memory()
    new_tokens = [torch.randint(0, spec.vocab, (n_new, ), generator=g).pin_memory() for _ in range(B)]
    from hcache_deepspeed_b200.hcache import LinkModel, plan_restore
    # measured input of the planner: how long the new turn's prefill occupies the GPU (it shares the restore window)
    eng.restore_kv([10 ** 9], [ctx_tokens], [latents], do_checks=False, defer=False, sharded=sharded)
    torch.cuda.synchronize()
    p0, p1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    eng.put([10 ** 9], [new_tokens[0]], do_checks=False)
    eng.flush(10 ** 9)
    eng.restore_kv([10 ** 9], [ctx_tokens], [latents], do_checks=False, defer=False, sharded=sharded)
    torch.cuda.synchronize()
    p0.record()
    eng.put([10 ** 9], [new_tokens[0]], do_checks=False)
    p1.record()
    torch.cuda.synchronize()
    eng.flush(10 ** 9)
    prefill_s = p0.elapsed_time(p1) / 1e3
    # adaptive: measure the host link as it is DURING the run (every rank copies at once, as in the timed region); when
    # the ranks contend for host memory bandwidth the planner trades transfers for a recomputed prefix of layers
    link = LinkModel(tp_size=tp)
    allow_recompute = a.plan == "auto"
    if a.plan == "adaptive" and tp == 1:
        from hcache_deepspeed_b200.hcache import measure_link
        link = measure_link(dev, tp_size=tp)
        allow_recompute = link.io_bytes_per_s < 0.8 * LinkModel().io_bytes_per_s     # only when the link is contended
    # planner: hidden states vs offloaded KV pages per layer by bytes moved (GQA models: KV is smaller than the hidden state)
    plan = plan_restore(spec, ctx, link, allow_kv=a.plan == "planner", allow_recompute=allow_recompute,
                        extra_compute_s=prefill_s)
    hidden_layers, recompute_layers, kv_layers = plan.layers_of("hidden"), plan.layers_of("none"), plan.layers_of("kv")
    if a.force_kv_layers > 0:
        n_kv_forced = min(a.force_kv_layers, spec.n_layers)
        hidden_layers, recompute_layers = list(range(spec.n_layers - n_kv_forced)), []
        kv_layers = list(range(spec.n_layers - n_kv_forced, spec.n_layers))
    lat_hidden = latents[hidden_layers[0]:hidden_layers[-1] + 1] if hidden_layers else None  # a contiguous run of layers
    kv_pages = None
    if kv_layers:
        bs_kv = model._state_manager.kv_block_size
        n_blk = (ctx + bs_kv - 1) // bs_kv
        # offloaded KV pages of THIS rank's KV heads (tensor parallel: every rank reloads its own shard, no collective)
        kv_pages = torch.empty((len(kv_layers), n_blk, 2, model.n_kv, bs_kv, spec.head_dim), dtype=torch.bfloat16,
                               pin_memory=True)
        for j in range(len(kv_layers)):
            kv_pages[j].copy_((torch.randn(kv_pages.shape[1:], generator=g) * 0.3).to(torch.bfloat16))
    interleave = plan.interleave
    nvme = None
    if a.source == "nvme" and lat_hidden is not None:
        # NVMe tier: the session's hidden states sit in a file; every restore reads them back (one request per layer, all
        # aio threads in flight, O_DIRECT so the page cache cannot serve them) into the pinned staging tensor first
        from hcache_deepspeed_b200.io.aio import AsyncIOHandle
        os.makedirs(a.nvme_dir, exist_ok=True)
        nvme = {"h": AsyncIOHandle(block_size=8 << 20, queue_depth=32, intra_op_parallelism=8, use_direct=True),
                "path": os.path.join(a.nvme_dir, f"hds_bench_rank{rank}.bin"),
                "layer_bytes": lat_hidden[0].numel() * 2}
        for l in range(lat_hidden.shape[0]):
            nvme["h"].sync_pwrite(lat_hidden[l], nvme["path"], l * nvme["layer_bytes"])
    rows_moved = hi - lo
    h2d_step = B * (len(hidden_layers) * rows_moved * spec.hidden * 2 + ctx * 4 + n_new * 4) + B * D * 4 + \
        B * (kv_pages.numel() * 2 if kv_pages is not None else 0)
    d2h_step = B * (D + 1) * 4

    ev = lambda: torch.cuda.Event(enable_timing=True)  # noqa: E731
    uid_base = [0]

    def one_step(record):
        uids = [uid_base[0] + i for i in range(B)]
        uid_base[0] += B
        first = []
        for i, uid in enumerate(uids):
            t0, e0, e1 = time.perf_counter(), ev(), ev()
            e0.record()
            layer_ready = None
            if nvme is not None:
                # every layer's read is queued now (O_DIRECT, kernel AIO: the page cache cannot serve it); the restore waits
                # for layer l only when it is about to copy layer l to the device, so disk, PCIe and the projections overlap
                reqs = [nvme["h"].async_pread(lat_hidden[l], nvme["path"], l * nvme["layer_bytes"])
                        for l in range(lat_hidden.shape[0])]
                layer_ready = lambda l, reqs=reqs: nvme["h"].wait(reqs[l])  # noqa: E731
            if sharded and not kv_layers and not recompute_layers:
                eng.restore_kv([uid], [ctx_tokens], [lat_hidden], do_checks=False, sharded=True, layer_ready=layer_ready)
            else:
                eng.restore_session(uid, ctx_tokens, hidden=lat_hidden, hidden_layers=hidden_layers, kv=kv_pages,
                                    kv_layers=kv_layers, kv_layer_major=True, recompute_layers=recompute_layers,
                                    interleave=interleave, sharded=sharded, layer_ready=layer_ready)
            logits, _ = eng.put([uid], [new_tokens[i]], do_checks=False)
            tok = kernels.argmax(logits)
            tok_host = tok.cpu()  # D2H of the first token: the TTFT point
            e1.record()
            e1.synchronize()
            first.append(tok_host)
            if record is not None:
                record["ttft_dev_ms"].append(e0.elapsed_time(e1))
                record["ttft_host_ms"].append((time.perf_counter() - t0) * 1e3)
        cur = [t.clone() for t in first]
        d0, d1, th0 = ev(), ev(), time.perf_counter()
        d0.record()
        for _ in range(D):
            logits, _ = eng.put(uids, cur, do_checks=False)
            nxt = kernels.argmax(logits).cpu()  # D2H of the sampled ids every iteration
            cur = [nxt[i:i + 1] for i in range(B)]
        d1.record()
        d1.synchronize()
        if record is not None:
            record["decode_dev_ms"].append(d0.elapsed_time(d1))
            record["decode_host_ms"].append((time.perf_counter() - th0) * 1e3)
        for uid in uids:
            eng.flush(uid)

    for _ in range(a.warmup):
        one_step(None)
    rec = {"ttft_dev_ms": [], "ttft_host_ms": [], "decode_dev_ms": [], "decode_host_ms": []}
    sampler = ClockSampler(local)
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    if rank == 0:
        sampler.start()
    launches0 = kernels.launch_count
    s_all, e_all = ev(), ev()
    s_all.record()
    for _ in range(a.steps):
        one_step(rec)
    e_all.record()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    clocks = sampler.stop() if rank == 0 else None
    if nvme is not None:
        try:
            os.remove(nvme["path"])
        except OSError:
            pass
    total_ms = s_all.elapsed_time(e_all)
    stats = torch.tensor([total_ms, sum(rec["decode_dev_ms"]), sum(rec["decode_host_ms"]),
                          statistics.median(rec["ttft_dev_ms"]), statistics.median(rec["ttft_host_ms"])], device=dev)
    launches = torch.tensor([float(kernels.launch_count - launches0)], device=dev)
    if world > 1:
        dist.all_reduce(stats, op=dist.ReduceOp.MAX)      # every time is the max over ranks
        dist.all_reduce(launches, op=dist.ReduceOp.SUM)
    total_ms, dec_dev_ms, dec_host_ms, ttft_dev, ttft_host = stats.tolist()
    if rank == 0:
        dec_tokens = a.steps * B * D * replicas           # whole job: every replica decodes its own sessions
        base_path = os.path.join(ROOT, "BASELINE.json")
        published = json.load(open(base_path)).get("published", {}) if os.path.exists(base_path) else {}
        ref_val = published.get("decode_tokens_per_sec")
        value = dec_tokens / (dec_dev_ms / 1e3)
        out = {
            "metric": "decode tokens/sec over restored multi-turn sessions (p50 restored-context TTFT in ttft_p50_ms)",
            "value": round(value, 2), "unit": "tokens/s", "n_gpus": world, "steps": a.steps, "warmup": a.warmup,
            "ms_per_step": round(total_ms / a.steps, 3), "higher_is_better": True, "scaling": "weak",
            "vs_baseline": (value / ref_val) if ref_val else None, "dtype": "bf16" if a.quant == "none" else "fp8-w8a8",
            "data": "synthetic tokens and hidden states, random-init weights", "impl": "ours",
            "ttft_p50_ms": round(ttft_dev, 3),
            "config": {"model": a.model, "global_batch": B * replicas, "seq_len": ctx + n_new + D,
                       "parallelism": f"tp{tp}" if replicas == 1 else f"dp{replicas}",
                       "restored_ctx": ctx, "new_turn_tokens": n_new, "decode_steps": D, "sessions_per_gpu": a.sessions},
            "details": {"numa_bound_cpus": len(numa_cpus), "hidden_state_source": a.source if nvme is not None or a.source == "dram" else "dram",
                        "restore_plan": {"recompute_prefix_layers": len(recompute_layers),
                                         "hidden_state_layers": len(hidden_layers),
                                         "kv_offload_layers": len(kv_layers), "tp_sharded_restore": sharded,
                                         "host_link_GBps": round(link.io_bytes_per_s / 1e9, 1)},
                        "tp_collectives": ("flag protocol: GEMM tail signal + " + ("multimem.ld_reduce/st" if getattr(model.comm, "use_nvls", False)
                                                                                 else "peer loads/stores") + " inside the norm kernel")
                        if bool(getattr(model.comm, "fused", False)) else ("nccl" if tp > 1 else "none"),
                        "decode_ms_per_iteration": round(dec_dev_ms / (a.steps * D), 4),
                        "l2_policy": f"working set per step (weights {model.w.n_bytes() / 2**30:.1f} GiB + KV) exceeds the 126 MB L2"},
            "e2e": {"value": round(dec_tokens / (dec_host_ms / 1e3), 2), "unit": "tokens/s", "ttft_p50_ms": round(ttft_host, 3),
                    "h2d_bytes_per_step": h2d_step * replicas, "d2h_bytes_per_step": d2h_step * replicas},
            "gpu_launches": int(launches.item()),
            "clocks": clocks,
        }
        print(json.dumps(out), flush=True)
    if world > 1:
        # leave without tearing down NCCL / symmetric-memory state rank by rank: destroy_process_group() was seen
        # to block at exit with peer-mapped buffers still alive
        dist.barrier()
        torch.cuda.synchronize()
        sys.stdout.flush()
        os._exit(0)


def run_reference(a):
    from baseline.reference_arm import run_reference_arm
    run_reference_arm(a)


if __name__ == "__main__":
    args = parse()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)
