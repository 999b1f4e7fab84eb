"""Reference arm of bench.py: the UNMODIFIED reference (baseline/_ref, installed with pip --no-deps from
/root/reference) driven through its own public API, `deepspeed.init_inference`, on the same workload.

What the reference can and cannot do on this box: its ragged v2 engine — the only place HCache
(`put` returning latents, `restore_kv`) exists — needs the external `deepspeed-kernels` wheel and refuses sm_100
(`blocked_flash.cpp:31-33`), so it cannot run. The stock path that does run is v1 `init_inference` over the
HuggingFace model (no kernel injection: the injected path's op would have to be JIT-built on the box). v1 has no
hidden-state restore, so a returning session's context can only be re-established by recomputing the prompt:
the TTFT this arm reports is recompute TTFT (ctx + new-turn tokens prefill). Decode is the same batched greedy
loop with the HF KV cache. None of this repository's kernels, models or engine are imported here.
"""
import json
import os
import statistics
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = os.path.join(ROOT, "baseline", "_ref")

LLAMA = {
    "llama-2-7b": dict(hidden_size=4096, intermediate_size=11008, num_hidden_layers=32, num_attention_heads=32, num_key_value_heads=32),
    "llama-2-13b": dict(hidden_size=5120, intermediate_size=13824, num_hidden_layers=40, num_attention_heads=40, num_key_value_heads=40),
    "llama-2-70b": dict(hidden_size=8192, intermediate_size=28672, num_hidden_layers=80, num_attention_heads=64, num_key_value_heads=8),
}


def unavailable(why: str) -> None:
    if int(os.environ.get("RANK", 0)) == 0:
        print(json.dumps({"impl": "reference", "unavailable": why}), flush=True)
    sys.exit(0)


def run_reference_arm(a) -> None:
    if not os.path.isdir(os.path.join(REF, "deepspeed")):
        unavailable("baseline/_ref is missing (pip install --no-deps --target baseline/_ref of /root/reference not done)")
    if a.model not in LLAMA:
        unavailable(f"reference arm only builds Llama-2 configs, not {a.model}")
    sys.path.insert(0, REF)
    os.environ.setdefault("DS_SKIP_CUDA_CHECK", "1")
    os.environ.setdefault("DS_ACCELERATOR", "cuda")
    try:
        import torch
        import torch.distributed as dist
        import deepspeed
        from transformers import LlamaConfig, LlamaForCausalLM
    except Exception as e:  # the reference or its imports are broken on this image
        unavailable(f"import failed: {type(e).__name__}: {e}")

    rank, world, local = (int(os.environ.get(k, d)) for k, d in (("RANK", 0), ("WORLD_SIZE", 1), ("LOCAL_RANK", 0)))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    os.environ.setdefault("MASTER_PORT", "29511")
    tp = world if getattr(a, "parallelism", "tp") == "tp" else 1   # same layout as the other arm (tp: stock AutoTP + NCCL)
    replicas = world // tp
    B, ctx, n_new, D = a.sessions * tp, a.ctx, a.new_tokens, a.decode_steps
    try:
        cfg = LlamaConfig(vocab_size=32000, max_position_embeddings=ctx + n_new + D + 64, **LLAMA[a.model])
        torch.manual_seed(1234)
        if world > 1:
            deepspeed.init_distributed("nccl")
        # build directly on this rank's GPU in bf16 (a host-side fp32 copy per rank would need ~27 GB of RAM each)
        torch.set_default_dtype(torch.bfloat16)
        with torch.device(dev):
            model = LlamaForCausalLM(cfg)
        torch.set_default_dtype(torch.float32)
        model = model.eval()
        engine = deepspeed.init_inference(model, dtype=torch.bfloat16, tensor_parallel={"tp_size": tp},
                                          replace_with_kernel_inject=False)
    except Exception as e:
        unavailable(f"deepspeed.init_inference failed: {type(e).__name__}: {str(e)[:200]}")

    g = torch.Generator().manual_seed(7)
    prompt = torch.randint(0, 32000, (ctx + n_new, ), generator=g).pin_memory()  # restored ctx + new turn, recomputed
    ev = lambda: torch.cuda.Event(enable_timing=True)  # noqa: E731
    host_stage = torch.empty((B, 1), dtype=torch.long).pin_memory()

    @torch.no_grad()
    def one_step(rec):
        for _ in range(B):  # per-session TTFT: recompute the whole context, read the first token back
            t0, e0, e1 = time.perf_counter(), ev(), ev()
            e0.record()
            ids = prompt.to(dev, non_blocking=True)[None]
            out = engine(input_ids=ids, use_cache=True)
            tok = out.logits[0, -1].argmax().cpu()
            e1.record()
            e1.synchronize()
            if rec is not None:
                rec["ttft_dev"].append(e0.elapsed_time(e1))
                rec["ttft_host"].append((time.perf_counter() - t0) * 1e3)
            del out
        # batched KV cache for the decode phase (one batched prefill; part of the step, not of TTFT/decode)
        ids = prompt.to(dev, non_blocking=True)[None].repeat(B, 1)
        out = engine(input_ids=ids, use_cache=True, logits_to_keep=1)   # last-position logits only (B x ctx x vocab would be 17 GB at B=64)
        cache = out.past_key_values
        cur = out.logits[:, -1].argmax(-1, keepdim=True)
        torch.cuda.synchronize()                       # the batched prefill above is not part of the decode clock
        d0, d1, th = ev(), ev(), time.perf_counter()
        d0.record()
        for _ in range(D):
            host_stage.copy_(cur, non_blocking=True)   # D2H of the sampled ids into the pinned staging buffer ...
            torch.cuda.current_stream().synchronize()
            out = engine(input_ids=host_stage.to(dev, non_blocking=True), past_key_values=cache, use_cache=True)  # ... and H2D
            cache = out.past_key_values
            cur = out.logits[:, -1].argmax(-1, keepdim=True)
        cur.cpu()
        d1.record()
        d1.synchronize()
        if rec is not None:
            rec["dec_dev"].append(d0.elapsed_time(d1))
            rec["dec_host"].append((time.perf_counter() - th) * 1e3)
        del out, cache

    try:
        for _ in range(a.warmup):
            one_step(None)
        rec = {"ttft_dev": [], "ttft_host": [], "dec_dev": [], "dec_host": []}
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        s, e = ev(), ev()
        s.record()
        for _ in range(a.steps):
            one_step(rec)
        e.record()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
    except Exception as ex:
        unavailable(f"reference run failed: {type(ex).__name__}: {str(ex)[:200]}")
    stats = torch.tensor([s.elapsed_time(e), sum(rec["dec_dev"]), sum(rec["dec_host"]), statistics.median(rec["ttft_dev"]),
                          statistics.median(rec["ttft_host"])], device=dev)
    if world > 1:
        dist.all_reduce(stats, op=dist.ReduceOp.MAX)
    total_ms, dec_dev, dec_host, ttft_dev, ttft_host = stats.tolist()
    if rank == 0:
        toks = a.steps * B * D * replicas
        print(json.dumps({
            "metric": "decode tokens/sec over restored multi-turn sessions (p50 restored-context TTFT in ttft_p50_ms)",
            "value": round(toks / (dec_dev / 1e3), 2), "unit": "tokens/s", "n_gpus": world, "steps": a.steps, "warmup": a.warmup,
            "ms_per_step": round(total_ms / a.steps, 3), "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "bf16", "data": "synthetic tokens, random-init weights", "impl": "reference",
            "ttft_p50_ms": round(ttft_dev, 3),
            "reference_path": "deepspeed.init_inference (v1, no kernel injection" + (f", AutoTP + NCCL all-reduce at tp={tp}" if tp > 1 else "") +
                              ") over HF LlamaForCausalLM; ttft_p50_ms is RECOMPUTE time-to-first-token: the context is re-established "
                              "by recomputing the prompt because the v2/HCache engine cannot run on sm_100 (see BASELINE.md §3)",
            "config": {"model": a.model, "global_batch": B * replicas, "seq_len": ctx + n_new + D,
                       "parallelism": f"tp{tp}" if replicas == 1 else f"dp{replicas}",
                       "restored_ctx": ctx, "new_turn_tokens": n_new, "decode_steps": D, "sessions_per_gpu": a.sessions},
            "e2e": {"value": round(toks / (dec_host / 1e3), 2), "unit": "tokens/s", "ttft_p50_ms": round(ttft_host, 3),
                    "h2d_bytes_per_step": (B * (ctx + n_new) * 8 * 2 + B * D * 8) * replicas,
                    "d2h_bytes_per_step": B * (D + 1) * 8 * replicas},
        }), flush=True)
