"""Engine-level GPU tests: prefill/decode numerics vs the CPU oracle path, HCache restore == live KV."""
import pytest
import torch

pytestmark = [pytest.mark.gpu, pytest.mark.timeout(600)]


def _cfg(T=512, S=8, ctx=1024, mem=512 * 1024 * 1024, **hc):
    from hcache_deepspeed_b200.inference.v2 import RaggedInferenceEngineConfig
    return RaggedInferenceEngineConfig(state_manager={
        "max_tracked_sequences": 32, "max_ragged_batch_size": T, "max_ragged_sequence_count": S, "max_context": ctx,
        "memory_config": {"mode": "allocate", "size": mem}}, hcache=hc or {"enabled": True})


@pytest.mark.parametrize("arch", ["tiny-llama-test", "tiny-opt-test", "tiny-mixtral-test"])
def test_gpu_matches_cpu_oracle(arch):
    from hcache_deepspeed_b200.inference.v2 import build_engine_from_arch
    from hcache_deepspeed_b200.models.arch import named_arch
    spec = named_arch(arch)
    gpu = build_engine_from_arch(spec, _cfg(), seed=5, dtype=torch.bfloat16, device=torch.device("cuda"))
    cpu = build_engine_from_arch(spec, _cfg(mem=64 * 1024 * 1024), seed=5, dtype=torch.float32, device=torch.device("cpu"))
    # same weights: copy the GPU (bf16) weights into the CPU engine
    for (n, t_cpu), (_, t_gpu) in zip(cpu.model().w.named_tensors().items(), gpu.model().w.named_tensors().items()):
        t_cpu.copy_(t_gpu.float().cpu())
    g = torch.Generator().manual_seed(1)
    a, b, c = (torch.randint(0, spec.vocab, (n, ), generator=g) for n in (150, 7, 129))
    lg, _ = gpu.put([1, 2, 3], [a, b, c])
    lc, _ = cpu.put([1, 2, 3], [a, b, c])
    torch.testing.assert_close(lg.float().cpu(), lc, rtol=5e-2, atol=5e-2)
    for _ in range(3):  # decode steps (split-KV attention + split-K GEMMs)
        nxt = lc.argmax(-1)
        toks = [nxt[i:i + 1] for i in range(3)]
        lg, _ = gpu.put([1, 2, 3], toks)
        lc, _ = cpu.put([1, 2, 3], toks)
        torch.testing.assert_close(lg.float().cpu(), lc, rtol=5e-2, atol=5e-2)


def test_restore_equals_live_kv_llama7b_shapes():
    from hcache_deepspeed_b200.inference.v2 import build_engine_from_arch
    from hcache_deepspeed_b200.models.arch import named_arch
    spec = named_arch("llama-2-7b").scaled(n_layers=2)
    eng = build_engine_from_arch(spec, _cfg(T=1024, ctx=2048, mem=1 << 30), seed=3, device=torch.device("cuda"))
    g = torch.Generator().manual_seed(0)
    p1, p2 = torch.randint(0, spec.vocab, (700, ), generator=g), torch.randint(0, spec.vocab, (45, ), generator=g)
    _, l1 = eng.put([1], [p1])
    _, l2 = eng.put([1], [p2])  # second chunk of the same sequence
    saved = torch.cat([l1[0], l2[0]], dim=1)
    assert saved.shape == (2, 745, spec.hidden) and l1[0].is_pinned()
    eng.restore_kv([9], [torch.cat([p1, p2])], [saved], defer=False)  # the test reads the pages directly
    kv = eng.state_manager.kv_cache.full_cache
    s1, s9 = eng.state_manager.get_sequence(1), eng.state_manager.get_sequence(9)
    assert s9.seen_tokens == 745
    live = kv[:, s1.all_block_ids.long().cuda()].float()
    rest = kv[:, s9.all_block_ids.long().cuda()].float()
    n_full = 745 // 64
    torch.testing.assert_close(rest[:, :n_full], live[:, :n_full], rtol=2e-2, atol=2e-2)
    probe = torch.randint(0, spec.vocab, (5, ), generator=g)
    la, _ = eng.put([1], [probe], capture_latents=False)
    lb, _ = eng.put([9], [probe], capture_latents=False)
    torch.testing.assert_close(la.float(), lb.float(), rtol=3e-2, atol=3e-2)


def test_deferred_restore_overlaps_next_put():
    """restore_kv(defer=True) + put == restore_kv(defer=False) + put, for one and for two sessions per call."""
    from hcache_deepspeed_b200.inference.v2 import build_engine_from_arch
    from hcache_deepspeed_b200.models.arch import named_arch
    spec = named_arch("llama-2-7b").scaled(n_layers=3)
    eng = build_engine_from_arch(spec, _cfg(T=1024, ctx=2048, mem=1 << 30), seed=4, device=torch.device("cuda"))
    g = torch.Generator().manual_seed(0)
    pa, pb = torch.randint(0, spec.vocab, (500, ), generator=g), torch.randint(0, spec.vocab, (130, ), generator=g)
    _, lat = eng.put([1, 2], [pa, pb])
    turn = [torch.randint(0, spec.vocab, (40, ), generator=g), torch.randint(0, spec.vocab, (3, ), generator=g)]
    eng.restore_kv([10, 11], [pa, pb], lat, defer=False)
    want, _ = eng.put([10, 11], turn, capture_latents=False)
    eng.restore_kv([20, 21], [pa, pb], lat, defer=True)
    got, _ = eng.put([20, 21], turn, capture_latents=False)
    # same kernels, same inputs; only the order of fp32 split-K atomics may differ
    torch.testing.assert_close(got.float(), want.float(), rtol=1e-2, atol=2e-2)
    live, _ = eng.put([1, 2], turn, capture_latents=False)
    torch.testing.assert_close(got.float(), live.float(), rtol=3e-2, atol=3e-2)
    # a deferred restore that is never followed by a forward is completed by flush
    eng.restore_kv([30], [pa], [lat[0]], defer=True)
    eng.flush(30)
    torch.cuda.synchronize()


def test_cuda_graph_decode_matches_eager():
    from hcache_deepspeed_b200.inference.v2 import RaggedInferenceEngineConfig, build_engine_from_arch
    from hcache_deepspeed_b200.models.arch import named_arch
    spec = named_arch("llama-2-7b").scaled(n_layers=2)
    engines = []
    for graphs in (False, True):
        cfg = _cfg(T=512, ctx=2048, mem=1 << 30)
        cfg.use_cuda_graph_decode = graphs
        engines.append(build_engine_from_arch(spec, cfg, seed=9, device=torch.device("cuda")))
    g = torch.Generator().manual_seed(2)
    prompts = [torch.randint(0, spec.vocab, (n, ), generator=g) for n in (100, 37, 260)]
    outs, fed = [], []
    for k, eng in enumerate(engines):
        lg, _ = eng.put([1, 2, 3], prompts, capture_latents=False)
        seq = [lg.float().cpu()]
        for step in range(8):  # crosses a KV block boundary and several graph replays
            if k == 0:
                fed.append(lg.argmax(-1).cpu())
            cur = fed[step]  # both engines are fed the eager engine's tokens
            lg, _ = eng.put([1, 2, 3], [cur[i:i + 1] for i in range(3)], capture_latents=False)
            seq.append(lg.float().cpu().clone())
        outs.append(seq)
    assert engines[1].graph_replays >= 5 and engines[0].graph_replays == 0
    # split-K partial sums meet through fp32 atomics whose order differs run to run: equal up to bf16 rounding
    for a, b in zip(*outs):
        torch.testing.assert_close(a, b, rtol=2e-2, atol=4e-2)


def test_async_capture_and_flush():
    from hcache_deepspeed_b200.inference.v2 import LatentHandle, build_engine_from_arch
    eng = build_engine_from_arch("tiny-llama-test", _cfg(enabled=True, capture="async"), seed=2, device=torch.device("cuda"))
    free0 = int(eng.free_blocks[0])
    _, lat = eng.put([4], [torch.arange(100)])
    assert isinstance(lat[0], LatentHandle)
    t = lat[0].wait()
    assert t.shape == (2, 100, 256) and torch.isfinite(t.float()).all()
    eng.flush(4)
    assert int(eng.free_blocks[0]) == free0


def test_smoke_entry():
    import __graft_entry__ as ge
    ge.smoke()


def test_graphed_callable_replays_and_tracks_inputs():
    """GraphedCallable (DSUNet / DSVAE / DSClipEncoder back end): capture once per signature, replay with fresh inputs."""
    from hcache_deepspeed_b200.model_implementations.diffusers import DSUNet
    from hcache_deepspeed_b200.model_implementations.features import GraphedCallable
    lin = torch.nn.Linear(64, 64).cuda().half()
    g = GraphedCallable(lambda x, scale=1.0: lin(x) * scale)
    for i in range(3):
        x = torch.randn(8, 64, device="cuda", dtype=torch.half)
        torch.testing.assert_close(g(x, scale=2.0), lin(x) * 2.0)
    assert len(g._graphs) == 1
    g(torch.randn(4, 64, device="cuda", dtype=torch.half), scale=2.0)
    assert len(g._graphs) == 2

    class UNet(torch.nn.Module):
        in_channels = 4

        def __init__(self):
            super().__init__()
            self.conv = torch.nn.Conv2d(4, 4, 3, padding=1)

        def forward(self, sample, t, ehs, return_dict=True):
            return self.conv(sample) * t + ehs.mean()

    u = DSUNet(UNet().cuda().half())
    for i in range(2):
        s, t, e = (torch.randn(1, 4, 16, 16, device="cuda", dtype=torch.half), torch.tensor(float(i + 1), device="cuda").half(),
                   torch.randn(1, 7, 8, device="cuda", dtype=torch.half))
        torch.testing.assert_close(u(s, t, e), u.unet(s, t, e), rtol=1e-2, atol=1e-2)
    assert u.cuda_graph_created


def test_fp8_engine_mode_gpu():
    """wf8af8 on the GPU: fp8 tensor-core GEMMs for every layer linear; logits track the bf16 engine, HCache restore
    reproduces the live KV, and small decode batches (the CUDA-graph path) run with the fp8 kernels inside."""
    from hcache_deepspeed_b200.inference.v2 import RaggedInferenceEngineConfig, build_engine_from_arch
    from hcache_deepspeed_b200.models.arch import named_arch
    spec = named_arch("llama-2-7b").scaled(n_layers=2)

    def build(mode):
        cfg = RaggedInferenceEngineConfig(state_manager={
            "max_tracked_sequences": 16, "max_ragged_batch_size": 512, "max_ragged_sequence_count": 8, "max_context": 1024,
            "memory_config": {"mode": "allocate", "size": 512 * 1024 * 1024}}, hcache={"enabled": True},
            quantization={"quantization_mode": mode})
        return build_engine_from_arch(spec, cfg, seed=3, dtype=torch.bfloat16, device=torch.device("cuda"))

    ref_eng, eng = build(None), build("wf8af8")
    g = torch.Generator().manual_seed(0)
    prompt = torch.randint(0, spec.vocab, (300, ), generator=g)
    want, _ = ref_eng.put([1], [prompt])
    got, lat = eng.put([1], [prompt])
    cos = torch.nn.functional.cosine_similarity(got.float().flatten(), want.float().flatten(), dim=0)
    assert cos > 0.98, float(cos)
    eng.restore_kv([2], [prompt], [lat[0]])
    nxt = torch.tensor([11])
    a, _ = eng.put([1], [nxt], capture_latents=False)
    b, _ = eng.put([2], [nxt], capture_latents=False)
    torch.testing.assert_close(a.float(), b.float(), rtol=2e-2, atol=6e-2)
    for _ in range(3):
        a, _ = eng.put([1, 2], [nxt, nxt], capture_latents=False)
    assert torch.isfinite(a.float()).all()


@pytest.mark.parametrize("kinds", [["kv"] * 6, ["hidden", "hidden", "hidden", "kv", "kv", "kv"], ["none", "hidden", "kv", "kv", "kv", "kv"]])
def test_layer_pipelined_kv_reload(kinds):
    """Mixed restore plans with offloaded KV pages streamed layer by layer (layer-major host tensor, deferred into the next
    put) reproduce the live session's logits; more KV layers than staging slots exercises slot reuse."""
    from hcache_deepspeed_b200.inference.v2 import build_engine_from_arch
    from hcache_deepspeed_b200.models.arch import named_arch
    spec = named_arch("llama-2-7b").scaled(n_layers=6)
    eng = build_engine_from_arch(spec, _cfg(T=1024, ctx=2048, mem=1 << 30), seed=4, device=torch.device("cuda"))
    g = torch.Generator().manual_seed(0)
    toks = torch.randint(0, spec.vocab, (700, ), generator=g)
    _, lat = eng.put([1], [toks])
    seq = eng.state_manager.get_sequence(1)
    hl = [i for i, k in enumerate(kinds) if k == "hidden"]
    kl = [i for i, k in enumerate(kinds) if k == "kv"]
    rl = [i for i, k in enumerate(kinds) if k == "none"]
    kv = eng.state_manager.kv_cache.offload_layers(seq.all_block_ids, kl)          # [n_blocks, len(kl), 2, kvh, bs, d] on host
    kv_major = kv.transpose(0, 1).contiguous().pin_memory()
    turn = torch.randint(0, spec.vocab, (40, ), generator=g)
    for uid, defer in ((10, False), (11, True)):
        eng.restore_session(uid, toks, hidden=lat[0][hl] if hl else None, hidden_layers=hl, kv=kv_major, kv_layers=kl,
                            kv_layer_major=True, recompute_layers=rl, defer=defer)
    want, _ = eng.put([1], [turn], capture_latents=False)
    for uid in (10, 11):
        got, _ = eng.put([uid], [turn], capture_latents=False)
        torch.testing.assert_close(got.float(), want.float(), rtol=3e-2, atol=3e-2)
    eng.restore_session(12, toks, hidden=lat[0][hl] if hl else None, hidden_layers=hl, kv=kv_major, kv_layers=kl,
                        kv_layer_major=True, recompute_layers=rl, defer=True)
    eng.flush(12)                                    # never followed by a forward: flush completes the transfers
    torch.cuda.synchronize()


def test_decode_graph_capture_survives_garbage_collection():
    """Regression: Python's cyclic collector finalising an old CUDAGraph inside an open capture invalidated the capture
    (cudaErrorStreamCaptureInvalidated, order-dependent). The capture helper collects first and keeps the collector off."""
    import gc
    from hcache_deepspeed_b200.inference.v2 import build_engine_from_arch
    from hcache_deepspeed_b200.models.arch import named_arch
    from hcache_deepspeed_b200.utils.cuda_graph import capture

    class Holder:
        pass

    x = torch.ones(64, device="cuda")
    for _ in range(4):           # dead reference cycles that own live CUDA graphs
        h = Holder()
        h.me = h
        h.graph, _ = capture(lambda: x * 2)
    del h
    spec = named_arch("llama-2-7b").scaled(n_layers=2)
    eng = build_engine_from_arch(spec, _cfg(T=256, ctx=512, mem=256 << 20), seed=1, device=torch.device("cuda"))
    eng.put([1], [torch.randint(0, spec.vocab, (20, ))], capture_latents=False)
    old = gc.get_threshold()
    gc.set_threshold(1, 1, 1)    # a full collection at (almost) every allocation: certain to fire inside the capture window
    try:
        for _ in range(5):
            out, _ = eng.put([1], [torch.tensor([3])], capture_latents=False)
    finally:
        gc.set_threshold(*old)
    assert eng.graph_replays >= 2 and torch.isfinite(out.float()).all()


def test_gds_cufile_roundtrip(tmp_path):
    """GPUDirect Storage path executed for real: device tensor -> file -> device tensor through libcufile, synchronous
    and asynchronous (pieces on worker threads), byte-exact; reports whether nvidia-fs or the library's compatibility
    mode served it. Skipped (with the reason) only when libcufile's driver cannot be opened on this machine."""
    from hcache_deepspeed_b200.io import gds
    if not gds.gds_available():     # probed in a child process with a time limit: a blocked cuFile call cannot hang the suite
        pytest.skip(f"cuFile is not usable on this machine: {gds.gds_unavailable_reason()}")
    h = gds.GDSHandle(block_size=1 << 20, threads=4)
    n = 8 << 20
    src = torch.randint(0, 255, (n, ), dtype=torch.uint8, device="cuda")
    path = str(tmp_path / "gds.bin")
    try:
        assert h.sync_pwrite(src, path) == n
    except (OSError, RuntimeError) as e:     # file system without O_DIRECT / cuFile support under the test directory
        pytest.skip(f"cuFile cannot use {tmp_path}: {e}")
    dst = torch.zeros_like(src)
    assert h.sync_pread(dst, path) == n
    torch.cuda.synchronize()
    assert torch.equal(src, dst)
    dst2 = torch.zeros_like(src)
    r1 = h.async_pread(dst2[:n // 2], path, 0)
    r2 = h.async_pread(dst2[n // 2:], path, n // 2)
    assert h.wait(r1) == n // 2 and h.wait(r2) == n // 2
    torch.cuda.synchronize()
    assert torch.equal(src, dst2)
    print(f"cuFile mode: {gds.gds_mode()}")


def _hf_model(kind):
    torch.manual_seed(0)
    if kind == "opt":
        from transformers import OPTConfig, OPTForCausalLM
        cfg = OPTConfig(vocab_size=2048, hidden_size=512, num_hidden_layers=3, ffn_dim=2048, num_attention_heads=8,
                        max_position_embeddings=1024, word_embed_proj_dim=512, do_layer_norm_before=True, dropout=0.0)
        return OPTForCausalLM(cfg).eval()
    if kind == "gpt2":
        from transformers import GPT2Config, GPT2LMHeadModel
        return GPT2LMHeadModel(GPT2Config(vocab_size=2048, n_embd=512, n_layer=3, n_head=8, n_positions=1024, resid_pdrop=0.0,
                                          embd_pdrop=0.0, attn_pdrop=0.0)).eval()
    if kind == "gptj":
        from transformers import GPTJConfig, GPTJForCausalLM
        return GPTJForCausalLM(GPTJConfig(vocab_size=2048, n_embd=512, n_layer=3, n_head=8, n_positions=1024, rotary_dim=32,
                                          resid_pdrop=0.0, embd_pdrop=0.0, attn_pdrop=0.0)).eval()
    if kind == "gpt_neox":
        from transformers import GPTNeoXConfig, GPTNeoXForCausalLM
        return GPTNeoXForCausalLM(GPTNeoXConfig(vocab_size=2048, hidden_size=512, num_hidden_layers=3, num_attention_heads=4,
                                                intermediate_size=2048, rotary_pct=0.25, max_position_embeddings=1024,
                                                hidden_dropout=0.0, attention_dropout=0.0)).eval()
    if kind == "gpt_neo":       # local layers: window 64 < the 200-token prompt
        from transformers import GPTNeoConfig, GPTNeoForCausalLM
        return GPTNeoForCausalLM(GPTNeoConfig(vocab_size=2048, hidden_size=512, num_layers=4, num_heads=8, window_size=64,
                                              attention_types=[[["global", "local"], 2]], max_position_embeddings=1024,
                                              resid_dropout=0.0, embed_dropout=0.0, attention_dropout=0.0)).eval()
    if kind == "mistral_window":
        from transformers import MistralConfig, MistralForCausalLM
        return MistralForCausalLM(MistralConfig(vocab_size=2048, hidden_size=512, intermediate_size=1408, num_hidden_layers=3,
                                                num_attention_heads=4, num_key_value_heads=2, sliding_window=96,
                                                max_position_embeddings=1024)).eval()
    if kind == "bloom":
        from transformers import BloomConfig, BloomForCausalLM
        return BloomForCausalLM(BloomConfig(vocab_size=2048, hidden_size=768, n_layer=3, n_head=6, hidden_dropout=0.0,
                                            attention_dropout=0.0)).eval()
    if kind == "falcon_alibi":
        from transformers import FalconConfig, FalconForCausalLM
        return FalconForCausalLM(FalconConfig(vocab_size=2048, hidden_size=512, num_hidden_layers=3, num_attention_heads=4,
                                              multi_query=False, parallel_attn=False, new_decoder_architecture=False, alibi=True,
                                              bias=True, hidden_dropout=0.0, attention_dropout=0.0)).eval()
    from transformers import LlamaConfig, LlamaForCausalLM
    cfg = LlamaConfig(vocab_size=2048, hidden_size=512, intermediate_size=1408, num_hidden_layers=3, num_attention_heads=4,
                      num_key_value_heads=2, max_position_embeddings=1024)
    return LlamaForCausalLM(cfg).eval()


# gpt2 / gptj / gpt_neox: the families the reference serves through its v1 transformer_inference op only
@pytest.mark.parametrize("kind", ["opt", "llama", "gpt2", "gptj", "gpt_neox", "bloom", "falcon_alibi", "gpt_neo",
                                  "mistral_window"])
def test_v1_init_inference_kernel_inject_gpu(kind):
    """``init_inference(replace_with_kernel_inject=True)`` on the GPU (v1 API): the injected engine runs the sm_100a
    kernels, its logits track the HuggingFace model's fp32 logits for a prompt and for a continuation over the live KV
    cache, and a flushed session restored from its saved hidden states continues identically."""
    import hcache_deepspeed_b200 as ds
    from hcache_deepspeed_b200.ops import kernels, native
    hf32 = _hf_model(kind)
    vocab = hf32.config.vocab_size
    g = torch.Generator().manual_seed(1)
    ctx = torch.randint(0, vocab, (1, 200), generator=g)
    turn = torch.randint(0, vocab, (1, 16), generator=g)
    with torch.no_grad():
        want_ctx = hf32(input_ids=ctx).logits
        want_turn = hf32(input_ids=torch.cat([ctx, turn], dim=1)).logits[:, 200:]
    import copy
    model = copy.deepcopy(hf32).to(torch.bfloat16).cuda()
    before = kernels.launch_count
    eng = ds.init_inference(model, dtype=torch.bfloat16, replace_with_kernel_inject=True,
                            hcache={"enabled": True, "max_context": 1024, "max_batch_tokens": 512, "kv_memory_bytes": 128 << 20})
    out = eng(input_ids=ctx.cuda(), session_ids=[7]).logits.float().cpu()
    assert native.cuda_lib_loaded() and kernels.launch_count > before, "the injected engine did not run the native kernels"
    cos = torch.nn.functional.cosine_similarity(out.flatten(), want_ctx.flatten(), dim=0)
    assert cos > 0.995, float(cos)
    lat = eng.hidden_states(7)
    assert tuple(lat.shape) == (hf32.config.num_hidden_layers, 200, hf32.config.hidden_size)
    live = eng(input_ids=turn.cuda(), session_ids=[7]).logits.float().cpu()
    cos = torch.nn.functional.cosine_similarity(live.flatten(), want_turn.flatten(), dim=0)
    assert cos > 0.995, float(cos)
    # end of turn: drop the KV cache, rebuild it from the saved hidden states instead of recomputing the prompt
    eng.flush(7)
    eng.restore_hidden_states(7, ctx, lat)
    restored = eng(input_ids=turn.cuda(), session_ids=[7]).logits.float().cpu()
    torch.testing.assert_close(restored, live, rtol=3e-2, atol=3e-2)


@pytest.mark.parametrize("mode", ["mapped", "wc"])
def test_fast_host_buffer_staging(mode, monkeypatch):
    """Natively allocated staging memory (``hds_host_buffer_alloc``: pinned + mapped (+ write-combined), ref
    ragged/csrc/fast_host_buffer.cu): the tensor view counts as pinned for torch, async copies out of it work, it has a
    device-side address, and an engine whose ragged batch stages through it produces bit-identical logits."""
    from hcache_deepspeed_b200.inference.v2 import build_engine_from_arch
    from hcache_deepspeed_b200.inference.v2.ragged.fast_host_buffer import FastHostBuffer, allocate_fast_host_buffer
    from hcache_deepspeed_b200.models.arch import named_arch
    torch.zeros(1, device="cuda")        # torch answers is_pinned() with False until it has initialised CUDA itself
    buf = FastHostBuffer(4096, torch.int32, mapped=True, write_combined=(mode == "wc"))
    assert buf.tensor.is_pinned() and buf.device_ptr()
    buf.tensor.numpy()[:] = torch.arange(4096, dtype=torch.int32).numpy()
    dev = torch.empty(4096, dtype=torch.int32, device="cuda")
    dev.copy_(buf.tensor, non_blocking=True)
    torch.cuda.synchronize()
    assert torch.equal(dev.cpu(), torch.arange(4096, dtype=torch.int32))
    buf.free()
    like = allocate_fast_host_buffer(torch.empty(3, 5, dtype=torch.float32))
    assert like.shape == (3, 5) and like.is_pinned()

    spec = named_arch("tiny-llama-test")
    g = torch.Generator().manual_seed(2)
    a, b = torch.randint(0, spec.vocab, (150, ), generator=g), torch.randint(0, spec.vocab, (9, ), generator=g)
    nxt = torch.randint(0, spec.vocab, (2, ), generator=g)
    plain = build_engine_from_arch(spec, _cfg(), seed=5, dtype=torch.bfloat16, device=torch.device("cuda"))
    want, _ = plain.put([1, 2], [a, b])
    want2, _ = plain.put([1, 2], [nxt[0:1], nxt[1:2]])
    monkeypatch.setenv("HDS_FAST_HOST_BUFFER", mode)
    eng = build_engine_from_arch(spec, _cfg(), seed=5, dtype=torch.bfloat16, device=torch.device("cuda"))
    assert getattr(eng._batch, "_fast", None) is not None
    got, _ = eng.put([1, 2], [a, b])
    got2, _ = eng.put([1, 2], [nxt[0:1], nxt[1:2]])
    torch.testing.assert_close(got.float(), want.float(), rtol=0, atol=0)
    torch.testing.assert_close(got2.float(), want2.float(), rtol=0, atol=0)
