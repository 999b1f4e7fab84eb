"""Worker functions for the multi-rank training-runtime tests (must be importable by spawned processes)."""
import os

import torch


class Stack(torch.nn.ModuleList):
    def forward(self, x):
        for l in self:
            x = torch.tanh(l(x))
        return x


class TinyNet(torch.nn.Module):
    def __init__(self, hidden=32, nlayers=3):
        super().__init__()
        self.layers = Stack([torch.nn.Linear(hidden, hidden) for _ in range(nlayers)])
        self.head = torch.nn.Linear(hidden, 1)

    def forward(self, x, y):
        return torch.nn.functional.mse_loss(self.head(self.layers(x)).squeeze(-1), y)


def make_batches(n, bs, hidden=32, seed=0):
    g = torch.Generator().manual_seed(seed)
    return [(torch.randn(bs, hidden, generator=g), torch.randn(bs, generator=g)) for _ in range(n)]


def reference_run(steps, global_bs, lr=1e-2, hidden=32, clip=0.0, wd=0.0):
    """Single-process fp32 AdamW on the full global batch — what every parallel configuration must reproduce."""
    torch.manual_seed(1234)
    m = TinyNet(hidden)
    opt = torch.optim.AdamW(m.parameters(), lr=lr, weight_decay=wd)
    losses = []
    for x, y in make_batches(steps, global_bs, hidden):
        loss = m(x, y)
        loss.backward()
        if clip > 0:
            torch.nn.utils.clip_grad_norm_(m.parameters(), clip)
        opt.step()
        opt.zero_grad()
        losses.append(float(loss))
    return losses, {k: v.clone() for k, v in m.state_dict().items()}


def zero_worker(rank, world, stage, steps, gas, offload, dtype_name, ckpt_dir):
    import hcache_deepspeed_b200 as ds
    ds.init_distributed(dist_backend="gloo")
    torch.manual_seed(1234)
    hidden, micro = 32, 4
    model = TinyNet(hidden)
    cfg = {"train_micro_batch_size_per_gpu": micro, "gradient_accumulation_steps": gas,
           "optimizer": {"type": "AdamW", "params": {"lr": 1e-2, "weight_decay": 0.0}},
           "zero_optimization": {"stage": stage, "stage3_param_persistence_threshold": 0}, "steps_per_print": 1000}
    if os.environ.get("HDS_TEST_MICS"):
        cfg["zero_optimization"]["mics_shard_size"] = int(os.environ["HDS_TEST_MICS"])
    if os.environ.get("HDS_TEST_Z3_REUSE") is not None:
        cfg["zero_optimization"]["stage3_max_reuse_distance"] = int(os.environ["HDS_TEST_Z3_REUSE"])
    if offload == "nvme":
        import tempfile
        cfg["zero_optimization"]["offload_optimizer"] = {"device": "nvme", "pipeline_read": True, "pipeline_write": True,
                                                         "nvme_path": tempfile.mkdtemp(prefix=f"hds_optswap_r{rank}_")}
        cfg["aio"] = {"block_size": 4096}
    elif offload:
        cfg["zero_optimization"]["offload_optimizer"] = {"device": "cpu"}
    if dtype_name == "bf16":
        cfg["bf16"] = {"enabled": True}
    elif dtype_name == "fp16":
        cfg["fp16"] = {"enabled": True, "initial_scale_power": 8}
    if os.environ.get("HDS_TEST_Z3_LEAF") == "1":
        from hcache_deepspeed_b200.utils.z3_leaf_module import get_z3_leaf_modules, set_z3_leaf_modules, z3_leaf_parameter
        assert set_z3_leaf_modules(model, ["Stack"]) == [model.layers]
    compile_on = os.environ.get("HDS_TEST_Z3_COMPILE") == "1"
    if compile_on:
        cfg["compile"] = {"deepcompile": True, "selective_gather_budget_bytes": 3000, "prefetch_bucket_size": 100000,
                          "offload_opt_states": True}
    engine, opt, _, _ = ds.initialize(model=model, config=cfg)
    if compile_on:
        assert engine.is_deepcompile_enabled() and not engine.is_compiled
        engine.compile()
    if os.environ.get("HDS_TEST_Z3_LEAF") == "1":
        assert get_z3_leaf_modules(engine.module) == [engine.module.layers]
        assert all(z3_leaf_parameter(p) for p in engine.module.layers.parameters())
        assert not any(z3_leaf_parameter(p) for p in engine.module.head.parameters())
    global_bs = micro * gas * world
    losses = []
    for step, (x, y) in enumerate(make_batches(steps, global_bs, hidden)):
        step_loss = 0.0
        for a in range(gas):
            lo = (a * world + rank) * micro
            xb, yb = x[lo:lo + micro], y[lo:lo + micro]
            if dtype_name != "fp32":
                xb = xb.to(torch.bfloat16 if dtype_name == "bf16" else torch.float16)
                yb = yb.to(xb.dtype)
            loss = engine(xb, yb)
            engine.backward(loss)
            engine.step()
            step_loss += float(loss) / gas
        losses.append(step_loss)
        if ckpt_dir and step == steps // 2 - 1:
            engine.save_checkpoint(ckpt_dir, tag="mid", client_state={"note": 7})
    out = {"losses": losses, "global_steps": engine.global_steps}
    if offload == "nvme":
        sw = engine.optimizer.optimizer_swapper
        out["swap_stats"] = dict(sw.stats)
        out["masters_resident"] = [m.numel() for m in engine.optimizer.fp32_partitioned_groups_flat]
    if compile_on:
        out["compile_report"] = engine.compile_report
        if stage == 3:
            out["prefetch_stats"] = dict(engine.optimizer.parameter_offload.stats)
            out["persistent"] = sum(1 for p in engine.module.parameters() if p.ds_persist)
    if stage == 3 and os.environ.get("HDS_TEST_MICS"):
        g = engine.optimizer.mics_comm_groups
        out["mics"] = (type(engine.optimizer).__name__, g.param_shard_size, g.param_repli_size,
                       max(p.ds_partition_numel * g.param_shard_size - p.ds_numel for p in engine.module.parameters()))
    if stage == 3:
        out["offload_stats"] = dict(engine.optimizer.parameter_offload.stats)
        from hcache_deepspeed_b200.runtime.zero import GatheredParameters
        sd = {}
        for n, p in engine.module.named_parameters():
            with GatheredParameters([p]):
                sd[n] = p.data.detach().float().clone()
        out["state"] = sd
    else:
        out["state"] = {k: v.detach().float().clone() for k, v in engine.module.state_dict().items()}
    if ckpt_dir:
        # resume from the mid-run checkpoint in a fresh engine and replay the second half
        torch.manual_seed(99)
        model2 = TinyNet(hidden)
        engine2, _, _, _ = ds.initialize(model=model2, config=cfg)
        path, client = engine2.load_checkpoint(ckpt_dir, tag="mid")
        assert path is not None and client["note"] == 7 and engine2.global_steps == steps // 2
        relosses = []
        for step, (x, y) in enumerate(make_batches(steps, global_bs, hidden)):
            if step < steps // 2:
                continue
            sl = 0.0
            for a in range(gas):
                lo = (a * world + rank) * micro
                loss = engine2(x[lo:lo + micro], y[lo:lo + micro])
                engine2.backward(loss)
                engine2.step()
                sl += float(loss) / gas
            relosses.append(sl)
        out["resumed_losses"] = relosses
    return out


class _Stage(torch.nn.Module):
    def __init__(self, hidden):
        super().__init__()
        self.l = torch.nn.Linear(hidden, hidden)

    def forward(self, x):
        return torch.tanh(self.l(x))


def pipe_worker(rank, world, stages, steps, micro_batches):
    import hcache_deepspeed_b200 as ds
    from hcache_deepspeed_b200.runtime.pipe import LayerSpec, PipelineModule
    ds.init_distributed(dist_backend="gloo")
    hidden, micro = 16, 4
    torch.manual_seed(7)
    specs = [LayerSpec(_Stage, hidden) for _ in range(4)] + [LayerSpec(torch.nn.Linear, hidden, 1)]
    net = PipelineModule(layers=specs, num_stages=stages, loss_fn=lambda out, y: torch.nn.functional.mse_loss(out.squeeze(-1), y),
                         partition_method="uniform", seed_layers=True, base_seed=100)
    dp = world // stages
    cfg = {"train_micro_batch_size_per_gpu": micro, "gradient_accumulation_steps": micro_batches,
           "optimizer": {"type": "Adam", "params": {"lr": 1e-2, "torch_adam": True}}, "steps_per_print": 1000}
    engine, _, _, _ = ds.initialize(model=net, config=cfg)
    dp_rank = engine.grid.get_data_parallel_id()
    losses = []
    for step in range(steps):
        g = torch.Generator().manual_seed(1000 + step)
        x = torch.randn(micro_batches * dp * micro, hidden, generator=g)
        y = torch.randn(micro_batches * dp * micro, generator=g)
        mbs = []
        for m in range(micro_batches):
            lo = (m * dp + dp_rank) * micro
            mbs.append((x[lo:lo + micro], y[lo:lo + micro]))
        losses.append(float(engine.train_batch(iter(mbs))))
    ev = float(engine.eval_batch(iter(mbs)))
    return {"losses": losses, "eval": ev, "stage": engine.stage_id, "parts": list(net.parts)}


def pipe_reference(steps, micro_batches, dp):
    hidden, micro = 16, 4
    layers = []
    for i in range(4):
        torch.manual_seed(100 + i)
        layers.append(_Stage(hidden))
    torch.manual_seed(104)
    layers.append(torch.nn.Linear(hidden, 1))
    net = torch.nn.Sequential(*layers)
    opt = torch.optim.Adam(net.parameters(), lr=1e-2)
    losses = []
    for step in range(steps):
        g = torch.Generator().manual_seed(1000 + step)
        x = torch.randn(micro_batches * dp * micro, hidden, generator=g)
        y = torch.randn(micro_batches * dp * micro, generator=g)
        loss = torch.nn.functional.mse_loss(net(x).squeeze(-1), y)
        loss.backward()
        opt.step()
        opt.zero_grad()
        losses.append(float(loss))
    return losses


def ckpt_tools_worker(rank, world, stage, save_dir, mode):
    """mode 'save': train a few steps with ZeRO and save; mode 'universal': resume a (different dp size) run from the
    universal checkpoint and return the parameters + one more step's loss."""
    import hcache_deepspeed_b200 as ds
    ds.init_distributed(dist_backend="gloo")
    torch.manual_seed(1234)
    model = TinyNet(32)
    cfg = {"train_micro_batch_size_per_gpu": 4, "optimizer": {"type": "AdamW", "params": {"lr": 1e-2, "weight_decay": 0.0}},
           "zero_optimization": {"stage": stage, "stage3_param_persistence_threshold": 0}, "steps_per_print": 1000}
    if mode == "universal_config":
        cfg["checkpoint"] = {"load_universal": True}
    engine, _, _, _ = ds.initialize(model=model, config=cfg)
    if mode == "universal_config":
        # save_dir is the ORIGINAL checkpoint folder: the tag comes from latest_universal, steps come back too
        path, client = engine.load_checkpoint(save_dir)
        m = engine.optimizer.single_partition_of_fp32_groups[0]
        return {"path": path, "steps": engine.global_steps, "state": {k: v.detach().float().clone() for k, v in engine.module.state_dict().items()},
                "exp_avg_norm": float(engine.optimizer.optimizer.state[m]["exp_avg"].norm())}
    if mode == "save":
        for x, y in make_batches(3, 4 * world, 32):
            lo = rank * 4
            loss = engine(x[lo:lo + 4], y[lo:lo + 4])
            engine.backward(loss)
            engine.step()
        engine.save_checkpoint(save_dir, tag="t3")
        if stage == 3:
            from hcache_deepspeed_b200.runtime.zero import GatheredParameters
            sd = {}
            for n, p in engine.module.named_parameters():
                with GatheredParameters([p]):
                    sd[n] = p.data.detach().float().clone()
            return sd
        return {k: v.detach().float().clone() for k, v in engine.module.state_dict().items()}
    from hcache_deepspeed_b200.checkpoint import load_universal_into_engine
    n = load_universal_into_engine(engine, save_dir)
    if stage == 3:
        from hcache_deepspeed_b200.runtime.zero import GatheredParameters
        state = {}
        for name, p in engine.module.named_parameters():
            with GatheredParameters([p]):
                state[name] = p.data.detach().float().clone()
        m = engine.optimizer.fp32_partitioned_groups_flat[0]
        sq = torch.tensor([float(engine.optimizer.optimizer.state[m]["exp_avg"].norm()) ** 2])
        ds.comm.all_reduce(sq)
        return {"restored": n, "state": state, "exp_avg_norm": float(sq.sqrt())}
    out = {"restored": n, "state": {k: v.detach().float().clone() for k, v in engine.module.state_dict().items()}}
    m = engine.optimizer.single_partition_of_fp32_groups[0]
    out["exp_avg_norm"] = float(engine.optimizer.optimizer.state[m]["exp_avg"].norm())
    return out


def zero3_frozen_worker(rank, world, ckpt_dir):
    """ZeRO-3 with a frozen layer and per-rank RNG seeds: zero.Init must take rank 0's initial values; a checkpoint must
    carry the frozen parameters (they are in no optimizer group) and restore them — full load and module-only load — and
    zero_to_fp32 must rebuild the complete fp32 state dict from the shards."""
    import hcache_deepspeed_b200 as ds
    from hcache_deepspeed_b200.checkpoint.zero_to_fp32 import get_fp32_state_dict_from_zero_checkpoint
    from hcache_deepspeed_b200.runtime.zero import GatheredParameters
    ds.init_distributed(dist_backend="gloo")
    hidden, micro = 32, 4
    cfg = {"train_micro_batch_size_per_gpu": micro, "optimizer": {"type": "AdamW", "params": {"lr": 1e-2}},
           "zero_optimization": {"stage": 3, "stage3_param_persistence_threshold": 40}, "steps_per_print": 1000}

    def gathered(engine):
        sd = {}
        for n, p in engine.module.named_parameters():
            with GatheredParameters([p]):
                sd[n] = p.data.detach().float().clone()
        return sd

    def build(seed):
        torch.manual_seed(seed)                  # DIFFERENT initial weights on every rank / engine
        model = TinyNet(hidden)
        for p in model.layers[0].parameters():
            p.requires_grad_(False)
        return ds.initialize(model=model, config=cfg)[0]

    engine = build(100 + rank)
    start = gathered(engine)
    torch.manual_seed(100)                       # what rank 0 drew
    want0 = {k: v.float() for k, v in TinyNet(hidden).state_dict().items()}
    init_matches_rank0 = all(torch.equal(start[k], want0[k]) for k in start)
    for x, y in make_batches(3, micro * world, hidden):
        loss = engine(x[rank * micro:(rank + 1) * micro], y[rank * micro:(rank + 1) * micro])
        engine.backward(loss)
        engine.step()
    trained = gathered(engine)
    frozen_unchanged = all(torch.equal(trained[k], start[k]) for k in trained if k.startswith("layers.0."))
    engine.save_checkpoint(ckpt_dir, tag="z3")
    out = {"init_matches_rank0": init_matches_rank0, "frozen_unchanged": frozen_unchanged, "trained": trained}
    e2 = build(500 + rank)
    e2.load_checkpoint(ckpt_dir, tag="z3")
    out["full_load"] = gathered(e2)
    e3 = build(900 + rank)
    e3.load_checkpoint(ckpt_dir, tag="z3", load_module_only=True)
    out["module_only_load"] = gathered(e3)
    if rank == 0:
        out["fp32_from_shards"] = {k: v.clone() for k, v in get_fp32_state_dict_from_zero_checkpoint(ckpt_dir, tag="z3").items()}
    return out


def zero_init_hf_worker(rank, world):
    """``zero.Init`` around a HuggingFace constructor: HF re-initialises every module after construction; the values must
    land in the partitions (ref: HF's deepspeed integration gathers params inside ``_init_weights``)."""
    import transformers as T
    import hcache_deepspeed_b200 as ds
    from hcache_deepspeed_b200.runtime.zero import GatheredParameters
    from hcache_deepspeed_b200.runtime.zero.partition_parameters import is_zero_param
    ds.init_distributed(dist_backend="gloo")
    cfg = T.GPT2Config(n_embd=64, n_layer=2, n_head=4, vocab_size=120, n_positions=128)
    torch.manual_seed(0)
    plain = T.GPT2LMHeadModel(cfg)
    torch.manual_seed(0 if rank == 0 else 77)          # only rank 0's initialisation counts
    with ds.zero.Init(config_dict_or_path={"train_micro_batch_size_per_gpu": 1, "zero_optimization": {"stage": 3}}):
        model = T.GPT2LMHeadModel(cfg)
    assert T.modeling_utils.PreTrainedModel._initialize_weights.__dict__.get("_ds_zero_wrapped") is None     # patch removed
    worst, n_zero = 0.0, 0
    for (n, p), (_, q) in zip(model.named_parameters(), plain.named_parameters()):
        n_zero += int(is_zero_param(p))
        with GatheredParameters([p]):
            worst = max(worst, float((p.data.float() - q.data.float()).abs().max()))
    with GatheredParameters([model.transformer.wte.weight]):
        std = float(model.transformer.wte.weight.data.float().std())
    out = {"worst": worst, "n_zero": n_zero, "wte_std": std}
    # from_pretrained inside the context: loads with partitioning suspended, then the whole model is partitioned
    import os
    import tempfile
    d = os.path.join(tempfile.gettempdir(), f"hds_zinit_{os.getppid()}")
    if rank == 0:
        plain.save_pretrained(d)
    torch.distributed.barrier()
    with ds.zero.Init(config_dict_or_path={"train_micro_batch_size_per_gpu": 1, "zero_optimization": {"stage": 3}}):
        loaded = T.AutoModelForCausalLM.from_pretrained(d)
    worst = 0.0
    for (n, p), (_, q) in zip(loaded.named_parameters(), plain.named_parameters()):
        assert is_zero_param(p), n
        with GatheredParameters([p]):
            worst = max(worst, float((p.data.float() - q.data.float()).abs().max()))
    engine, _, _, _ = ds.initialize(model=loaded, config={"train_micro_batch_size_per_gpu": 1, "zero_optimization": {"stage": 3},
                                                          "optimizer": {"type": "AdamW", "params": {"lr": 1e-3}}})
    ids = torch.randint(1, 120, (1, 9))
    engine.eval()           # GPT-2 has dropout
    plain.eval()
    with torch.no_grad():
        out["pretrained_logits"] = float((engine(input_ids=ids).logits - plain(ids).logits).abs().max())
    out["pretrained_worst"] = worst
    torch.distributed.barrier()
    if rank == 0:
        import shutil
        shutil.rmtree(d, ignore_errors=True)
    return out


def sparse_gradients_worker(rank, world, sparse_module):
    """``"sparse_gradients": true``: embedding gradients are exchanged as (rows, values); the result equals the dense
    all-reduce average."""
    import hcache_deepspeed_b200 as ds
    ds.init_distributed(dist_backend="gloo")

    class Net(torch.nn.Module):
        def __init__(self):
            super().__init__()
            self.emb = torch.nn.Embedding(50, 8, sparse=sparse_module)
            self.out = torch.nn.Linear(8, 1)

        def forward(self, ids, y):
            return (self.out(self.emb(ids).mean(1)) - y).pow(2).mean()

    def run(sparse_cfg):
        torch.manual_seed(0)
        net = Net()
        cfg = {"train_micro_batch_size_per_gpu": 4, "optimizer": {"type": "SGD", "params": {"lr": 0.1}}, "steps_per_print": 1000,
               "sparse_gradients": sparse_cfg}
        engine, _, _, _ = ds.initialize(model=net, config=cfg)
        g = torch.Generator().manual_seed(10 + rank)
        for _ in range(3):
            ids, y = torch.randint(0, 50, (4, 3), generator=g), torch.randn(4, 1, generator=g)
            loss = engine(ids, y)
            engine.backward(loss)
            engine.step()
        return engine, net.emb.weight.detach().clone()

    eng_s, w_sparse = run(True)
    stats = dict(eng_s.sparse_allreduce_stats)
    if sparse_module:
        return {"stats": stats, "w": w_sparse}
    _, w_dense = run(False)
    return {"stats": stats, "w": w_sparse, "w_dense": w_dense}


def moe_checkpoint_worker(rank, world, ckpt_dir, mode):
    """Mixture-of-experts checkpoints: one file per GLOBAL expert, so a run saved with expert parallelism 2 can be loaded
    with expert parallelism 1 (ref: engine.py _save_moe_checkpoint / load_moe_state_dict)."""
    import hcache_deepspeed_b200 as ds
    from hcache_deepspeed_b200.moe.layer import MoE
    from hcache_deepspeed_b200.moe.utils import split_params_into_different_moe_groups_for_optimizer
    ds.init_distributed(dist_backend="gloo")
    torch.manual_seed(0)

    class Net(torch.nn.Module):
        def __init__(self):
            super().__init__()
            self.inp = torch.nn.Linear(8, 16)
            expert = torch.nn.Sequential(torch.nn.Linear(16, 32), torch.nn.ReLU(), torch.nn.Linear(32, 16))
            self.moe = MoE(hidden_size=16, expert=expert, num_experts=4, ep_size=world, k=1, capacity_factor=4.0, min_capacity=8)
            self.out = torch.nn.Linear(16, 1)

        def forward(self, x, y=None):
            h, aux, _ = self.moe(torch.relu(self.inp(x)))
            pred = self.out(h)
            return pred if y is None else (pred - y).pow(2).mean() + 0.01 * aux

    net = Net()
    groups = split_params_into_different_moe_groups_for_optimizer({"params": list(net.parameters()), "name": "all"})
    engine, _, _, _ = ds.initialize(model=net, optimizer=torch.optim.AdamW(groups, lr=1e-2),
                                    config={"train_micro_batch_size_per_gpu": 8, "steps_per_print": 1000})
    assert engine.has_moe_layers and engine.num_experts == [4]
    probe = torch.randn(8, 8, generator=torch.Generator().manual_seed(99))
    if mode == "save":
        g = torch.Generator().manual_seed(3 + rank)
        x, y = torch.randn(8, 8, generator=g), torch.randn(8, 1, generator=g)
        losses = []
        for _ in range(10):
            loss = engine(x, y)
            engine.backward(loss)
            engine.step()
            losses.append(float(loss.detach()))
        engine.save_checkpoint(ckpt_dir, tag="moe")
        engine.eval()
        with torch.no_grad():
            pred = engine(probe)
        experts = {f"{rank * 2 + i}": [p.detach().clone() for p in e.parameters()]
                   for i, e in enumerate(net.moe.deepspeed_moe.experts.deepspeed_experts)}
        # same-degree reload restores the optimizer state of THIS rank's experts too
        for p in engine.module.parameters():
            p.data.add_(1.0)
        engine.load_checkpoint(ckpt_dir, tag="moe")
        with torch.no_grad():
            again = engine(probe)
        st = engine.optimizer.state_dict()["state"]
        return {"losses": losses, "pred": pred, "again": again, "experts": experts, "files": sorted(os.listdir(os.path.join(ckpt_dir, "moe"))),
                "opt_entries": len(st)}
    engine.load_checkpoint(ckpt_dir, tag="moe", load_optimizer_states=False)
    engine.eval()
    with torch.no_grad():
        pred = engine(probe)
    experts = {f"{i}": [p.detach().clone() for p in e.parameters()] for i, e in enumerate(net.moe.deepspeed_moe.experts.deepspeed_experts)}
    return {"pred": pred, "experts": experts}


def pipe_checkpoint_worker(rank, world, ckpt_dir, mode):
    """Pipeline checkpoints are written one file per LAYER: a run saved with 2 stages reloads with 1 stage (and with 2).
    Also covers activation checkpointing on the first stage (token ids in: the embedding must still get its gradient) and
    tied embedding / head weights across stages."""
    import hcache_deepspeed_b200 as ds
    from hcache_deepspeed_b200.pipe import LayerSpec, PipelineModule, TiedLayerSpec
    ds.init_distributed(dist_backend="gloo")

    class Emb(torch.nn.Module):
        def __init__(self, v, h):
            super().__init__()
            self.weight = torch.nn.Parameter(torch.randn(v, h) * 0.1)

        def forward(self, ids):
            return torch.nn.functional.embedding(ids, self.weight)

    class Block(torch.nn.Module):
        def __init__(self, h):
            super().__init__()
            self.l1, self.l2 = torch.nn.Linear(h, 2 * h), torch.nn.Linear(2 * h, h)

        def forward(self, x):
            return x + self.l2(torch.relu(self.l1(x)))

    def head(emb, x):
        return x @ emb.weight.t()

    def loss_fn(logits, labels):
        return torch.nn.functional.cross_entropy(logits.reshape(-1, logits.shape[-1]), labels.reshape(-1))

    torch.manual_seed(0)
    layers = [TiedLayerSpec("emb", Emb, 50, 16)] + [LayerSpec(Block, 16) for _ in range(4)] + [TiedLayerSpec("emb", Emb, 50, 16, forward_fn=head)]
    net = PipelineModule(layers=layers, num_stages=world, loss_fn=loss_fn, partition_method="uniform",
                         activation_checkpoint_interval=1 if mode == "save" else 0)
    cfg = {"train_batch_size": 8, "train_micro_batch_size_per_gpu": 2, "optimizer": {"type": "Adam", "params": {"lr": 1e-2}},
           "steps_per_print": 1000}
    engine, _, _, _ = ds.initialize(model=net, config=cfg, model_parameters=[p for p in net.parameters() if p.requires_grad])
    g = torch.Generator().manual_seed(7)
    data = [(torch.randint(0, 50, (2, 6), generator=g), ) * 2 for _ in range(4)]

    def batches():
        while True:
            for d in data:
                yield d

    if mode == "save":
        emb0 = net.tied_modules["emb"].weight.detach().clone()
        losses = [float(engine.train_batch(data_iter=batches())) for _ in range(10)]
        moved = float((net.tied_modules["emb"].weight.detach() - emb0).abs().max())
        ev = float(engine.eval_batch(batches()))
        engine.save_checkpoint(ckpt_dir, tag="p")
        return {"losses": losses, "eval": ev, "emb_moved": moved, "files": sorted(os.listdir(os.path.join(ckpt_dir, "p"))),
                "emb_sum": float(net.tied_modules["emb"].weight.detach().sum())}
    engine.load_checkpoint(ckpt_dir, tag="p", load_optimizer_states=False)
    return {"eval": float(engine.eval_batch(batches())), "emb_sum": float(net.tied_modules["emb"].weight.detach().sum())}


def zero3_nvme_param_worker(rank, world, device, nvme_dir):
    """ZeRO-Infinity parameter tier: with ``offload_param.device = "nvme"`` partitions above the direct-I/O threshold live in
    files and are staged through pinned buffers for every gather / update; results equal host offload exactly."""
    import hcache_deepspeed_b200 as ds
    from hcache_deepspeed_b200.runtime.zero.partition_parameters import get_param_swapper, set_param_swapper
    ds.init_distributed(dist_backend="gloo")
    set_param_swapper(None)
    torch.manual_seed(0)

    class Net(torch.nn.Module):
        def __init__(self):
            super().__init__()
            self.emb = torch.nn.Embedding(6000, 128)            # 3 MB fp32: 1.5 MB per rank, above the 1 MB threshold
            self.mid = torch.nn.Linear(128, 128)                 # small: stays in host memory
            self.out = torch.nn.Linear(128, 6000, bias=False)

        def forward(self, ids):
            return torch.nn.functional.cross_entropy(self.out(torch.tanh(self.mid(self.emb(ids)))).reshape(-1, 6000), ids.reshape(-1))

    off = {"device": device, "pin_memory": False}
    if device == "nvme":
        off.update(nvme_path=os.path.join(nvme_dir, f"r{rank}"), buffer_count=4, buffer_size=1_000_000)
    cfg = {"train_micro_batch_size_per_gpu": 4, "optimizer": {"type": "AdamW", "params": {"lr": 1e-2}}, "steps_per_print": 1000,
           "zero_optimization": {"stage": 3, "stage3_param_persistence_threshold": 0, "offload_param": off},
           "aio": {"block_size": 65536, "queue_depth": 4}}
    net = Net()
    engine, _, _, _ = ds.initialize(model=net, config=cfg)
    g = torch.Generator().manual_seed(5 + rank)
    ids = torch.randint(0, 6000, (4, 12), generator=g)
    losses = []
    for _ in range(4):
        loss = engine(ids)
        engine.backward(loss)
        engine.step()
        losses.append(float(loss.detach()))
    out = {"losses": losses, "nvme_params": [n for n, p in net.named_parameters() if getattr(p, "ds_nvme", False)]}
    if device == "nvme":
        sw = get_param_swapper()
        files = sorted(os.listdir(sw.swap_folder))
        out.update(files=files, free_buffers=sw.available_swap_in_buffers(), host_slices=[int(p.ds_tensor.numel()) for p in net.parameters()
                                                                                           if getattr(p, "ds_nvme", False)])
        engine.save_checkpoint(os.path.join(nvme_dir, "ck"), tag="t")
        engine.load_checkpoint(os.path.join(nvme_dir, "ck"), tag="t")
        out["after_reload"] = float(engine(ids).detach())
        set_param_swapper(None)
    return out


def zeropp_worker(rank, world, qw, qg, loco=None):
    """ZeRO++ on one node: int8 weight all-gather (qwZ) and quantised gradient reduce via all-to-all (qgZ). Training tracks
    the unquantised run closely; traffic accounting shows the saving."""
    import hcache_deepspeed_b200 as ds
    ds.init_distributed(dist_backend="gloo")
    torch.manual_seed(0)
    model = torch.nn.Sequential(torch.nn.Linear(64, 256), torch.nn.Tanh(), torch.nn.Linear(256, 64), torch.nn.Tanh(), torch.nn.Linear(64, 1))
    cfg = {"train_micro_batch_size_per_gpu": 8, "optimizer": {"type": "AdamW", "params": {"lr": 3e-3}}, "steps_per_print": 1000,
           "zero_optimization": {"stage": 3, "stage3_param_persistence_threshold": 0, "zero_quantized_weights": qw,
                                 "zero_quantized_gradients": qg}}
    if loco is not None:
        cfg["zero_optimization"]["zeropp_loco_param"] = loco
    engine, _, _, _ = ds.initialize(model=model, config=cfg)
    g = torch.Generator().manual_seed(5 + rank)
    x = torch.randn(8, 64, generator=g)
    y = x[:, :1] * 2.0 - x[:, 1:2]
    losses = []
    for _ in range(40):
        loss = (engine(x) - y).pow(2).mean()
        engine.backward(loss)
        engine.step()
        losses.append(float(loss.detach()))
    flagged = sum(1 for p in model.parameters() if getattr(p, "ds_quantized_gather", False))
    out = {"losses": losses, "flagged": flagged, "qgz": dict(engine.optimizer.qgz_stats)}
    if loco is not None:
        from hcache_deepspeed_b200.runtime.comm.coalesced_collectives import all_to_all_loco_quant_reduce
        out["loco_buffers"] = len(engine.optimizer._loco_err)
        out["loco_ages"] = sorted({age for _, age in engine.optimizer._loco_err.values()})
        # the standalone form: over repeated calls on the same gradient, the error fed back makes the running mean of the
        # results approach the true mean more closely than a single quantised exchange does
        p = torch.nn.Parameter(torch.zeros(64, 32))
        gen = torch.Generator().manual_seed(100 + rank)
        p.grad = torch.randn(64, 32, generator=gen)
        grads = [torch.randn(64, 32, generator=torch.Generator().manual_seed(100 + r)) for r in range(world)]
        want = (sum(grads) / world).reshape(-1).chunk(world)[rank]
        outs = [all_to_all_loco_quant_reduce([p], {"local_0": None}, {"err_beta": 1.0, "reset_T": 1000})[0][:want.numel()] for _ in range(8)]
        out["one_shot_err"] = float((outs[0] - want).abs().mean())
        out["fed_back_err"] = float((sum(outs) / len(outs) - want).abs().mean())
    return out


def partition_activations_worker(rank, world):
    """``checkpointing.configure(mpu, partition_activations=True)``: each model-parallel rank keeps 1/mp of every
    checkpointed input and all-gathers it for the recomputation; gradients equal the plain run (ref:
    activation_checkpointing/checkpointing.py partition_activations / gather_partitioned_activations)."""
    import hcache_deepspeed_b200 as ds
    from hcache_deepspeed_b200.runtime.activation_checkpointing import checkpointing
    ds.init_distributed(dist_backend="gloo")

    class MPU:
        def get_model_parallel_group(self):
            return torch.distributed.group.WORLD

        def get_model_parallel_world_size(self):
            return world

        def get_model_parallel_rank(self):
            return rank

        def get_data_parallel_group(self):
            return None

        def get_data_parallel_world_size(self):
            return 1

        def get_data_parallel_rank(self):
            return 0

    torch.manual_seed(0)                      # replicated weights and inputs, as in a tensor-parallel group
    net = torch.nn.Sequential(torch.nn.Linear(16, 16), torch.nn.Tanh(), torch.nn.Linear(16, 16))
    x = torch.randn(6, 16, requires_grad=True)
    y = net(x).pow(2).sum()
    y.backward()
    want = [p.grad.clone() for p in net.parameters()] + [x.grad.clone()]
    out = {}
    for cpu in (False, True):
        for p in net.parameters():
            p.grad = None
        x.grad = None
        checkpointing.configure(MPU(), partition_activations=True, checkpoint_in_cpu=cpu)
        saved = []
        orig = checkpointing.partition_activation

        def spy(t, _orig=orig):
            part = _orig(t)
            saved.append((t.numel(), part.numel()))
            return part

        checkpointing.partition_activation = spy
        try:
            y2 = checkpointing.checkpoint(net, x).pow(2).sum()
            y2.backward()
        finally:
            checkpointing.partition_activation = orig
        got = [p.grad.clone() for p in net.parameters()] + [x.grad.clone()]
        out[f"cpu={cpu}"] = {"err": max(float((a - b).abs().max()) for a, b in zip(got, want)), "loss": float((y - y2).abs()),
                             "saved": saved}
    checkpointing.configure(None, partition_activations=False, checkpoint_in_cpu=False)
    checkpointing.reset()
    return out


def safe_accessors_worker(rank, world, stage):
    """``deepspeed.utils.safe_get_* / safe_set_*`` under ZeRO 1 / 2 / 3: full fp32 parameter, gradient (between backward
    and step) and Adam state of a partitioned parameter, setters that every rank sees afterwards, and the partition-local
    variants of ZeRO-3 (ref: utils/tensor_fragment.py:118-330)."""
    import hcache_deepspeed_b200 as ds
    from hcache_deepspeed_b200 import utils as U
    ds.init_distributed(dist_backend="gloo")
    torch.manual_seed(0)
    model = TinyNet(32)
    ref = TinyNet(32)
    ref.load_state_dict(model.state_dict())
    cfg = {"train_micro_batch_size_per_gpu": 4, "optimizer": {"type": "Adam", "params": {"lr": 1e-2}}, "steps_per_print": 1000,
           "zero_optimization": {"stage": stage, "stage3_param_persistence_threshold": 0}}
    engine, _, _, _ = ds.initialize(model=model, config=cfg)
    opt_ref = torch.optim.Adam(ref.parameters(), lr=1e-2)
    x, y = next(iter(make_batches(1, 4 * world, 32)))
    lo = rank * 4
    # reference: the averaged gradient of the global batch
    (sum(ref(x[r * 4:r * 4 + 4], y[r * 4:r * 4 + 4]) for r in range(world)) / world).backward()
    want_grad = {n: p.grad.clone() for n, p in ref.named_parameters()}
    loss = engine(x[lo:lo + 4], y[lo:lo + 4])
    engine.backward(loss)
    out = {}
    def is_matrix(q):
        return len(getattr(q, "ds_shape", q.shape)) == 2

    name, p = next((n, q) for n, q in engine.module.named_parameters() if is_matrix(q))
    g = U.safe_get_full_grad(p)
    out["grad_err"] = float((g - want_grad[name]).abs().max())
    U.safe_set_full_grad(p, torch.zeros_like(want_grad[name]))          # freeze this parameter for one step
    before = U.safe_get_full_fp32_param(p).clone()
    engine.step()
    after = U.safe_get_full_fp32_param(p)
    out["frozen_by_zero_grad"] = float((after - before).abs().max())
    other_name, other = next((n, q) for n, q in engine.module.named_parameters() if is_matrix(q) and q is not p)
    opt_ref.step()
    out["other_param_err"] = float((U.safe_get_full_fp32_param(other) - dict(ref.named_parameters())[other_name].detach()).abs().max())
    ea = U.safe_get_full_optimizer_state(other, "exp_avg")
    out["exp_avg_err"] = float((ea - opt_ref.state[dict(ref.named_parameters())[other_name]]["exp_avg"]).abs().max())
    # setters: a new value set collectively is what every rank reads back, and what the next forward uses
    new = torch.full(tuple(after.shape), 0.25)
    U.safe_set_full_fp32_param(p, new)
    U.safe_set_full_optimizer_state(p, torch.ones_like(new), "exp_avg")
    out["set_param"] = float((U.safe_get_full_fp32_param(p) - new).abs().max())
    out["set_state"] = float((U.safe_get_full_optimizer_state(p, "exp_avg") - 1.0).abs().max())
    if stage == 3:
        from hcache_deepspeed_b200.runtime.zero import GatheredParameters
        with GatheredParameters([p]):
            out["lp_follows"] = float((p.data.float() - 0.25).abs().max())
        loc = U.safe_get_local_fp32_param(p)
        out["local_numel"] = (int(loc.numel()), int(p.ds_partition_numel))
        U.safe_set_local_fp32_param(p, torch.full_like(loc, float(rank + 1)))
        full = U.safe_get_full_fp32_param(p).reshape(-1)
        out["local_set"] = (float(full[0]), float(full[-1]))
        out["local_state"] = int(U.safe_get_local_optimizer_state(p, "exp_avg").numel())
    else:
        out["lp_follows"] = float((p.data.float() - 0.25).abs().max())
    return out


def engine_accessors_worker(rank, world, stage):
    """Public helpers of the engine that other frameworks call directly: bucketed all-reduce, sparse all-reduce, scalar
    all-gather, state offload / reload, partition-cache release, module-state loading (ref: engine.py:2560-2720, 3943)."""
    import hcache_deepspeed_b200 as ds
    from hcache_deepspeed_b200.runtime.sparse_tensor import SparseTensor
    from hcache_deepspeed_b200.runtime.zero.offload_states import OffloadStateTypeEnum, offloaded_state_kinds
    ds.init_distributed(dist_backend="gloo")
    torch.manual_seed(0)
    model = TinyNet(32)
    cfg = {"train_micro_batch_size_per_gpu": 4, "optimizer": {"type": "Adam", "params": {"lr": 1e-2}}, "steps_per_print": 1000,
           "zero_optimization": {"stage": stage, "stage3_param_persistence_threshold": 0}}
    engine, _, _, _ = ds.initialize(model=model, config=cfg)
    group = engine.data_parallel_group
    out = {}
    # bucket helpers
    a, b = torch.full((3, ), float(rank + 1)), torch.full((2, 2), float(10 * (rank + 1)))
    engine.allreduce_no_retain([a, b], group, numel_per_bucket=3)
    mean = sum(range(1, world + 1)) / world
    out["bucket"] = (float(a[0]), float(b[0, 0]), mean)
    out["scalars"] = [int(t) for t in engine.all_gather_scalar(torch.tensor([rank + 5]), group)]
    dense = torch.zeros(6, 2)
    dense[rank] = 1.0 + rank
    st = engine.sparse_allreduce(SparseTensor(dense), group)
    out["sparse_rows"] = sorted(st.indices.tolist())
    out["sparse_dense"] = st.to_dense()[:world, 0].tolist()
    # one training step, then offload everything, reload, and train on: same result as without the round trip
    x, y = next(iter(make_batches(1, 4 * world, 32)))
    lo = rank * 4

    def step(e):
        loss = e(x[lo:lo + 4], y[lo:lo + 4])
        e.backward(loss)
        e.step()
        return float(loss)

    step(engine)
    engine.offload_states(pin_memory=False)
    out["kinds"] = sorted(k.value for k in offloaded_state_kinds(engine))
    engine.reload_states()
    out["kinds_after"] = len(offloaded_state_kinds(engine))
    engine.offload_states(include=[OffloadStateTypeEnum.optim_states, OffloadStateTypeEnum.hp_params], pin_memory=False)
    engine.reload_states()
    step(engine)
    l_round_trip = step(engine)        # the second step after the round trip sees the first one's update
    torch.manual_seed(0)
    model2 = TinyNet(32)
    engine2, _, _, _ = ds.initialize(model=model2, config=cfg)
    step(engine2)
    step(engine2)
    out["loss_pair"] = (l_round_trip, step(engine2))
    # partition cache + module state loading
    if stage == 3:
        from hcache_deepspeed_b200.runtime.zero import GatheredParameters, ZeroParamStatus
        from hcache_deepspeed_b200.runtime.zero.partition_parameters import all_gather_params
        ps = list(engine.module.parameters())
        all_gather_params(ps)
        engine.empty_partition_cache()
        out["released"] = all(p.ds_status == ZeroParamStatus.NOT_AVAILABLE for p in ps)
        sd = {k: torch.full_like(v, 0.5) for k, v in TinyNet(32).state_dict().items()}
        engine.load_module_state_dict({"module": sd}, fetch_z3_params=True)
        with GatheredParameters(ps):
            out["loaded"] = all(float((p.data - 0.5).abs().max()) == 0.0 for p in ps)
    else:
        sd = {k: torch.full_like(v, 0.5) for k, v in engine.module.state_dict().items()}
        engine.load_module_state_dict({"module": sd})
        out["loaded"] = all(float((p.data.float() - 0.5).abs().max()) == 0.0 for p in engine.module.parameters())
        out["released"] = True
    if stage >= 1:
        # the fp32 masters were refreshed by the load too (else the next step would bring the old weights back)
        opt = engine.optimizer
        full = opt.get_full_hp_param(next(iter(engine.module.parameters())))
        out["master_follows_load"] = float((full - 0.5).abs().max())
        # optimizer-level public methods
        opt.set_lr(0.125)
        out["lr"] = (opt.get_lr(), engine.get_lr()[0])
        loss = engine(x[lo:lo + 4], y[lo:lo + 4])
        engine.backward(loss)
        if stage < 2:
            opt.reduce_gradients()
        grads = opt._api_local_grads()
        norm = opt.get_grad_norm_direct(grads)
        out["overflow"] = opt.check_overflow()
        p0 = next(iter(engine.module.parameters()))
        g_full = opt.get_fp32_grad_for_param(p0)
        out["grad_shape_ok"] = tuple(g_full.shape) == tuple(getattr(p0, "ds_shape", p0.shape))
        engine.step()
        out["norm_pair"] = (norm / opt.loss_scale, engine.get_global_grad_norm())
        parts = opt.get_data_parallel_partitions(torch.arange(7))
        out["parts"] = [t.tolist() for t in parts]
        opt.override_loss_scale(4.0)
        out["scale"] = opt.loss_scale
        inside, outside, first = opt.get_partition_info([torch.zeros(3), torch.zeros(4), torch.zeros(5)], 6, 1)
        out["pinfo"] = ([t.numel() for t in inside], [t.numel() for t in outside], first)
        lean = opt.get_lean_optimizer_state()
        out["lean_keys"] = sorted(k for k in lean[0] if k.startswith("exp_avg"))
    return out


def pipe_api_worker(rank, world):
    """Smaller public methods of the pipeline engine / module: named side losses, changing the batch size and the
    activation shape between batches, per-layer file listing, loading layer files through ``load_module_state_dict``."""
    import tempfile
    import hcache_deepspeed_b200 as ds
    from hcache_deepspeed_b200.runtime.pipe import LayerSpec, PipelineModule
    ds.init_distributed(dist_backend="gloo")
    hidden, micro = 16, 4

    class Net(PipelineModule):
        aux = None

        def get_additional_losses(self):
            return {"l1": Net.aux if Net.aux is not None else torch.zeros(())}

    def loss_fn(out, y):
        Net.aux = out.detach().abs().mean()
        return torch.nn.functional.mse_loss(out.squeeze(-1), y)

    torch.manual_seed(7)
    specs = [LayerSpec(_Stage, hidden) for _ in range(3)] + [LayerSpec(torch.nn.Linear, hidden, 1)]
    net = Net(layers=specs, num_stages=2, loss_fn=loss_fn, partition_method="uniform", seed_layers=True, base_seed=100)
    cfg = {"train_micro_batch_size_per_gpu": micro, "gradient_accumulation_steps": 2,
           "optimizer": {"type": "Adam", "params": {"lr": 1e-2, "torch_adam": True}}, "steps_per_print": 1000}
    engine, _, _, _ = ds.initialize(model=net, config=cfg)
    out = {"stage": engine.stage_id}

    def batches(n, rows=micro, seed=0):
        g = torch.Generator().manual_seed(seed)
        return iter([(torch.randn(rows, hidden, generator=g), torch.randn(rows, generator=g)) for _ in range(n)])

    engine.train_batch(batches(2))
    extra = engine.get_additional_losses()
    out["extra"] = {k: float(v) for k, v in extra.items()}
    engine.set_train_batch_size(micro * 3)                     # 3 micro-batches now
    out["micro_batches"] = (engine.micro_batches, engine.gradient_accumulation_steps())
    engine.train_batch(batches(3, seed=1))
    engine.reset_activation_shape()                            # other row count per micro-batch from here on
    out["loss_new_shape"] = float(engine.train_batch(batches(3, rows=2, seed=2)))
    net.set_checkpoint_interval(1)
    out["interval"] = net.checkpoint_interval
    out["loss_ckpt"] = float(engine.train_batch(batches(3, rows=2, seed=3)))
    # per-layer files
    d = os.environ.get("HDS_TEST_PIPE_DIR")
    engine.save_checkpoint(d, tag="t")
    ds.comm.barrier()
    out["files"] = [os.path.basename(p) for p in net.ckpt_layer_path_list(os.path.join(d, "t"), 0)]
    before = [p.detach().clone() for p in net.parameters()]
    with torch.no_grad():
        for p in net.parameters():
            p.zero_()
    engine.load_module_state_dict({"module": os.path.join(d, "t")})
    out["reloaded"] = all(torch.equal(a, b) for a, b in zip(before, net.parameters()))
    engine.tput_log("pipeline api worker done")
    engine.log_for_device("stage", engine.stage_id)
    out["mem"] = engine.mem_status("end")
    return out


def runtime_utils_worker(rank, world):
    """``PartitionedTensor`` round trips (even / uneven split, rebuilt from its wire form), per-group ZeRO all-gathers and
    the mpu-style names of ``utils.groups`` (ref: runtime/utils.py:624-720, 961-1008; utils/groups.py:60-130)."""
    import hcache_deepspeed_b200 as ds
    from hcache_deepspeed_b200.runtime.utils import PartitionedTensor, all_gather_dp_groups, all_gather_into_tensor_dp_groups
    from hcache_deepspeed_b200.utils import groups
    ds.init_distributed(dist_backend="gloo")
    out = {}
    for name, shape in (("even", (4, 6)), ("uneven", (3, 5))):
        full = torch.arange(shape[0] * shape[1], dtype=torch.float32).reshape(shape) + 1
        pt = PartitionedTensor(full, group=None)
        meta, local = pt.to_meta(), pt.data()
        rebuilt = PartitionedTensor.from_meta(meta, local, group=None, device="cpu")
        out[name] = (bool(torch.equal(pt.full(), full)), bool(torch.equal(rebuilt.full(), full)), int(local.numel()),
                     list(rebuilt.full_size()))
    flat = torch.zeros(8)
    parts = list(flat.chunk(world))
    parts[rank].fill_(float(rank + 1))
    all_gather_into_tensor_dp_groups([flat], [parts], [None])
    out["gathered"] = flat.tolist()
    flat2 = torch.zeros(8)
    parts2 = list(flat2.chunk(world))
    parts2[rank].copy_(torch.arange(4.0) + 10 * rank)
    all_gather_dp_groups([flat2], [parts2], [None], start_alignment_factor=1, allgather_bucket_size=4)
    out["gathered_sharded"] = flat2.tolist()
    out["mpu"] = (groups.get_model_parallel_world_size(), groups.get_model_parallel_rank(), groups.get_data_parallel_world_size(),
                  groups.get_data_parallel_rank(), groups.get_tensor_model_parallel_src_rank())
    groups.set_tensor_model_parallel_world_size(2)
    groups.set_tensor_model_parallel_rank(rank)
    out["mpu_pinned"] = (groups.get_tensor_model_parallel_world_size(), groups.get_tensor_model_parallel_rank(),
                         groups.get_tensor_model_parallel_src_rank())
    groups.set_tensor_model_parallel_world_size(None)
    groups.set_tensor_model_parallel_rank(None)
    objs = [{"a": rank}] if rank == 0 else [None]
    ds.comm.broadcast_object_list(objs, src=0)
    out["obj"] = objs[0]
    return out


def mics_hierarchy_worker(rank, world):
    """MiCS groups for a shard group spanning two 'nodes' of two ranks, and the two-hop all-gather in rank order."""
    os.environ["HDS_MICS_NODE_SIZE"] = "2"
    import hcache_deepspeed_b200 as ds
    from hcache_deepspeed_b200.runtime.zero import mics
    ds.init_distributed(dist_backend="gloo")
    g = mics.create_mics_comm_groups(4, hierarchical_allgather=True)
    flat = mics.create_mics_comm_groups(4, hierarchical_allgather=False)
    local = torch.tensor([10.0 * rank, 10.0 * rank + 1])
    two_hop = mics.hierarchical_all_gather(local, g)
    one_hop = mics.hierarchical_all_gather(local, flat)
    return {"hier": mics.has_hierarchical_all_gather_groups(g), "flat": mics.has_hierarchical_all_gather_groups(flat),
            "two_hop": two_hop.tolist(), "one_hop": one_hop.tolist(),
            "intra": ds.comm.get_world_size(group=g.param_intra_node_group), "inter": ds.comm.get_world_size(group=g.param_inter_node_shard_group)}
