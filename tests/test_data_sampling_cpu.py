"""Indexed datasets, the offline data analyzer and token-budget batching (reference tests:
tests/unit/runtime/test_data_efficiency.py, data_sampling/data_analyzer.py::test_compare_both_data_analyzers)."""
import numpy as np
import pytest
import torch

from hcache_deepspeed_b200.runtime.data_pipeline.data_sampling import (DataAnalyzer, DistributedDataAnalyzer, MMapIndexedDataset,
                                                                        MMapIndexedDatasetBuilder, batch_by_seqlens,
                                                                        get_dataloader_and_lr_scheduler_for_variable_batch_size,
                                                                        load_sample_to_metric, scale_lr)
from hcache_deepspeed_b200.runtime.data_pipeline.data_sampling import indexed_dataset as ixd
from hcache_deepspeed_b200.runtime.data_pipeline.data_sampling.utils import find_fit_int_dtype, split_index


def test_mmap_indexed_dataset_roundtrip(tmp_path):
    rng = np.random.default_rng(0)
    docs = [rng.integers(0, 50000, size=n) for n in (5, 1, 17, 0, 9)]
    prefix = str(tmp_path / "ds")
    b = ixd.make_builder(ixd.data_file_path(prefix), "mmap", vocab_size=50000)
    for d in docs:
        b.add_item(torch.from_numpy(d))
        b.end_document()
    b.finalize(ixd.index_file_path(prefix))
    assert ixd.infer_dataset_impl(prefix) == "mmap" and ixd.dataset_exists(prefix)
    ds = ixd.make_dataset(prefix, "infer")
    assert len(ds) == 5 and ds.dtype == np.uint16 and ds.sizes.tolist() == [5, 1, 17, 0, 9]
    assert ds.doc_idx.tolist() == [0, 1, 2, 3, 4, 5]
    for i, d in enumerate(docs):
        assert ds[i].tolist() == d.tolist()
    assert ds.get(2, offset=3, length=4).tolist() == docs[2][3:7].tolist()
    assert [x.tolist() for x in ds[1:3]] == [docs[1].tolist(), docs[2].tolist()]
    # header layout is the Megatron one
    raw = open(ixd.index_file_path(prefix), "rb").read()
    assert raw[:9] == b"MMIDIDX\x00\x00" and raw[9:17] == (1).to_bytes(8, "little") and raw[17] == 6
    # merge
    b2 = MMapIndexedDatasetBuilder(ixd.data_file_path(prefix + "2"), dtype=np.uint16)
    b2.add_item_numpy(np.array([7, 8]))
    b2.end_document()
    b2.merge_file_(prefix)
    b2.finalize(ixd.index_file_path(prefix + "2"))
    m = MMapIndexedDataset(prefix + "2")
    assert len(m) == 6 and m[0].tolist() == [7, 8] and m[3].tolist() == docs[2].tolist() and m.doc_idx.tolist() == list(range(7))
    import pickle
    assert len(pickle.loads(pickle.dumps(m))) == 6


class _Seqs(torch.utils.data.Dataset):
    def __init__(self, n=203, seed=0):
        self.lens = np.random.default_rng(seed).integers(1, 40, size=n)

    def __len__(self):
        return len(self.lens)

    def __getitem__(self, i):
        return torch.full((int(self.lens[i]), ), i % 7)


def _metrics():
    names = ["seqlen", "vocab"]
    fns = [lambda batch: torch.tensor([len(s) for s in batch]),
           lambda batch: torch.bincount(torch.cat(batch), minlength=7)]
    return names, fns, ["single_value_per_sample", "accumulate_value_over_samples"]


@pytest.mark.parametrize("workers", [1, 3])
def test_data_analyzer_map_reduce(tmp_path, workers):
    ds = _Seqs()
    names, fns, types_ = _metrics()
    for w in range(workers):
        DataAnalyzer(ds, num_workers=workers, worker_id=w, batch_size=16, metric_names=names, metric_functions=fns,
                     metric_types=types_, save_path=str(tmp_path), collate_fn=list).run_map()
    DataAnalyzer(ds, num_workers=workers, worker_id=0, metric_names=names, metric_functions=fns, metric_types=types_,
                 save_path=str(tmp_path)).run_reduce()
    assert load_sample_to_metric(str(tmp_path), "seqlen").tolist() == ds.lens.tolist()
    i2m = MMapIndexedDataset(f"{tmp_path}/seqlen/seqlen_index_to_metric")
    i2s = MMapIndexedDataset(f"{tmp_path}/seqlen/seqlen_index_to_sample")
    uniq = sorted(set(ds.lens.tolist()))
    assert [int(i2m[j][0]) for j in range(len(i2m))] == uniq
    for j, v in enumerate(uniq):
        assert sorted(i2s[j].tolist()) == np.nonzero(ds.lens == v)[0].tolist()
    merged = MMapIndexedDataset(f"{tmp_path}/seqlen/seqlen_index_to_sample_percentile_merged")
    allidx = np.concatenate([merged[j] for j in range(len(merged))])
    assert sorted(allidx.tolist()) == list(range(len(ds))) and (np.diff(ds.lens[allidx]) >= 0).all()
    vocab = MMapIndexedDataset(f"{tmp_path}/vocab/vocab_metric_value")[0]
    want = np.zeros(7, dtype=np.int64)
    for i, n in enumerate(ds.lens):
        want[i % 7] += n
    assert vocab.tolist() == want.tolist()


def test_distributed_data_analyzer_single_process(tmp_path):
    ds = _Seqs(50)
    names, fns, types_ = _metrics()
    DistributedDataAnalyzer(ds, batch_size=8, metric_names=names, metric_functions=fns, metric_types=types_,
                            save_path=str(tmp_path), collate_fn=list).run_map_reduce()
    assert load_sample_to_metric(str(tmp_path), "seqlen").tolist() == ds.lens.tolist()


def test_small_helpers():
    assert find_fit_int_dtype(0, 255) == np.uint8 and find_fit_int_dtype(0, 256) == np.uint16
    assert find_fit_int_dtype(-1, 100) == np.int8 and find_fit_int_dtype(-1, 40000) == np.int32
    assert split_index(0, 10, 3) == [(0, 3), (3, 6), (6, 10)]
    assert scale_lr(8, 16, 0.1) == pytest.approx(0.2) and scale_lr(8, 32, 0.1, "sqrt") == pytest.approx(0.2)
    assert scale_lr(8, 32, 0.1, None) == 0.1


@pytest.mark.parametrize("order", ["dataloader", "seqlen", "random"])
@pytest.mark.parametrize("same_size", [False, True])
def test_batch_by_seqlens(order, same_size):
    ds = _Seqs(300, seed=3)
    ebs = 4
    mb_ids, batch_sizes, max_lens = batch_by_seqlens(ds.lens, max_tokens=100, min_batch_size=1, max_batch_size=8,
                                                     sequence_picking_order=order, effective_batch_size=ebs,
                                                     required_microbatches_of_same_size=same_size, seed=1)
    assert len(mb_ids) == ebs * len(batch_sizes) == ebs * len(max_lens)
    seen = []
    for k, (batch_id, ids) in enumerate(mb_ids):
        assert batch_id == k // ebs and 1 <= len(ids) <= 8 and ds.lens[ids].sum() <= 100
        seen += ids
    assert len(seen) == len(set(seen))                      # nothing is used twice
    assert len(seen) >= 0.9 * len(ds)                       # and the packing is not wasteful
    for b in range(len(batch_sizes)):
        mbs = [ids for bid, ids in mb_ids if bid == b]
        assert sum(len(m) for m in mbs) == batch_sizes[b] and max(ds.lens[i] for m in mbs for i in m) == max_lens[b]
        if same_size:
            assert len({len(m) for m in mbs}) == 1


def test_variable_batch_dataloader_and_lr():
    ds = _Seqs(120, seed=5)
    model = torch.nn.Linear(2, 2)
    opt = torch.optim.SGD(model.parameters(), lr=0.1)
    sched = torch.optim.lr_scheduler.StepLR(opt, step_size=1, gamma=0.5)
    seen = []
    for rank in range(2):
        loader, lr_sched, io = get_dataloader_and_lr_scheduler_for_variable_batch_size(
            ds, ds.lens, max_tokens=64, base_batch_size=8, lr_scheduler_or_optimizer=sched if rank == 0 else None,
            effective_batch_size=2, dataloader_rank=rank, dataloader_num_replicas=2, dataloader_collate_fn=lambda b: b)
        for batch in loader:
            assert sum(len(s) for s in batch) <= 64
            seen.append(len(batch))
        if rank == 0:
            sizes = lr_sched.batch_sizes
            assert opt.param_groups[0]["lr"] == pytest.approx(0.1 * sizes[0] / 8)
            for step in range(1, 4):
                opt.step()
                lr_sched.step()
                assert opt.param_groups[0]["lr"] == pytest.approx(0.1 * 0.5 ** step * sizes[step % len(sizes)] / 8)
            sd = lr_sched.state_dict()
            lr_sched.load_state_dict(sd)
            assert lr_sched.last_epoch == 3
    assert sum(seen) >= 0.9 * 120 - 8


def test_sampler_reads_analyzer_indexes(tmp_path):
    from hcache_deepspeed_b200.runtime.data_pipeline.data_sampler import DeepSpeedDataSampler
    ds = _Seqs(200, seed=9)
    DataAnalyzer(ds, batch_size=32, metric_names=["seqlen"], metric_functions=[_metrics()[1][0]],
                 metric_types=["single_value_per_sample"], save_path=str(tmp_path), collate_fn=list).run_map_reduce()
    cfg = {"data_sampling": {"curriculum_learning": {"enabled": True, "curriculum_metrics": {"seqlen": {
        "index_to_sample_path": f"{tmp_path}/seqlen/seqlen_index_to_sample",
        "index_to_metric_path": f"{tmp_path}/seqlen/seqlen_index_to_metric",
        "min_difficulty": 10, "max_difficulty": 40, "schedule_type": "fixed_linear",
        "schedule_config": {"total_curriculum_step": 10, "difficulty_step": 1}}}}}}
    s = DeepSpeedDataSampler(cfg, len(ds), micro_batch_size=4, data_parallel_rank=0, data_parallel_size=2)
    assert s.difficulties["seqlen"].tolist() == ds.lens.tolist()
    first = s.get_next_global_batch()
    assert ds.lens[first].max() <= 13          # early batches only draw easy (short) samples


def test_convert_to_random_ltd_and_clean_state_dict():
    from hcache_deepspeed_b200.runtime.data_pipeline.data_routing import (RandomLayerTokenDrop, convert_to_random_ltd,
                                                                        save_without_random_ltd)
    model = torch.nn.Sequential(torch.nn.Linear(4, 4), torch.nn.Sequential(torch.nn.Linear(4, 4), torch.nn.ReLU()))
    keys = list(model.state_dict())
    convert_to_random_ltd(model, torch.nn.Linear)
    assert isinstance(model[0], RandomLayerTokenDrop) and isinstance(model[1][0], RandomLayerTokenDrop)
    assert model[0].layer_id == 0 and model[1][0].layer_id == 1
    assert any("random_ltd_layer" in k for k in model.state_dict())
    assert list(save_without_random_ltd(model)) == keys
    assert model(torch.randn(2, 3, 4)).shape == (2, 3, 4)


def test_data_efficiency_config_tree_and_routing_helpers():
    import torch

    from hcache_deepspeed_b200.runtime.data_pipeline.config import (get_curriculum_params_legacy, get_data_efficiency_config,
                                                                    get_random_ltd_enabled)
    from hcache_deepspeed_b200.runtime.data_pipeline.data_routing.scheduler import BaseScheduler
    from hcache_deepspeed_b200.runtime.data_pipeline.data_routing.utils import bsh_decoder_gather, bsh_decoder_scatter
    c = get_data_efficiency_config({"data_efficiency": {"enabled": True, "data_routing": {"enabled": True, "random_ltd": {"enabled": True}}}})
    assert c["enabled"] and c["seed"] == 1234 and c["data_sampling"]["num_epochs"] == 1000
    assert c["data_sampling"]["curriculum_learning"]["enabled"] is False and c["data_routing"]["random_ltd"]["enabled"]
    assert get_random_ltd_enabled({"data_efficiency": {"data_routing": {"random_ltd": {"enabled": True}}}})
    assert get_curriculum_params_legacy({}) is False
    assert get_curriculum_params_legacy({"curriculum_learning": {"enabled": True, "min_difficulty": 8}}) == {"min_difficulty": 8}
    h = torch.randn(2, 10, 4)
    part, idx, _ = bsh_decoder_gather(6, h, None)
    assert part.shape == (2, 6, 4) and bool((idx[:, 1:] > idx[:, :-1]).all())
    out = bsh_decoder_scatter(h, part * 0, idx)
    assert int((out.abs().sum(-1) == 0).sum()) == 12
    b = BaseScheduler()
    b.state = {"schedule_type": "fixed_linear", "min_value": 16, "max_value": 128, "schedule_config": {"require_steps": 100, "seq_per_step": 16}}
    assert [b.get_value(s) for s in (0, 50, 1000)] == [16, 64, 128]


def test_engine_builds_curriculum_sampler_from_data_efficiency_config(tmp_path):
    """``initialize(training_data=...)`` with ``data_efficiency.data_sampling.curriculum_learning``: the engine's loader
    feeds batches whose difficulty follows the schedule, and the sampler state rides in checkpoints (ref: engine.py
    deepspeed_io, data_sampler.py state_dict)."""
    import numpy as np
    import torch
    import hcache_deepspeed_b200 as ds

    class DS(torch.utils.data.Dataset):
        def __len__(self):
            return 120

        def __getitem__(self, i):
            return torch.full((4, ), float(i % 10)), torch.tensor([float(i % 10)])

    difficulty = np.arange(120) % 10            # the feature value IS the difficulty
    cfg = {"train_micro_batch_size_per_gpu": 4, "optimizer": {"type": "SGD", "params": {"lr": 0.0}}, "steps_per_print": 1000,
           "data_efficiency": {"enabled": True, "seed": 7, "data_sampling": {"enabled": True, "num_epochs": 1, "curriculum_learning": {
               "enabled": True, "curriculum_metrics": {"len": {"min_difficulty": 1, "max_difficulty": 9, "schedule_type": "fixed_linear",
                                                               "schedule_config": {"total_curriculum_step": 8, "difficulty_step": 1}}}}}}}
    model = torch.nn.Linear(4, 1)
    engine, _, loader, _ = ds.initialize(model=model, config=cfg, training_data=DS())
    engine.training_data_sampler.difficulties["len"] = difficulty
    it = iter(loader)
    maxima = []
    for step in range(8):
        x, y = next(it)
        assert x.shape == (4, 4)
        maxima.append(float(x.max()))
        loss = (engine(x) - y).pow(2).mean()
        engine.backward(loss)
        engine.step()
        if step == 3:
            engine.save_checkpoint(str(tmp_path), tag="c")
            saved = dict(engine.training_data_sampler.state_dict())
    assert maxima[0] <= 2 and maxima[-1] >= 5 and all(m <= 9 for m in maxima)          # easy samples first, harder later
    assert maxima == sorted(maxima) or max(maxima[:3]) < max(maxima[5:])
    engine2, _, loader2, _ = ds.initialize(model=torch.nn.Linear(4, 1), config=cfg, training_data=DS())
    engine2.load_checkpoint(str(tmp_path), tag="c")
    st = engine2.training_data_sampler.state_dict()
    assert st["consumed_samples"] == saved["consumed_samples"] > 0 and st["curriculum_step"] == saved["curriculum_step"]


def test_engine_runs_random_ltd_schedule_and_checkpoints_it(tmp_path):
    """``data_efficiency.data_routing.random_ltd`` on the engine: the scheduler is created by the engine, handed to the
    wrapped layers by ``convert_to_random_ltd`` -> ``engine.random_ltd_initialize()``, advanced every forward, saved and
    restored with the checkpoint, and checkpoints carry the unwrapped key names (ref: engine.py:365-369, 691, 2002, 2766)."""
    import hcache_deepspeed_b200 as ds
    from hcache_deepspeed_b200.runtime.data_pipeline.data_routing import RandomLayerTokenDrop, convert_to_random_ltd

    class Block(torch.nn.Module):
        def __init__(self):
            super().__init__()
            self.fc = torch.nn.Linear(8, 8)
            self.seen = []

        def forward(self, x):
            self.seen.append(x.shape[1])
            return x + torch.tanh(self.fc(x))

    class Net(torch.nn.Module):
        def __init__(self):
            super().__init__()
            self.blocks = torch.nn.ModuleList([Block() for _ in range(3)])

        def forward(self, x):
            for b in self.blocks:
                x = b(x)
            return x.pow(2).mean()

    def build():
        torch.manual_seed(0)
        cfg = {"train_micro_batch_size_per_gpu": 2, "optimizer": {"type": "SGD", "params": {"lr": 0.01}}, "steps_per_print": 1000,
               "data_efficiency": {"enabled": True, "data_routing": {"enabled": True, "random_ltd": {
                   "enabled": True, "total_layer_num": 3, "random_ltd_layer_num": 2, "random_ltd_layer_id": [1, 2],
                   "random_ltd_schedule": {"min_value": 4, "max_value": 16, "schedule_type": "fixed_linear",
                                           "schedule_config": {"require_steps": 4, "seq_per_step": 4}}}}}}
        engine, _, _, _ = ds.initialize(model=Net(), config=cfg)
        convert_to_random_ltd(engine, Block)
        return engine

    engine = build()
    assert engine.random_ltd_enabled() and engine.random_ltd_scheduler is not None
    wrapped = [m for m in engine.module.modules() if isinstance(m, RandomLayerTokenDrop)]
    assert len(wrapped) == 3 and [w.scheduler is engine.random_ltd_scheduler for w in wrapped] == [False, True, True]
    x = torch.randn(2, 16, 8)
    kept = []
    for _ in range(5):
        loss = engine(x)
        engine.backward(loss)
        engine.step()
        kept.append(engine.module.blocks[1].random_ltd_layer.seen[-1])
        assert engine.module.blocks[0].random_ltd_layer.seen[-1] == 16      # not in random_ltd_layer_id: every token
    assert kept == [4, 4, 8, 12, 16], kept
    assert not any("random_ltd_layer" in k for k in engine.module_state_dict())
    engine.save_checkpoint(str(tmp_path), tag="t")
    fresh = build()
    fresh.load_checkpoint(str(tmp_path), tag="t")
    assert fresh.random_ltd_scheduler.state_dict() == engine.random_ltd_scheduler.state_dict()
    for a, b in zip(fresh.module.parameters(), engine.module.parameters()):
        assert torch.equal(a, b)


def test_sampler_clusters_unlock_with_difficulty_and_drain_without_replacement():
    """Cluster scheme of the reference's sampler (data_sampler.py:171-340): a move of the difficulty forms a new cluster of
    the newly eligible samples, batches draw from all clusters in proportion to size, each cluster is drained in a
    shuffled order before any of its samples repeats; the stream resumes exactly from ``state_dict``."""
    from hcache_deepspeed_b200.runtime.data_pipeline.data_sampler import DeepSpeedDataSampler
    n = 240
    diff = np.repeat(np.arange(1, 7), n // 6)                       # six difficulty levels, 40 samples each
    cfg = {"seed": 5, "data_sampling": {"curriculum_learning": {"enabled": True, "curriculum_metrics": {"len": {
        "difficulty_type": "value", "clustering_type": "schedule_based", "min_difficulty": 1, "max_difficulty": 6,
        "schedule_type": "fixed_discrete", "schedule_config": {"difficulty": [1, 3, 6], "max_step": [4, 8]}}}}}}

    def make():
        return DeepSpeedDataSampler(cfg, n, micro_batch_size=5, data_parallel_rank=1, data_parallel_size=2,
                                    gradient_accumulation_steps=2, difficulties={"len": diff})

    s = make()
    assert s.global_batch_size == 20 and s.get_start_end_idx() == (5, 10) and s.get_start_end_idx(7) == (4, 7)
    assert s.get_sample_based_on_metric_value("len", 1, 3).tolist() == list(range(40, 120))
    seen = []
    for step in range(1, 5):                                        # difficulty 1: one cluster of 40 samples
        b = s.get_next_global_batch()
        assert diff[b].max() == 1
        seen += b.tolist()
    assert len(s.data_clusters) == 1 and len(set(seen[:40])) == 40 and set(seen[40:]) <= set(range(40))   # drained, then reshuffled
    b = s.get_next_global_batch()                                   # step 5: difficulty 3 unlocks 80 more as a second cluster
    assert s.data_cluster_sizes == [40, 80] and diff[b].max() <= 3
    counts = np.zeros(2)
    for _ in range(3):
        counts += s.sample_from_clusters()
    assert counts.sum() == 60 and counts[1] > counts[0]            # proportional to cluster size
    state = s.state_dict()
    nxt = [s.get_next_global_batch().tolist() for _ in range(5)]    # crosses into difficulty 6 (third cluster)
    assert s.data_cluster_sizes == [40, 80, 120] and int(s._clustered.sum()) == n
    r = make()
    r.load_state_dict(state)
    assert [r.get_next_global_batch().tolist() for _ in range(5)] == nxt
    # percentile-based difficulty: the easiest k/max_difficulty of the data, whatever the metric's values are
    pc = {"data_sampling": {"curriculum_learning": {"enabled": True, "curriculum_metrics": {"len": {
        "difficulty_type": "percentile", "min_difficulty": 25, "max_difficulty": 100, "schedule_type": "fixed_discrete",
        "schedule_config": {"difficulty": [25, 100], "max_step": [2]}}}}}}
    p = DeepSpeedDataSampler(pc, n, micro_batch_size=10, data_parallel_rank=0, data_parallel_size=1, difficulties={"len": diff * 1000})
    first = p.get_next_global_batch()
    assert p.data_cluster_sizes == [50] and set(first.tolist()) <= set(np.argsort(diff, kind="stable")[:50].tolist())   # 25 % of 240 = 50 (2 per step)
    micro = list(iter(make()))
    assert len(micro) == (n // 20) * 2 and all(len(m) == 5 for m in micro)
