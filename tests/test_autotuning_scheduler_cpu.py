"""Autotuning experiment scheduler: slot reservation, parallel subprocess experiments, result parsing, resume."""
import json
import os
import sys
import textwrap

from hcache_deepspeed_b200.autotuning import Node, ResourceManager, get_autotuning_config
from hcache_deepspeed_b200.autotuning.utils import (canonical_name, get_all_configs, get_val_by_key, memory_to_string,
                                                    number_to_string, prune_configs, set_val_by_key)

SCRIPT = textwrap.dedent("""
    import argparse, json, os, sys
    p = argparse.ArgumentParser()
    p.add_argument("--deepspeed_config")
    p.add_argument("--fail_on", type=int, default=-1)
    a = p.parse_args()
    cfg = json.load(open(a.deepspeed_config))
    mbs = cfg["train_micro_batch_size_per_gpu"]
    if mbs == a.fail_on:
        sys.stderr.write("RuntimeError: out of memory\\n")
        sys.exit(1)
    dev = os.environ.get("CUDA_VISIBLE_DEVICES", "")
    json.dump({"throughput": 100.0 * mbs - mbs * mbs, "latency": 1.0 / mbs, "devices": dev},
              open(cfg["autotuning"]["metric_path"], "w"))
""")


def test_node_slots():
    n = Node("h", 4)
    a = n.reserve_slots(3)
    assert a == [0, 1, 2] and n.reserve_slots(2) is None
    n.restore_slots(a)
    assert n.idle_slots == [0, 1, 2, 3]


def test_resource_manager_runs_queue_in_parallel(tmp_path):
    script = tmp_path / "train.py"
    script.write_text(SCRIPT)
    exps = []
    for mbs in (1, 2, 4, 8):
        path = tmp_path / f"exp_{mbs}.json"
        path.write_text(json.dumps({"name": f"z0_tmbspg{mbs}", "num_gpus": 1, "num_nodes": 1, "user_script": str(script),
                                    "user_args": ["--fail_on", "8"],
                                    "ds_config": {"train_micro_batch_size_per_gpu": mbs}}))
        exps.append(str(path))
    rm = ResourceManager(hosts=["localhost"], num_gpus_per_node=2, results_dir=str(tmp_path / "res"), exps_dir=str(tmp_path / "exps"))
    rm.schedule_experiments(exps)
    assert rm.status() == {"queued": 4, "running": 0, "finished": 0}
    rm.run(poll_s=0.05)
    assert rm.status() == {"queued": 0, "running": 0, "finished": 4} and rm.nodes[0].idle_slots == [0, 1]
    best, val = rm.parse_results("throughput")
    assert best["name"] == "z0_tmbspg4" and val == 384.0
    best, val = rm.parse_results("latency")
    assert best["name"] == "z0_tmbspg4" and val == 0.25          # mbs 8 failed
    errs = {e["name"]: err for e, err in rm.finished_experiments.values()}
    assert "out of memory" in errs["z0_tmbspg8"] and errs["z0_tmbspg1"] is None
    used = {json.load(open(tmp_path / "res" / f"z0_tmbspg{m}" / "metrics.json"))["devices"] for m in (1, 2, 4)}
    assert used <= {"0", "1"}
    # a second manager over the same result directory takes finished experiments as done
    rm2 = ResourceManager(hosts=["localhost"], num_gpus_per_node=2, results_dir=str(tmp_path / "res"), exps_dir=str(tmp_path / "exps"))
    rm2.schedule_experiments(exps)
    assert rm2.status()["finished"] == 3 and rm2.status()["queued"] == 1
    # an experiment that can never fit is reported, not waited for
    big = tmp_path / "big.json"
    big.write_text(json.dumps({"name": "big", "num_gpus": 4, "num_nodes": 1, "user_script": str(script), "ds_config":
                               {"train_micro_batch_size_per_gpu": 1}}))
    rm3 = ResourceManager(hosts=["localhost"], num_gpus_per_node=2, results_dir=str(tmp_path / "r3"), exps_dir=str(tmp_path / "e3"))
    rm3.schedule_experiments([str(big)])
    rm3.run(poll_s=0.01)
    assert "slots" in list(rm3.finished_experiments.values())[0][1]


def test_autotuning_utils_and_config():
    space = {"zero_optimization": {"stage": 2, "overlap_comm": [True, False]}, "train_micro_batch_size_per_gpu": [1, 2]}
    cfgs = get_all_configs(space)
    assert len(cfgs) == 4 and all(c["zero_optimization"]["stage"] == 2 for c in cfgs)
    assert len(prune_configs(cfgs + cfgs)) == 4 and len(prune_configs(cfgs, ignored_keys=("overlap_comm",))) == 2
    assert canonical_name(cfgs[0], tuning_keys=["stage", "train_micro_batch_size_per_gpu"]) == "z2_tmbspg1"
    assert get_val_by_key(cfgs[0], "overlap_comm") is True
    set_val_by_key(cfgs[0], "stage", 3)
    assert cfgs[0]["zero_optimization"]["stage"] == 3
    assert memory_to_string(3 * 2 ** 30) == "3.0 G" and number_to_string(2.5e6) == "2.5 M"
    c = get_autotuning_config({"autotuning": {"enabled": True, "metric": "latency"}})
    assert c.enabled and c.metric == "latency" and c.tuner_type == "gridsearch"


def test_run_experiment_goes_through_the_launcher(tmp_path):
    """``run_experiment`` (module level, ref: autotuning/scheduler.py:310): one experiment on reserved slots through this
    package's launcher - 2 local ranks - with the metric file of the autotuning block filled by the script."""
    import json
    import os
    from hcache_deepspeed_b200.autotuning.scheduler import Node, Reservation, clean_up, get_job_id, get_user, run_experiment
    script = tmp_path / "train.py"
    script.write_text(
        "import argparse, json, os\n"
        "p = argparse.ArgumentParser(); p.add_argument('--deepspeed_config'); p.add_argument('--local_rank', type=int, default=0)\n"
        "a, _ = p.parse_known_args()\n"
        "cfg = json.load(open(a.deepspeed_config))\n"
        "if int(os.environ.get('RANK', 0)) == 0:\n"
        "    json.dump({'throughput': 10.0 * cfg['train_micro_batch_size_per_gpu'], 'world': int(os.environ['WORLD_SIZE'])},\n"
        "              open(cfg['autotuning']['metric_path'], 'w'))\n")
    node = Node("localhost", 2)
    res = Reservation(node, node.reserve_slots(2))
    exp = {"exp_id": 0, "name": "e0", "result_dir": str(tmp_path / "e0"), "ds_config": {"train_micro_batch_size_per_gpu": 4},
           "master_port": 29733}
    assert run_experiment(exp, [res], str(script), ["--deepspeed_config", "placeholder.json"]) == 0
    assert json.load(open(os.path.join(exp["result_dir"], "metrics.json"))) == {"throughput": 40.0, "world": 2}
    assert exp["launcher_args"][:2] == ["--include", "localhost:0,1"] and os.path.exists(os.path.join(exp["result_dir"], "exp.json"))
    clean_up(exp, [res])                                   # nothing left to stop: must not raise
    assert get_job_id() and get_user()
