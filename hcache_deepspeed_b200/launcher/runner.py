"""Job launcher front end (ref: deepspeed/launcher/runner.py + multinode_runner.py).

``hds [--hostfile F] [--include ...] [--exclude ...] [--num_nodes N] [--num_gpus G] [--master_addr A]
[--master_port P] [--launcher pdsh|ssh|local] script.py args...``

Single node: spawns ``launcher.launch`` directly. Multi node: builds the per-node command line and runs it over
pdsh / ssh with the resource map base64-encoded, exactly as the reference does. One process per GPU; ranks
rendezvous through ``torch.distributed`` (NCCL).
"""
from __future__ import annotations

import argparse
import base64
import collections
import json
import os
import re
import shlex
import subprocess
import sys
from typing import Dict, List, Optional

from ..utils.logging import logger

DLTS_HOSTFILE = "/job/hostfile"
EXPORT_ENVS = ["NCCL", "PYTHON", "MV2", "UCX", "HDS", "CUDA"]
PDSH_MAX_FAN_OUT = 1024


def parse_args(args=None):
    p = argparse.ArgumentParser(description="hcache_deepspeed_b200 runner: launch one process per GPU on one or more nodes")
    p.add_argument("-H", "--hostfile", type=str, default=DLTS_HOSTFILE)
    p.add_argument("-i", "--include", type=str, default="")
    p.add_argument("-e", "--exclude", type=str, default="")
    p.add_argument("--num_nodes", type=str, default="-1", help="N, or MIN:MAX with --elastic_training")
    p.add_argument("--min_elastic_nodes", type=int, default=-1)
    p.add_argument("--max_elastic_nodes", type=int, default=-1)
    p.add_argument("--num_gpus", "--num_accelerators", type=int, default=-1)
    p.add_argument("--master_port", default=29500, type=int)
    p.add_argument("--master_addr", default="", type=str)
    p.add_argument("--launcher", default="pdsh", type=str, help="pdsh | ssh | local")
    p.add_argument("--launcher_args", default="", type=str)
    p.add_argument("--module", action="store_true")
    p.add_argument("--no_python", action="store_true")
    p.add_argument("--no_local_rank", action="store_true")
    p.add_argument("--no_ssh_check", action="store_true")
    p.add_argument("--force_multi", action="store_true")
    p.add_argument("--save_pid", action="store_true")
    p.add_argument("--enable_each_rank_log", default="None", type=str)
    p.add_argument("--bind_cores_to_rank", action="store_true")
    p.add_argument("--elastic_training", action="store_true")
    p.add_argument("--autotuning", default="", choices=["", "tune", "run"],
                   help="tune: search ZeRO stage x micro-batch size with short runs of the script and write the best config; "
                        "run: tune, then launch the script with that config")
    p.add_argument("user_script", type=str)
    p.add_argument("user_args", nargs=argparse.REMAINDER)
    return p.parse_args(args=args)


def fetch_hostfile(hostfile_path: str) -> Optional["collections.OrderedDict[str, int]"]:
    if not os.path.isfile(hostfile_path):
        return None
    with open(hostfile_path) as fd:
        return _parse_hostfile(fd.readlines())


def _parse_hostfile(lines: List[str]) -> "collections.OrderedDict[str, int]":
    pattern = r"^(\S+)\s+slots=(\d+)"
    pool = collections.OrderedDict()
    for line in lines:
        line = line.strip()
        if not line or line.startswith("#"):
            continue
        m = re.search(pattern, line)
        if not m:
            raise ValueError(f"Hostfile contains a bad entry: {line}, unable to proceed with launching")
        host, slots = m.group(1), int(m.group(2))
        if host in pool:
            raise ValueError(f"Hostfile contains multiple entries for {host}, unable to proceed with launching")
        pool[host] = slots
    if not pool:
        raise ValueError("Hostfile is empty or not formatted correctly, unable to proceed with launching.")
    return pool


def parse_resource_filter(host_info: Dict[str, List[int]], include_str: str = "", exclude_str: str = "") -> Dict[str, List[int]]:
    """``host1@host2:0,2`` style include / exclude filters (mutually exclusive), as in the reference."""
    if include_str and exclude_str:
        raise ValueError("include_str and exclude_str are mutually exclusive.")
    if not include_str and not exclude_str:
        return host_info
    filtered: Dict[str, List[int]] = {}
    parse_str = include_str or exclude_str
    if exclude_str:
        filtered = {h: list(s) for h, s in host_info.items()}
    for node_config in parse_str.split("@"):
        if ":" in node_config:
            hostname, slots = node_config.split(":")
            slots = [int(x) for x in slots.split(",")]
            if hostname not in host_info:
                raise ValueError(f"Hostname '{hostname}' not found in hostfile")
            for s in slots:
                if s not in host_info[hostname]:
                    raise ValueError(f"No slot '{s}' specified on host '{hostname}'")
            if include_str:
                filtered.setdefault(hostname, [])
                filtered[hostname] += slots
            else:
                for s in slots:
                    filtered[hostname].remove(s)
        else:
            hostname = node_config
            if hostname not in host_info:
                raise ValueError(f"Hostname '{hostname}' not found in hostfile")
            if include_str:
                filtered[hostname] = list(host_info[hostname])
            else:
                filtered[hostname] = []
    filtered = {h: sorted(set(s)) for h, s in filtered.items() if s}
    ordered = collections.OrderedDict()
    for host in host_info:
        if host in filtered:
            ordered[host] = filtered[host]
    return ordered


def parse_node_config(node_config: str):
    """``host`` -> (host, []); ``host:0,2`` -> (host, [0, 2])."""
    if ":" not in node_config:
        return node_config, []
    hostname, slots = node_config.split(":")
    return hostname, [int(x) for x in slots.split(",")]


def parse_node_config_list(node_config_list: str) -> Dict[str, List[int]]:
    """``host1:0,1@host2@host1:3`` -> {host1: [0, 1, 3], host2: []} (slots merged, sorted, de-duplicated)."""
    out: Dict[str, List[int]] = collections.OrderedDict()
    for node_config in node_config_list.split("@"):
        hostname, slots = parse_node_config(node_config)
        out.setdefault(hostname, [])
        out[hostname] += slots
    return {k: sorted(set(v)) for k, v in out.items()}


def parse_num_nodes(str_num_nodes, elastic_training: bool):
    """``--num_nodes``: ``N`` -> (N, -1); ``MIN:MAX`` (elastic training only) -> (MIN, MAX)."""
    parts = str(str_num_nodes).split(":")
    if len(parts) == 1:
        return int(parts[0]), -1
    if len(parts) == 2 and elastic_training:
        return int(parts[0]), int(parts[1])
    if len(parts) == 2:
        raise RuntimeError("MIN:MAX format is only supported in elastic training")
    raise RuntimeError(f"num_nodes {str_num_nodes} is not in MIN:MAX format")


def parse_inclusion_exclusion(resource_pool, inclusion: str, exclusion: str):
    active = collections.OrderedDict((h, list(range(s))) for h, s in resource_pool.items())
    return parse_resource_filter(active, include_str=inclusion, exclude_str=exclusion)


def encode_world_info(world_info) -> str:
    return base64.urlsafe_b64encode(json.dumps(world_info).encode()).decode()


def build_launch_cmd(args, world_info_b64: str, node_rank: str = "%n") -> List[str]:
    cmd = [sys.executable, "-u", "-m", "hcache_deepspeed_b200.launcher.launch", f"--world_info={world_info_b64}",
           f"--node_rank={node_rank}", f"--master_addr={args.master_addr}", f"--master_port={args.master_port}"]
    if args.no_python:
        cmd.append("--no_python")
    if args.module:
        cmd.append("--module")
    if args.no_local_rank:
        cmd.append("--no_local_rank")
    if args.save_pid:
        cmd += ["--save_pid", str(os.getpid())]
    if args.enable_each_rank_log != "None":
        cmd.append(f"--enable_each_rank_log={args.enable_each_rank_log}")
    return cmd + [args.user_script] + list(args.user_args)


def build_multinode_cmd(args, active_resources, world_info_b64: str, env: Dict[str, str]) -> List[str]:
    exports = " ".join(f"export {k}={shlex.quote(v)};" for k, v in env.items() if any(k.startswith(p) for p in EXPORT_ENVS))
    if args.launcher == "pdsh":
        hosts = ",".join(active_resources.keys())
        launch = build_launch_cmd(args, world_info_b64, node_rank="%n")
        return ["pdsh", "-S", "-f", str(PDSH_MAX_FAN_OUT), "-w", hosts] + shlex.split(args.launcher_args) + \
            [exports + f" cd {os.path.abspath('.')}; " + " ".join(map(shlex.quote, launch))]
    if args.launcher == "ssh":
        script = []
        for rank, host in enumerate(active_resources):
            launch = build_launch_cmd(args, world_info_b64, node_rank=str(rank))
            script.append(f"ssh {shlex.quote(host)} {shlex.quote(exports + ' cd ' + os.path.abspath('.') + '; ' + ' '.join(map(shlex.quote, launch)))} &")
        return ["bash", "-c", "\n".join(script) + "\nwait"]
    raise NotImplementedError(f"Unknown launcher {args.launcher}")


def _config_path_of(user_args: List[str]) -> Optional[str]:
    for flag in ("--deepspeed_config", "--deepspeed"):
        for i, a in enumerate(user_args):
            if a == flag and i + 1 < len(user_args) and user_args[i + 1].endswith(".json"):
                return user_args[i + 1]
            if a.startswith(flag + "="):
                return a.split("=", 1)[1]
    return None


def _with_config_path(user_args: List[str], path: str) -> List[str]:
    out, skip = [], False
    for a in user_args:
        if skip:
            skip = False
            continue
        if a in ("--deepspeed_config", ):
            skip = True
            continue
        if a.startswith("--deepspeed_config="):
            continue
        out.append(a)
    return out + ["--deepspeed_config", path]


def run_autotuning(args, active_resources) -> Optional[str]:
    """``--autotuning tune|run`` (ref: launcher/runner.py run_autotuning + autotuning/autotuner.py): a first short run of
    the script reports the model size, then the experiments (ZeRO stage x micro-batch size, pruned by the memory model)
    run through THIS launcher on the active resources; the best config is written to ``<results_dir>/ds_config_optimal.json``."""
    import json
    from ..autotuning.autotuner import Autotuner
    cfg_path = _config_path_of(list(args.user_args))
    if cfg_path is None:
        raise ValueError("--autotuning needs the script to be given its config with --deepspeed_config <file.json>")
    with open(cfg_path) as f:
        user_cfg = json.load(f)
    n_gpus = sum(len(s) for s in active_resources.values())
    base_cmd = [sys.executable, os.path.abspath(sys.argv[0])]
    for flag, val in (("--num_gpus", args.num_gpus), ("--num_nodes", args.num_nodes)):
        if val and val > 0:
            base_cmd += [flag, str(val)]
    if args.hostfile and os.path.isfile(args.hostfile):
        base_cmd += ["-H", args.hostfile]
    if args.include:
        base_cmd += ["--include", args.include]
    if args.master_addr:
        base_cmd += ["--master_addr", args.master_addr]

    def launch(cfg_file: str, port_offset: int) -> int:
        cmd = base_cmd + ["--master_port", str(args.master_port + 1 + port_offset), args.user_script] + \
            _with_config_path(list(args.user_args), cfg_file)
        return subprocess.run(cmd, stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL).returncode

    at_cfg = dict(user_cfg.get("autotuning") or {})
    results_dir = at_cfg.get("results_dir", "autotuning_results")
    os.makedirs(results_dir, exist_ok=True)
    info_path = os.path.join(results_dir, "model_info.json")
    probe = dict(user_cfg)
    probe["autotuning"] = dict(at_cfg, enabled=True, model_info_path=info_path, model_info={"profile": True})
    probe_file = os.path.join(results_dir, "ds_config_model_info.json")
    with open(probe_file, "w") as f:
        json.dump(probe, f)
    launch(probe_file, 0)
    if not os.path.exists(info_path):
        raise RuntimeError("autotuning: the model-info run wrote no model_info.json (does the script call initialize() with "
                           "--deepspeed_config and take at least one optimizer step?)")
    num_params = json.load(open(info_path))["num_params"]
    counter = {"n": 0}

    def run_fn(cfg):
        counter["n"] += 1
        exp_file = os.path.join(results_dir, f"exp_{counter['n']}.json")
        metric_file = os.path.join(results_dir, f"exp_{counter['n']}_metrics.json")
        full = dict(cfg)
        full["autotuning"] = dict(at_cfg, enabled=True, metric_path=metric_file)
        with open(exp_file, "w") as f:
            json.dump(full, f)
        launch(exp_file, counter["n"])
        if not os.path.exists(metric_file):
            return None
        return json.load(open(metric_file)).get(at_cfg.get("metric", "throughput"))

    tuner = Autotuner(user_cfg, num_params=num_params, num_gpus=n_gpus, run_fn=run_fn, results_dir=results_dir)
    best = tuner.tune()
    tuner.print_tuning_results()
    if best is None:
        return None
    path = tuner.write_optimal_config()
    print(f"autotuning: best config written to {path}")
    return path


def main(args=None):
    args = parse_args(args)
    if args.elastic_training:
        assert args.master_addr != "", "Master Addr is required when elastic training is enabled"
    args.num_nodes, max_nodes = parse_num_nodes(args.num_nodes, args.elastic_training)
    if args.elastic_training:
        if args.min_elastic_nodes < 0:
            args.min_elastic_nodes = args.num_nodes
        if args.max_elastic_nodes < 0:
            args.max_elastic_nodes = max_nodes if max_nodes > 0 else args.num_nodes
    resource_pool = fetch_hostfile(args.hostfile)
    visible = os.environ.get("CUDA_VISIBLE_DEVICES", "")
    if not resource_pool and visible:
        # one node, devices restricted by the environment: keep that restriction unless explicit filters were given
        if args.include or args.exclude or args.num_nodes > 1 or args.num_gpus > 0:
            logger.info(f"Detected CUDA_VISIBLE_DEVICES={visible} but ignoring it because one or several of --include/"
                        "--exclude/--num_gpus/--num_nodes cl args were used. If you want to use CUDA_VISIBLE_DEVICES don't pass "
                        "any of these arguments to the launcher.")
        else:
            args.include = f"localhost:{visible}"
            logger.info(f"Detected CUDA_VISIBLE_DEVICES={visible}: setting --include={args.include}")
            del os.environ["CUDA_VISIBLE_DEVICES"]
    if not resource_pool:
        import torch
        n = args.num_gpus if args.num_gpus > 0 else max(1, torch.cuda.device_count())
        if args.include.startswith("localhost:"):
            n = max(n, max(parse_node_config(args.include)[1]) + 1)
        resource_pool = collections.OrderedDict(localhost=n)
        args.master_addr = args.master_addr or "127.0.0.1"
    active = parse_inclusion_exclusion(resource_pool, args.include, args.exclude)
    if args.num_nodes > 0:
        active = collections.OrderedDict(list(active.items())[:args.num_nodes])
    if args.num_gpus > 0:
        active = collections.OrderedDict((h, s[:args.num_gpus]) for h, s in active.items())
    multi_node = args.force_multi or len(active) > 1
    if not args.master_addr:
        first = next(iter(active))
        args.master_addr = "127.0.0.1" if not multi_node else \
            subprocess.check_output(["ssh", first, "hostname -I"]).decode().split()[0]
    world_info = encode_world_info(active)
    if args.autotuning:
        best = run_autotuning(args, active)
        if args.autotuning == "tune" or best is None:
            return
        args.user_args = _with_config_path(list(args.user_args), best)
    env = os.environ.copy()
    # a source checkout that is not installed: the per-node launcher and the ranks import the package through PYTHONPATH
    # (exported to remote nodes too, "PYTHON" is in EXPORT_ENVS)
    pkg_root = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    paths = [p for p in env.get("PYTHONPATH", "").split(os.pathsep) if p]
    if pkg_root not in paths:
        env["PYTHONPATH"] = os.pathsep.join([pkg_root] + paths)
    if not multi_node:
        cmd = build_launch_cmd(args, world_info, node_rank="0")
    else:
        cmd = build_multinode_cmd(args, active, world_info, env)
    result = subprocess.Popen(cmd, env=env)
    result.wait()
    if result.returncode != 0:
        sys.exit(result.returncode)


if __name__ == "__main__":
    main()
