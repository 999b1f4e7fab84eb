"""Per-node process spawner (ref: deepspeed/launcher/launch.py:133-348): one process per local GPU with
RANK / LOCAL_RANK / WORLD_SIZE / MASTER_ADDR / MASTER_PORT set; if any rank exits non-zero the whole local process
tree is terminated (a dead tensor-parallel rank would otherwise leave its peers spinning on a barrier)."""
from __future__ import annotations

import argparse
import base64
import json
import os
import signal
import subprocess
import sys
import time
from collections import defaultdict

import psutil


def parse_args(args=None):
    p = argparse.ArgumentParser()
    p.add_argument("--node_rank", type=int, default=0)
    p.add_argument("--master_addr", default="127.0.0.1", type=str)
    p.add_argument("--master_port", default=29500, type=int)
    p.add_argument("--world_info", default="None", type=str)
    p.add_argument("--module", action="store_true")
    p.add_argument("--no_python", action="store_true")
    p.add_argument("--no_local_rank", action="store_true")
    p.add_argument("--save_pid", type=int, default=0)
    p.add_argument("--enable_each_rank_log", default="None", type=str)
    p.add_argument("--enable_elastic_training", action="store_true",
                   help="supervise the local ranks with the elastic agent: restart the group on failure")
    p.add_argument("--max_elastic_restarts", type=int, default=3)
    p.add_argument("training_script", type=str)
    p.add_argument("training_script_args", nargs=argparse.REMAINDER)
    return p.parse_args(args=args)


def terminate_process_tree(pid: int) -> None:
    try:
        parent = psutil.Process(pid)
    except psutil.NoSuchProcess:
        return
    children = parent.children(recursive=True) + [parent]
    for c in children:
        try:
            c.terminate()
        except psutil.NoSuchProcess:
            pass
    _, alive = psutil.wait_procs(children, timeout=30)
    for c in alive:
        try:
            c.kill()
        except psutil.NoSuchProcess:
            pass


def decode_world_info(world_info: str):
    return json.loads(base64.urlsafe_b64decode(world_info))


def main(args=None):
    args = parse_args(args)
    assert args.world_info != "None", "must provide world info dict"
    world_info = decode_world_info(args.world_info)
    node_list = list(world_info.keys())
    local_node = node_list[args.node_rank]
    local_slots = world_info[local_node]
    global_rank_mapping = defaultdict(list)
    curr = 0
    for node in node_list:
        for _ in world_info[node]:
            global_rank_mapping[node].append(curr)
            curr += 1
    world_size = curr
    env = os.environ.copy()
    env["MASTER_ADDR"] = args.master_addr
    env["MASTER_PORT"] = str(args.master_port)
    env["WORLD_SIZE"] = str(world_size)
    env["CROSS_RANK"] = str(args.node_rank)
    env["CROSS_SIZE"] = str(len(node_list))
    env["LOCAL_SIZE"] = str(len(local_slots))
    env.setdefault("CUDA_VISIBLE_DEVICES", ",".join(map(str, local_slots)))
    processes = []
    log_dir = None if args.enable_each_rank_log == "None" else args.enable_each_rank_log
    if log_dir:
        os.makedirs(log_dir, exist_ok=True)
    if args.enable_elastic_training:
        # ref launch.py:190-208: hand the local ranks to the elastic agent (restart on failure, bounded)
        from ..elasticity.elastic_agent import DSElasticAgent, WorkerState
        cmd = ([sys.executable, "-u"] + (["-m"] if args.module else []) if not args.no_python else []) + \
            [args.training_script] + args.training_script_args
        agent = DSElasticAgent(cmd, len(local_slots), env={k: env[k] for k in ("CROSS_RANK", "CROSS_SIZE", "LOCAL_SIZE",
                                                                               "CUDA_VISIBLE_DEVICES") if k in env},
                               max_restarts=args.max_elastic_restarts, master_addr=args.master_addr,
                               master_port=args.master_port, node_rank=args.node_rank,
                               membership_fn=lambda: len(node_list), log_dir=log_dir)
        sys.exit(0 if agent.run().state == WorkerState.SUCCEEDED else 1)
    for local_rank, _slot in enumerate(local_slots):
        rank = global_rank_mapping[local_node][local_rank]
        e = env.copy()
        e["RANK"] = str(rank)
        e["LOCAL_RANK"] = str(local_rank)
        cmd = []
        if not args.no_python:
            cmd = [sys.executable, "-u"] + (["-m"] if args.module else [])
        cmd.append(args.training_script)
        if not args.no_local_rank:
            cmd.append(f"--local_rank={local_rank}")
        cmd += args.training_script_args
        if log_dir:
            f = open(os.path.join(log_dir, f"rank_{rank}.log"), "w")
            processes.append(subprocess.Popen(cmd, env=e, stdout=f, stderr=f))
        else:
            processes.append(subprocess.Popen(cmd, env=e))

    def sigkill_handler(signum, frame):
        for p in processes:
            terminate_process_tree(p.pid)
        sys.exit(1)

    signal.signal(signal.SIGINT, sigkill_handler)
    signal.signal(signal.SIGTERM, sigkill_handler)
    alive = set(processes)
    last_rc = 0
    while alive:
        finished = {p for p in alive if p.poll() is not None}
        for p in finished:
            if p.returncode != 0:
                last_rc = p.returncode
                for q in processes:
                    if q.poll() is None:
                        terminate_process_tree(q.pid)
                sys.exit(last_rc)
        alive -= finished
        time.sleep(0.5)
    sys.exit(last_rc)


if __name__ == "__main__":
    main()
