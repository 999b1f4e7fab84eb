"""Elastic worker-group supervisor: start the local ranks, watch them, and on a failure (or a change in the set of
participating nodes) tear the whole group down and start it again — up to ``max_restarts`` times — with the restart
count and world size exported to the workers so they can resume from their last checkpoint.

Reference: ``deepspeed/elasticity/elastic_agent.py:32-190`` (``DSElasticAgent(LocalElasticAgent)``: torch-elastic agent
whose ``_invoke_run`` loop polls worker state and the rendezvous, restarting on ``FAILED`` / ``UNHEALTHY`` or when nodes
are waiting to join) and ``launcher/launch.py:190-208``. This agent is self-contained (plain subprocesses, no
torch-elastic rendezvous): membership comes from a ``membership_fn`` callback — e.g. re-reading a hostfile — so it
also works where no etcd / c10d store is reachable.
"""
from __future__ import annotations

import os
import signal
import subprocess
import sys
import time
from dataclasses import dataclass, field
from enum import Enum
from typing import Callable, Dict, List, Optional, Sequence


class WorkerState(str, Enum):
    INIT = "INIT"
    HEALTHY = "HEALTHY"
    SUCCEEDED = "SUCCEEDED"
    FAILED = "FAILED"
    STOPPED = "STOPPED"


@dataclass
class RunResult:
    state: WorkerState
    restarts: int
    return_codes: Dict[int, int] = field(default_factory=dict)
    world_sizes: List[int] = field(default_factory=list)      # world size of every attempt


class DSElasticAgent:

    def __init__(self, cmd: Sequence[str], nproc_per_node: int, env: Optional[Dict[str, str]] = None, max_restarts: int = 3,
                 monitor_interval: float = 0.2, master_addr: str = "127.0.0.1", master_port: int = 29500, node_rank: int = 0,
                 membership_fn: Optional[Callable[[], int]] = None, exit_barrier_timeout: float = 300.0,
                 log_dir: Optional[str] = None, run_id: str = "hds_elastic"):
        """``cmd``: worker command line; ``membership_fn``: current number of nodes (polled every interval; a change
        restarts the group with the new world size and does not count against ``max_restarts``)."""
        self.cmd, self.nproc, self.ds_env = list(cmd), nproc_per_node, dict(env or {})
        self.max_restarts, self.monitor_interval = max_restarts, monitor_interval
        self.master_addr, self.master_port, self.node_rank = master_addr, master_port, node_rank
        self.membership_fn = membership_fn or (lambda: 1)
        self.exit_barrier_timeout, self.log_dir, self.run_id = exit_barrier_timeout, log_dir, run_id
        self._procs: List[subprocess.Popen] = []
        self._remaining_restarts = max_restarts
        self.state = WorkerState.INIT

    # ---- worker group ------------------------------------------------------------------------------------------
    def _worker_env(self, local_rank: int, n_nodes: int, restart_count: int) -> Dict[str, str]:
        world = n_nodes * self.nproc
        env = dict(os.environ)
        env.update(self.ds_env)
        env.update({
            "LOCAL_RANK": str(local_rank), "RANK": str(self.node_rank * self.nproc + local_rank), "GROUP_RANK": str(self.node_rank),
            "LOCAL_WORLD_SIZE": str(self.nproc), "WORLD_SIZE": str(world), "GROUP_WORLD_SIZE": str(n_nodes),
            "MASTER_ADDR": self.master_addr, "MASTER_PORT": str(self.master_port),
            "TORCHELASTIC_RESTART_COUNT": str(restart_count), "TORCHELASTIC_MAX_RESTARTS": str(self.max_restarts),
            "TORCHELASTIC_RUN_ID": self.run_id, "NCCL_ASYNC_ERROR_HANDLING": os.getenv("NCCL_ASYNC_ERROR_HANDLING", "1"),
        })
        return env

    def _start_workers(self, n_nodes: int, restart_count: int) -> None:
        self._procs = []
        for lr in range(self.nproc):
            out = None
            if self.log_dir:
                d = os.path.join(self.log_dir, f"attempt_{restart_count}")
                os.makedirs(d, exist_ok=True)
                out = open(os.path.join(d, f"rank_{lr}.log"), "w")
            self._procs.append(subprocess.Popen(self.cmd, env=self._worker_env(lr, n_nodes, restart_count), stdout=out,
                                                stderr=subprocess.STDOUT if out else None, start_new_session=True))
        self.state = WorkerState.HEALTHY

    def _stop_workers(self, grace: float = 5.0) -> None:
        """SIGTERM every worker's process group, SIGKILL what is still alive after ``grace`` seconds."""
        for p in self._procs:
            if p.poll() is None:
                try:
                    os.killpg(p.pid, signal.SIGTERM)
                except ProcessLookupError:
                    pass
        deadline = time.time() + grace
        for p in self._procs:
            try:
                p.wait(timeout=max(0.0, deadline - time.time()))
            except subprocess.TimeoutExpired:
                try:
                    os.killpg(p.pid, signal.SIGKILL)
                except ProcessLookupError:
                    pass
                p.wait()

    def _monitor(self) -> WorkerState:
        codes = [p.poll() for p in self._procs]
        if any(c is not None and c != 0 for c in codes):
            return WorkerState.FAILED
        if all(c == 0 for c in codes):
            return WorkerState.SUCCEEDED
        return WorkerState.HEALTHY

    # ---- main loop ------------------------------------------------------------------------------------------------
    def run(self) -> RunResult:
        n_nodes = self.membership_fn()
        restart_count = 0
        result = RunResult(WorkerState.INIT, 0)
        self._start_workers(n_nodes, restart_count)
        result.world_sizes.append(n_nodes * self.nproc)
        try:
            while True:
                time.sleep(self.monitor_interval)
                state = self._monitor()
                if state == WorkerState.SUCCEEDED:
                    self.state = state
                    break
                if state == WorkerState.FAILED:
                    if self._remaining_restarts <= 0:
                        self._stop_workers()
                        self.state = WorkerState.FAILED
                        break
                    self._remaining_restarts -= 1
                    restart_count += 1
                    self._stop_workers()
                    n_nodes = self.membership_fn()
                    self._start_workers(n_nodes, restart_count)
                    result.world_sizes.append(n_nodes * self.nproc)
                    continue
                now = self.membership_fn()
                if now != n_nodes:                      # scale-up / scale-down: same attempt number, new world size
                    self._stop_workers()
                    n_nodes = now
                    self._start_workers(n_nodes, restart_count)
                    result.world_sizes.append(n_nodes * self.nproc)
        except KeyboardInterrupt:
            self._stop_workers()
            self.state = WorkerState.STOPPED
        result.state, result.restarts = self.state, restart_count
        result.return_codes = {i: (p.returncode if p.returncode is not None else -1) for i, p in enumerate(self._procs)}
        return result


def main(argv=None) -> int:
    """``python -m hcache_deepspeed_b200.elasticity.elastic_agent --nproc 2 --max_restarts 3 -- python train.py ...``"""
    import argparse
    ap = argparse.ArgumentParser(description="elastic supervisor for one node's ranks")
    ap.add_argument("--nproc", type=int, default=1)
    ap.add_argument("--max_restarts", type=int, default=3)
    ap.add_argument("--master_addr", default="127.0.0.1")
    ap.add_argument("--master_port", type=int, default=29500)
    ap.add_argument("--node_rank", type=int, default=0)
    ap.add_argument("--hostfile", default=None, help="membership = number of non-empty lines, re-read while running")
    ap.add_argument("--log_dir", default=None)
    ap.add_argument("cmd", nargs=argparse.REMAINDER)
    a = ap.parse_args(argv)
    cmd = a.cmd[1:] if a.cmd and a.cmd[0] == "--" else a.cmd
    assert cmd, "no worker command given"

    def members() -> int:
        if not a.hostfile or not os.path.exists(a.hostfile):
            return 1
        with open(a.hostfile) as f:
            return max(1, sum(1 for ln in f if ln.strip() and not ln.lstrip().startswith("#")))

    res = DSElasticAgent(cmd, a.nproc, max_restarts=a.max_restarts, master_addr=a.master_addr, master_port=a.master_port,
                         node_rank=a.node_rank, membership_fn=members, log_dir=a.log_dir).run()
    return 0 if res.state == WorkerState.SUCCEEDED else 1


if __name__ == "__main__":
    sys.exit(main())
