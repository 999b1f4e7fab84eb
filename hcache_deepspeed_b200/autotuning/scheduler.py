"""Experiment scheduler of the autotuner (``deepspeed/autotuning/scheduler.py:33-440``: ``ResourceManager`` with
``Node`` / ``Reservation`` / ``Job``): experiments are queued, each reserves ``num_gpus`` device slots on the known
hosts, runs as a subprocess with exactly those devices visible, and writes its metrics file; finished experiments free
their slots and the next queued one starts. Several experiments therefore run side by side on one 8-GPU node."""
from __future__ import annotations

import json
import os
import subprocess
import sys
import time
from dataclasses import dataclass, field
from typing import Dict, List, Optional, Tuple

from ..utils.logging import logger


@dataclass
class Node:
    host: str
    max_slots: int
    idle_slots: List[int] = field(default_factory=list)

    def __post_init__(self):
        self.idle_slots = list(range(self.max_slots))

    def reserve_slots(self, slot_request: int) -> Optional[List[int]]:
        if len(self.idle_slots) < slot_request:
            return None
        taken, self.idle_slots = self.idle_slots[:slot_request], self.idle_slots[slot_request:]
        return taken

    def restore_slots(self, slots: List[int]) -> None:
        self.idle_slots = sorted(set(self.idle_slots) | set(slots))


@dataclass
class Reservation:
    node: Node
    slots: List[int]

    def restore_slots(self) -> None:
        self.node.restore_slots(self.slots)

    def desc(self) -> str:
        return f"{self.node.host}:{','.join(map(str, self.slots))}@"


@dataclass
class Job:
    exp: Dict
    reservations: List[Reservation]
    process: Optional[subprocess.Popen] = None
    started: float = 0.0


class ResourceManager:

    def __init__(self, args=None, hosts: Optional[List[str]] = None, num_gpus_per_node: int = 1, results_dir: str = "autotuning_results",
                 exps_dir: str = "autotuning_exps", arg_mappings: Optional[Dict[str, str]] = None, timeout_s: float = 3600.0):
        self.args = args
        self.nodes = [Node(h, num_gpus_per_node) for h in (hosts or ["localhost"])]
        self.results_dir, self.exps_dir = results_dir, exps_dir
        self.arg_mappings = arg_mappings or {}
        self.timeout_s = timeout_s
        self.experiment_queue: List[Dict] = []
        self.running_experiments: Dict[int, Job] = {}
        self.finished_experiments: Dict[int, Tuple[Dict, Optional[str]]] = {}    # exp_id -> (exp, error)
        self.experiment_count = 0
        self.exp_paths = set()
        os.makedirs(results_dir, exist_ok=True)
        os.makedirs(exps_dir, exist_ok=True)

    # ---- queue -----------------------------------------------------------------------------------------------------
    def schedule_experiments(self, exp_paths: List[str]) -> None:
        """Queue experiment description files (``{"name", "ds_config", "num_gpus", "num_nodes"}``); an experiment whose
        result directory already holds metrics is taken as finished (resume of an interrupted tuning run)."""
        for path in exp_paths:
            if path in self.exp_paths:
                continue
            self.exp_paths.add(path)
            with open(path) as f:
                exp = json.load(f)
            exp["exp_id"] = self.experiment_count
            self.experiment_count += 1
            exp["result_dir"] = os.path.join(self.results_dir, exp["name"])
            metric_file = os.path.join(exp["result_dir"], "metrics.json")
            exp["ds_config"].setdefault("autotuning", {})["metric_path"] = metric_file
            if os.path.exists(metric_file):
                self.finished_experiments[exp["exp_id"]] = (exp, None)
                continue
            self.experiment_queue.append(exp)

    def resource_request(self, exp: Dict) -> Optional[List[Reservation]]:
        num_nodes, slots = exp.get("num_nodes", 1), exp.get("num_gpus", 1)
        got: List[Reservation] = []
        for node in self.nodes:
            if len(got) == num_nodes:
                break
            s = node.reserve_slots(slots)
            if s is not None:
                got.append(Reservation(node, s))
        if len(got) < num_nodes:
            for r in got:
                r.restore_slots()
            return None
        return got

    # ---- running ---------------------------------------------------------------------------------------------------
    def _command(self, exp: Dict, reservations: List[Reservation]) -> Tuple[List[str], Dict[str, str]]:
        os.makedirs(exp["result_dir"], exist_ok=True)
        cfg_path = os.path.join(exp["result_dir"], "ds_config.json")
        with open(cfg_path, "w") as f:
            json.dump(exp["ds_config"], f)
        with open(os.path.join(exp["result_dir"], "exp.json"), "w") as f:
            json.dump({k: v for k, v in exp.items() if k != "process"}, f)
        user_script = exp.get("user_script") or getattr(self.args, "user_script", None)
        user_args = list(exp.get("user_args") or getattr(self.args, "user_args", []) or [])
        # values of tuned keys that the script takes on its command line (``arg_mappings``) are rewritten in place
        for key, flag in self.arg_mappings.items():
            val = exp["ds_config"]
            for part in key.split("."):
                val = val.get(part) if isinstance(val, dict) else None
            if val is not None and flag in user_args:
                user_args[user_args.index(flag) + 1] = str(val)
        env = dict(os.environ)
        env["CUDA_VISIBLE_DEVICES"] = ",".join(map(str, reservations[0].slots))
        n = len(reservations[0].slots)
        if n > 1:
            cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={n}", "--master-addr",
                   "127.0.0.1", "--master-port", str(29600 + exp["exp_id"] % 300), user_script]
        else:
            cmd = [sys.executable, user_script]
        return cmd + user_args + ["--deepspeed_config", cfg_path], env

    def run_job(self, exp: Dict, reservations: List[Reservation]) -> None:
        cmd, env = self._command(exp, reservations)
        out = open(os.path.join(exp["result_dir"], "stdout.log"), "w")
        err = open(os.path.join(exp["result_dir"], "stderr.log"), "w")
        proc = subprocess.Popen(cmd, env=env, stdout=out, stderr=err)
        self.running_experiments[exp["exp_id"]] = Job(exp, reservations, proc, time.time())
        logger.info(f"autotuning: started exp {exp['exp_id']} ({exp['name']}) on {''.join(r.desc() for r in reservations)}")

    def experiment_check(self, pbar=None) -> None:
        for exp_id, job in list(self.running_experiments.items()):
            rc = job.process.poll()
            timed_out = rc is None and time.time() - job.started > self.timeout_s
            if rc is None and not timed_out:
                continue
            if timed_out:
                job.process.kill()
                job.process.wait()
            err = None
            if timed_out:
                err = "timeout"
            elif rc != 0:
                err = self._stderr_tail(job.exp) or f"exit code {rc}"
            self.finished_experiments[exp_id] = (job.exp, err)
            for r in job.reservations:
                r.restore_slots()
            del self.running_experiments[exp_id]
            if pbar is not None:
                pbar.update(1)

    @staticmethod
    def _stderr_tail(exp: Dict) -> Optional[str]:
        path = os.path.join(exp["result_dir"], "stderr.log")
        if not os.path.exists(path):
            return None
        lines = [l.strip() for l in open(path, errors="replace") if l.strip()]
        return lines[-1] if lines else None

    def run(self, poll_s: float = 0.2) -> None:
        """Drain the queue."""
        while self.experiment_queue or self.running_experiments:
            started = False
            if self.experiment_queue:
                exp = self.experiment_queue[0]
                res = self.resource_request(exp)
                if res is not None:
                    self.experiment_queue.pop(0)
                    self.run_job(exp, res)
                    started = True
                elif not self.running_experiments:
                    self.experiment_queue.pop(0)
                    self.finished_experiments[exp["exp_id"]] = (exp, "not enough device slots for this experiment")
                    continue
            self.experiment_check()
            if not started:
                time.sleep(poll_s)

    # ---- results ---------------------------------------------------------------------------------------------------
    def parse_results(self, metric: str) -> Tuple[Optional[Dict], Optional[float]]:
        """Best finished experiment and its metric value (``latency`` is minimised, everything else maximised)."""
        best, best_val = None, None
        for exp, err in self.finished_experiments.values():
            if err:
                logger.info(f"autotuning: exp {exp['exp_id']} ({exp['name']}) failed: {err}")
                continue
            path = exp["ds_config"]["autotuning"]["metric_path"]
            if not os.path.exists(path):
                continue
            with open(path) as f:
                val = json.load(f).get(metric)
            exp["result"] = val
            if val is None:
                continue
            if best_val is None or (val < best_val if metric == "latency" else val > best_val):
                best, best_val = exp, val
        return best, best_val

    def status(self) -> Dict[str, int]:
        return {"queued": len(self.experiment_queue), "running": len(self.running_experiments),
                "finished": len(self.finished_experiments)}

    def clear(self) -> None:
        for job in self.running_experiments.values():
            job.process.kill()
            job.process.wait()
            for r in job.reservations:
                r.restore_slots()
        self.experiment_queue, self.running_experiments, self.finished_experiments = [], {}, {}
        self.exp_paths = set()


# ---- module-level entry points of the reference's scheduler (autotuning/scheduler.py:288-430) ------------------------------
def get_job_id() -> str:
    """Job id given by the cluster infrastructure (DLWS / DLTS), else ``unknown-job-id``."""
    return os.environ.get("DLWS_JOB_ID") or os.environ.get("DLTS_JOB_ID") or "unknown-job-id"


def get_user() -> str:
    return os.environ.get("USER", "unknown-user")


def run_experiment(exp: Dict, reservations, user_script: str, user_args: List[str]) -> int:
    """Run ONE experiment to completion on the reserved (host, slots) through this package's launcher - the multi-node
    form of ``ResourceManager.run_job``: the config is written to ``<result_dir>/ds_config.json`` (with the autotuning
    block pointing its metric file there), the launcher gets ``--include host:slots@...``, output goes to
    ``stdout.log`` / ``stderr.log``. Returns the launcher's exit code; a non-zero code is also recorded in ``stderr.log``."""
    include = "@".join(f"{r.node.host}:{','.join(map(str, sorted(r.slots)))}" for r in reservations)
    os.makedirs(exp["result_dir"], exist_ok=True)
    cfg = dict(exp["ds_config"])
    at = dict(cfg.get("autotuning") or {})
    at.setdefault("enabled", True)
    at.setdefault("metric_path", os.path.join(exp["result_dir"], "metrics.json"))
    at.setdefault("model_info_path", os.path.join(exp["result_dir"], "model_info.json"))
    cfg["autotuning"] = at
    cfg_path = os.path.join(exp["result_dir"], "ds_config.json")
    with open(cfg_path, "w") as f:
        json.dump(cfg, f)
    with open(os.path.join(exp["result_dir"], "exp.json"), "w") as f:
        json.dump({k: v for k, v in exp.items() if k not in ("process", "ds_config")} | {"ds_config_path": cfg_path}, f)
    args = list(user_args)
    if "--deepspeed_config" in args:
        args[args.index("--deepspeed_config") + 1] = cfg_path
    else:
        args += ["--deepspeed_config", cfg_path]
    cmd = [sys.executable, "-m", __package__.rsplit(".", 1)[0] + ".launcher.runner"]
    if exp.get("hostfile"):
        cmd += ["--hostfile", str(exp["hostfile"])]
    cmd += ["--include", include, "--master_port", str(exp.get("master_port", 29500))] + list(exp.get("launcher_args") or [])
    cmd += [user_script] + args
    exp["launcher_args"] = cmd[3:cmd.index(user_script)]
    with open(os.path.join(exp["result_dir"], "stdout.log"), "w") as out, open(os.path.join(exp["result_dir"], "stderr.log"), "w") as err:
        proc = subprocess.Popen(cmd, stdout=out, stderr=err, start_new_session=True)
        exp["pid"] = proc.pid
        rc = proc.wait()
        if rc != 0:
            err.write(f"\\nlauncher exited with code {rc}\\n")
    clean_up(exp, reservations)
    return rc


def clean_up(exp: Dict, reservations) -> None:
    """Make sure nothing of a finished (or abandoned) experiment keeps running: the launcher was started as its own
    session, so its whole process group - launcher, per-node launch, ranks - is signalled by id. (The reference sweeps
    the nodes with a name-pattern kill; an id cannot hit a process that is not the experiment's.)"""
    import signal
    pid = exp.get("pid")
    if not pid:
        return
    for sig in (signal.SIGTERM, signal.SIGKILL):
        try:
            os.killpg(pid, sig)
        except (ProcessLookupError, PermissionError):
            return
        time.sleep(0.2)
