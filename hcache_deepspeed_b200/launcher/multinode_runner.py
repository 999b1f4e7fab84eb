"""Multi-node launch back ends: each builds the command line that starts ``launcher.launch`` (or the user script
directly, for MPI-style launchers that spawn one process per rank) on every node.
Reference: ``deepspeed/launcher/multinode_runner.py`` (PDSHRunner :51, OpenMPIRunner :117, MPICHRunner :196,
IMPIRunner :252, SlurmRunner :334, MVAPICHRunner :385)."""
from __future__ import annotations

import os
import shutil
import sys
from abc import ABC, abstractmethod
from typing import Dict, List

PDSH_MAX_FAN_OUT = 1024


class MultiNodeRunner(ABC):

    def __init__(self, args, world_info_base64: str):
        self.args = args
        self.user_arguments = list(getattr(args, "user_args", []))
        self.user_script = args.user_script
        self.world_info_base64 = world_info_base64
        self.exports: Dict[str, str] = {}

    @abstractmethod
    def backend_exists(self) -> bool:
        ...

    @abstractmethod
    def get_cmd(self, environment: Dict[str, str], active_resources: Dict[str, List[int]]) -> List[str]:
        ...

    def parse_user_args(self) -> List[str]:
        """The user script's arguments as they go on the remote command line (PDSH quotes those with spaces)."""
        return list(self.user_arguments)

    def add_export(self, key: str, var: str):
        self.exports[key.strip()] = var.strip()

    @property
    def name(self) -> str:
        return self.__class__.__name__

    def validate_args(self):
        pass

    def _launch_module(self) -> List[str]:
        return [sys.executable, "-u", "-m", "hcache_deepspeed_b200.launcher.launch", f"--world_info={self.world_info_base64}",
                f"--master_addr={self.args.master_addr}", f"--master_port={self.args.master_port}"]


class PDSHRunner(MultiNodeRunner):

    def backend_exists(self) -> bool:
        return shutil.which("pdsh") is not None

    @property
    def name(self):
        return "pdsh"

    def get_cmd(self, environment, active_resources):
        environment["PDSH_RCMD_TYPE"] = "ssh"
        hosts = ",".join(active_resources.keys())
        exports = "".join(f"export {k}={v}; " for k, v in self.exports.items())
        launch = self._launch_module() + ["--node_rank=%n"]
        return ["pdsh", "-S", "-f", str(PDSH_MAX_FAN_OUT), "-w", hosts, exports + f"cd {os.path.abspath('.')};",
                *launch, self.user_script, *self.parse_user_args()]

    def parse_user_args(self) -> List[str]:
        """pdsh hands the command to a remote shell, which would split an argument at its spaces: such arguments are
        wrapped in double quotes (embedded quotes escaped)."""
        out = []
        for arg in self.user_arguments:
            if " " in arg:
                arg = '"{}"'.format(arg.replace('"', '\\"'))
            out.append(arg)
        return out


class _MPIStyle(MultiNodeRunner):
    """Launchers that start one process per rank themselves: the user script runs directly and reads its rank from the
    MPI environment (``comm.mpi_discovery``)."""

    def _total_procs(self, active_resources) -> int:
        return sum(len(v) for v in active_resources.values())

    def _exports(self, flag: str, sep: str = "=") -> List[str]:
        out = []
        for k, v in self.exports.items():
            out += [flag, k, v] if sep == " " else [flag, f"{k}{sep}{v}"]
        return out

    def _script(self) -> List[str]:
        py = [] if getattr(self.args, "no_python", False) else [sys.executable, "-u"]
        if getattr(self.args, "module", False):
            py.append("-m")
        return py + [self.user_script] + self.user_arguments


class OpenMPIRunner(_MPIStyle):

    def __init__(self, args, world_info_base64, resource_pool=None):
        super().__init__(args, world_info_base64)
        self.resource_pool = resource_pool
        self.add_export("UCX_TLS", "tcp")

    def backend_exists(self):
        return shutil.which("ompi_info") is not None

    @property
    def name(self):
        return "openmpi"

    def validate_args(self):
        if getattr(self.args, "include", "") or getattr(self.args, "exclude", ""):
            raise ValueError(f"{self.name} backend does not support worker include/exclusion")
        if getattr(self.args, "num_nodes", -1) != -1 or getattr(self.args, "num_gpus", -1) != -1:
            raise ValueError(f"{self.name} backend does not support limiting num nodes/gpus")

    def get_cmd(self, environment, active_resources):
        cmd = ["mpirun", "-n", str(self._total_procs(active_resources)), "-hostfile", str(self.args.hostfile), "--mca", "btl",
               "^openib", "--mca", "btl_tcp_if_include", "eth0"]
        cmd += getattr(self.args, "launcher_args", "").split()
        return cmd + self._exports("-x") + self._script()


class MPICHRunner(_MPIStyle):

    def backend_exists(self):
        return shutil.which("mpirun") is not None

    @property
    def name(self):
        return "mpich"

    def get_cmd(self, environment, active_resources):
        per_node = {len(v) for v in active_resources.values()}
        if len(per_node) != 1:
            raise ValueError("MPICH requires same number of devices per node")
        ppn = per_node.pop()
        cmd = ["mpirun", "-n", str(self._total_procs(active_resources)), "-ppn", str(ppn)]
        cmd += getattr(self.args, "launcher_args", "").split()
        return cmd + self._exports("-genv", " ") + self._script()


class IMPIRunner(MPICHRunner):

    @property
    def name(self):
        return "impi"

    def get_cmd(self, environment, active_resources):
        cmd = super().get_cmd(environment, active_resources)
        return cmd[:1] + ["-hosts", ",".join(active_resources.keys())] + cmd[1:]


class SlurmRunner(_MPIStyle):

    def backend_exists(self):
        return shutil.which("sinfo") is not None

    @property
    def name(self):
        return "slurm"

    def get_cmd(self, environment, active_resources):
        assert not getattr(self.args, "detect_nvlink_pairs", False), "slurm backend does not support remapping visible devices"
        cmd = ["srun", "-n", str(self._total_procs(active_resources))] + getattr(self.args, "launcher_args", "").split()
        if getattr(self.args, "comment", ""):
            cmd += ["--comment", self.args.comment]
        if getattr(self.args, "include", ""):
            cmd += ["--include", self.args.include]
        if getattr(self.args, "exclude", ""):
            cmd += ["--exclude", self.args.exclude]
        if getattr(self.args, "num_nodes", -1) > 0:
            cmd += ["--nodes", str(self.args.num_nodes)]
        if getattr(self.args, "num_gpus", -1) > 0:
            cmd += ["--gpus", str(self.args.num_gpus)]
        exports = "--export=ALL" + "".join(f",{k}={v}" for k, v in self.exports.items())
        return cmd + [exports] + self._script()


class MVAPICHRunner(_MPIStyle):

    def __init__(self, args, world_info_base64, resource_pool=None):
        super().__init__(args, world_info_base64)
        self.resource_pool = resource_pool
        for k, v in (("MV2_SMP_USE_CMA", "0"), ("MV2_DEBUG_SHOW_BACKTRACE", "1"), ("MV2_USE_CUDA", "1"),
                     ("MV2_SUPPORT_DL", "1"), ("MV2_ENABLE_AFFINITY", "0"), ("MV2_INTER_ALLGATHER_TUNING", "5"),
                     ("MV2_CUDA_USE_NAIVE", "0")):
            self.add_export(k, v)

    def backend_exists(self):
        return shutil.which("mpiname") is not None

    @property
    def name(self):
        return "mvapich"

    def get_cmd(self, environment, active_resources):
        per_node = {len(v) for v in active_resources.values()}
        if len(per_node) != 1:
            raise ValueError("mvapich requires same number of devices per node")
        hostfile = "/tmp/hds_mvapich_hostfile"
        with open(hostfile, "w") as f:
            f.write("\n".join(active_resources.keys()) + "\n")
        cmd = ["mpirun", "-np", str(self._total_procs(active_resources)), "-ppn", str(per_node.pop()), "--hostfile", hostfile]
        cmd += getattr(self.args, "launcher_args", "").split()
        return cmd + self._exports("-env", "=") + self._script()


RUNNERS = {"pdsh": PDSHRunner, "openmpi": OpenMPIRunner, "mpich": MPICHRunner, "impi": IMPIRunner, "slurm": SlurmRunner,
           "mvapich": MVAPICHRunner}


def expand_slurm_nodelist(nodelist: str):
    """``node[01-03,07],gpu5`` -> [node01, node02, node03, node07, gpu5] (what ``scontrol show hostnames`` prints, without
    needing scontrol on the rank's node)."""
    hosts, depth, start = [], 0, 0
    parts = []
    for i, ch in enumerate(nodelist):
        depth += ch == "["
        depth -= ch == "]"
        if ch == "," and depth == 0:
            parts.append(nodelist[start:i])
            start = i + 1
    parts.append(nodelist[start:])
    for part in filter(None, (p.strip() for p in parts)):
        lb = part.find("[")
        if lb < 0:
            hosts.append(part)
            continue
        rb = part.index("]", lb)
        prefix, suffix = part[:lb], part[rb + 1:]
        for rng in part[lb + 1:rb].split(","):
            lo, _, hi = rng.partition("-")
            if not hi:
                hosts.append(f"{prefix}{lo}{suffix}")
                continue
            width = len(lo)
            hosts.extend(f"{prefix}{n:0{width}d}{suffix}" for n in range(int(lo), int(hi) + 1))
    return hosts


def first_slurm_host(nodelist: str) -> str:
    return expand_slurm_nodelist(nodelist)[0]
