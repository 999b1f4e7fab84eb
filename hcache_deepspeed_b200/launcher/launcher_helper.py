"""Start a user script under an MPI-style launcher (``deepspeed/launcher/launcher_helper.py:14-110``): the outer
launcher (mpirun / srun / impi) exports ITS rank variables; this helper maps them onto ``RANK`` / ``LOCAL_RANK`` and
execs the script, optionally pinning the process to a core range."""
from __future__ import annotations

import argparse
import os
import subprocess
import sys
from typing import List, Optional

RANK_NAMES = ["OMPI_COMM_WORLD_RANK", "PMI_RANK", "PMIX_RANK", "SLURM_PROCID", "MV2_COMM_WORLD_RANK"]
LOCAL_RANK_NAMES = ["OMPI_COMM_WORLD_LOCAL_RANK", "MPI_LOCALRANKID", "SLURM_LOCALID", "MV2_COMM_WORLD_LOCAL_RANK"]


def parse_args(args=None):
    p = argparse.ArgumentParser(description="launcher helper: map the outer launcher's environment and start the user script")
    p.add_argument("--launcher", type=str, default="MPICH", help="outer launcher: MPICH, IMPI, OPENMPI, SLURM, MVAPICH")
    p.add_argument("--module", action="store_true", help="run the script as `python -m`")
    p.add_argument("--no_python", action="store_true", help="exec the script directly")
    p.add_argument("--bind_cores_to_rank", action="store_true")
    p.add_argument("--bind_core_list", type=str, default=None)
    p.add_argument("user_script", type=str)
    p.add_argument("user_args", nargs=argparse.REMAINDER)
    return p.parse_args(args=args)


def env_mapping(env: dict, rank_name_list: Optional[List[str]] = None, local_rank_name_list: Optional[List[str]] = None) -> dict:
    """Set ``RANK`` / ``LOCAL_RANK`` from whichever launcher variables are present; they must agree with each other."""
    def pick(names, what):
        found = {n: env[n] for n in names if n in env}
        if not found:
            raise EnvironmentError(f"{what} could not be determined: none of {names} is set")
        if len(set(found.values())) > 1:
            raise EnvironmentError(f"{what} is ambiguous, the launcher variables disagree: {found}")
        return next(iter(found.values()))

    env["RANK"] = pick(rank_name_list or RANK_NAMES, "rank")
    env["LOCAL_RANK"] = pick(local_rank_name_list or LOCAL_RANK_NAMES, "local rank")
    return env


def build_command(args) -> List[str]:
    cmd: List[str] = []
    if args.bind_cores_to_rank:
        from ..utils.numa import get_numactl_cmd
        local, local_size = int(os.environ["LOCAL_RANK"]), int(os.environ.get("LOCAL_SIZE", 1))
        _, numa = get_numactl_cmd(args.bind_core_list, local_size, local)
        cmd += numa
    if not args.no_python:
        cmd += [sys.executable, "-u"] + (["-m"] if args.module else [])
    elif args.module:
        raise ValueError("Don't use both the '--no_python' flag and the '--module' flag at the same time.")
    return cmd + [args.user_script] + list(args.user_args)


def main(args=None) -> int:
    args = parse_args(args)
    env_mapping(os.environ)
    return subprocess.Popen(build_command(args), env=os.environ.copy()).wait()


if __name__ == "__main__":
    sys.exit(main())
