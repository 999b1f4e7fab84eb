"""``deepspeed.comm``-compatible facade over ``torch.distributed``.

One process per GPU; NCCL (NVLink 5 / NVSwitch) on CUDA, gloo on CPU. Mirrors the module-level API of
``deepspeed/comm/comm.py:222-806`` (same function names and argument order) including ``timed_op`` /
``CommsLogger`` accounting and ``inference_all_reduce``. Hot-path tensor-parallel reductions do NOT go through
here: they are fused into GEMM epilogues over symmetric memory (``parallel/tp_comm.py``); this facade is the
bootstrap / cold-path plumbing and the NCCL baseline.
"""
from __future__ import annotations

import functools

import os
from datetime import timedelta
from typing import Any, List, Optional

import torch
import torch.distributed as tdist

from ..utils.logging import logger
from .comms_logging import CommsLogger, get_caller_func
from .reduce_op import ReduceOp

comms_logger = CommsLogger()
cdb = None  # kept for source compatibility: truthy once initialised
DEFAULT_TIMEOUT_MIN = int(os.getenv("DEEPSPEED_TIMEOUT", os.getenv("HDS_TIMEOUT", 30)))
default_pg_timeout = timedelta(minutes=DEFAULT_TIMEOUT_MIN)

_REDUCE = {ReduceOp.SUM: tdist.ReduceOp.SUM, ReduceOp.PRODUCT: tdist.ReduceOp.PRODUCT, ReduceOp.MIN: tdist.ReduceOp.MIN,
           ReduceOp.MAX: tdist.ReduceOp.MAX, ReduceOp.BAND: tdist.ReduceOp.BAND, ReduceOp.BOR: tdist.ReduceOp.BOR,
           ReduceOp.BXOR: tdist.ReduceOp.BXOR, ReduceOp.AVG: tdist.ReduceOp.AVG}


def _op(op):
    return _REDUCE.get(op, op)


def _msg_size(args, kwargs) -> int:
    for a in list(args) + list(kwargs.values()):
        if isinstance(a, torch.Tensor):
            return a.numel() * a.element_size()
        if isinstance(a, (list, tuple)) and a and isinstance(a[0], torch.Tensor):
            return sum(t.numel() * t.element_size() for t in a)
    return 0


def timed_op(func):
    """Record latency + message size of a collective when the comms logger is on (forces a device sync)."""

    @functools.wraps(func)
    def wrapper(*args, **kwargs):
        prof = comms_logger.enabled and (kwargs.pop("prof", False) or comms_logger.prof_all
                                         or kwargs.get("log_name", func.__name__) in comms_logger.prof_ops)
        log_name = kwargs.pop("log_name", func.__name__)
        kwargs.pop("debug", None)
        if not prof:
            return func(*args, **kwargs)
        cuda = torch.cuda.is_available()
        if cuda:
            start, end = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            start.record()
        else:
            import time
            t0 = time.perf_counter()
        try:
            return func(*args, **kwargs)
        finally:
            if cuda:
                end.record()
                end.synchronize()
                ms = start.elapsed_time(end)
            else:
                ms = (time.perf_counter() - t0) * 1e3
            name = log_name if not comms_logger.debug else f"{log_name} | [Caller Func: {get_caller_func()}]"
            comms_logger.append(func.__name__, name, ms, _msg_size(args, kwargs), get_world_size(kwargs.get("group")))

    wrapper.__name__ = func.__name__
    wrapper.__doc__ = func.__doc__
    return wrapper


def configure(deepspeed_config=None, enabled=None, prof_all=None, prof_ops=None, verbose=None, debug=None):
    comms_logger.configure(getattr(deepspeed_config, "comms_config", None), enabled, prof_all, prof_ops, verbose, debug)


def log_summary(show_straggler: bool = False):
    if is_initialized():
        barrier(log_name="log_summary_barrier")
    if get_rank() == 0:
        comms_logger.log_all(print_log=True, show_straggler=show_straggler)
    elif show_straggler:
        comms_logger.log_all(print_log=False, show_straggler=True)


# ---- state ------------------------------------------------------------------------------------------
def is_initialized() -> bool:
    return tdist.is_available() and tdist.is_initialized()


def is_available() -> bool:
    return tdist.is_available()


def get_rank(group=None) -> int:
    return tdist.get_rank(group=group) if is_initialized() else int(os.environ.get("RANK", 0))


def get_world_size(group=None) -> int:
    return tdist.get_world_size(group=group) if is_initialized() else int(os.environ.get("WORLD_SIZE", 1))


def get_local_rank() -> int:
    return int(os.environ.get("LOCAL_RANK", 0))


def get_global_rank(group=None, group_rank: int = 0) -> int:
    if group is None or not is_initialized():
        return group_rank
    return tdist.get_global_rank(group, group_rank)


def get_world_group():
    return tdist.group.WORLD


def get_all_ranks_from_group(group=None) -> List[int]:
    n = get_world_size(group)
    return [get_global_rank(group, r) for r in range(n)]


def new_group(ranks=None, **kwargs):
    return tdist.new_group(ranks=ranks, **kwargs)


def destroy_process_group(group=None):
    if is_initialized():
        tdist.destroy_process_group(group)


def set_backend():
    return None


def has_all_reduce_coalesced() -> bool:
    return hasattr(tdist, "all_reduce_coalesced")


def has_coalescing_manager() -> bool:
    return hasattr(tdist.distributed_c10d, "_coalescing_manager")


def all_gather_coalesced_available() -> bool:
    return False


# ---- collectives ---------------------------------------------------------------------------------------
@timed_op
def broadcast(tensor, src, group=None, async_op=False):
    eng = _host_engine(group, tensor)
    if eng is not None:
        eng.broadcast(tensor, src if group is None else tdist.get_group_rank(group, src))
        return _Done() if async_op else None
    return tdist.broadcast(tensor, src, group=group, async_op=async_op)


@timed_op
def all_gather(tensor_list, tensor, group=None, async_op=False):
    eng = _host_engine(group, tensor)
    if eng is not None and all(t.shape == tensor.shape and t.dtype == tensor.dtype for t in tensor_list):
        flat = torch.empty((len(tensor_list), ) + tuple(tensor.shape), dtype=tensor.dtype)
        eng.all_gather_into_tensor(flat, tensor)
        for dst, src in zip(tensor_list, flat):
            dst.copy_(src)
        return _Done() if async_op else None
    return tdist.all_gather(tensor_list, tensor, group=group, async_op=async_op)


@timed_op
def all_gather_into_tensor(output_tensor, tensor, group=None, async_op=False):
    eng = _host_engine(group, output_tensor, tensor)
    if eng is not None:
        eng.all_gather_into_tensor(output_tensor, tensor)
        return _Done() if async_op else None
    return tdist.all_gather_into_tensor(output_tensor, tensor, group=group, async_op=async_op)


def allgather_fn(output_tensor, input_tensor, group=None, async_op=False, debug=None):
    return all_gather_into_tensor(output_tensor, input_tensor, group=group, async_op=async_op)


def all_gather_coalesced(output_tensors, input_tensors, group=None, async_op=False):
    handles = [tdist.all_gather_into_tensor(o, i, group=group, async_op=async_op) for o, i in zip(output_tensors, input_tensors)]
    return handles[-1] if handles else None


@timed_op
def reduce_scatter_tensor(output_tensor, tensor, op=ReduceOp.SUM, group=None, async_op=False):
    eng = _host_engine(group, tensor, output_tensor, op=op)
    if eng is not None:
        eng.reduce_scatter_tensor(output_tensor, tensor, op)
        return _Done() if async_op else None
    return tdist.reduce_scatter_tensor(output_tensor, tensor, op=_op(op), group=group, async_op=async_op)


def reduce_scatter_fn(output_tensor, tensor, op=ReduceOp.SUM, group=None, async_op=False, debug=None):
    return reduce_scatter_tensor(output_tensor, tensor, op=op, group=group, async_op=async_op)


@timed_op
def reduce_scatter(output, input_list, op=ReduceOp.SUM, group=None, async_op=False):
    return tdist.reduce_scatter(output, input_list, op=_op(op), group=group, async_op=async_op)


@timed_op
def all_to_all_single(output, tensor, output_split_sizes=None, input_split_sizes=None, group=None, async_op=False):
    return tdist.all_to_all_single(output, tensor, output_split_sizes=output_split_sizes,
                                   input_split_sizes=input_split_sizes, group=group, async_op=async_op)


@timed_op
def all_to_all(output_tensor_list, input_tensor_list, group=None, async_op=False):
    if tdist.get_backend(group) != "gloo":
        return tdist.all_to_all(output_tensor_list, input_tensor_list, group=group, async_op=async_op)
    # gloo has no list all-to-all: exchange with non-blocking point-to-point pairs
    ws, me = tdist.get_world_size(group), tdist.get_rank(group)
    output_tensor_list[me].copy_(input_tensor_list[me])
    reqs = []
    for r in range(ws):
        if r == me:
            continue
        peer = tdist.get_global_rank(group, r) if group is not None else r
        reqs.append(tdist.isend(input_tensor_list[r].contiguous(), peer, group=group))
        reqs.append(tdist.irecv(output_tensor_list[r], peer, group=group))
    for q in reqs:
        q.wait()
    return None


@timed_op
def send(tensor, dst, group=None, tag=0):
    return tdist.send(tensor, dst, group=group, tag=tag)


@timed_op
def recv(tensor, src=None, group=None, tag=0):
    return tdist.recv(tensor, src=src, group=group, tag=tag)


@timed_op
def isend(tensor, dst, group=None, tag=0):
    return tdist.isend(tensor, dst, group=group, tag=tag)


@timed_op
def irecv(tensor, src=None, group=None, tag=0):
    return tdist.irecv(tensor, src=src, group=group, tag=tag)


@timed_op
def gather(tensor, gather_list=None, dst=0, group=None, async_op=False):
    return tdist.gather(tensor, gather_list=gather_list, dst=dst, group=group, async_op=async_op)


@timed_op
def scatter(tensor, scatter_list=None, src=0, group=None, async_op=False):
    return tdist.scatter(tensor, scatter_list=scatter_list, src=src, group=group, async_op=async_op)


@timed_op
def barrier(group=None, async_op=False, device_ids=None):
    eng = _host_engine(group) if device_ids is None else None
    if eng is not None:
        eng.barrier()
        return _Done() if async_op else None
    return tdist.barrier(group=group, async_op=async_op, device_ids=device_ids)


@timed_op
def monitored_barrier(group=None, timeout=None, wait_all_ranks=False):
    return tdist.monitored_barrier(group=group, timeout=timeout, wait_all_ranks=wait_all_ranks)


@timed_op
def reduce(tensor, dst, op=ReduceOp.SUM, group=None, async_op=False):
    return tdist.reduce(tensor, dst, op=_op(op), group=group, async_op=async_op)


class _Done:
    """Work handle of a collective that has already completed (the shared-memory engine is synchronous)."""

    def wait(self, timeout=None) -> bool:
        return True

    def is_completed(self) -> bool:
        return True


def _host_engine(group, *tensors, op=None, always: bool = False):
    """The shared-memory engine attached to ``group`` when it can carry this call (contiguous CPU tensors, a dtype / op it
    has), else None. General collectives use it only after a ``CCLBackend`` switched routing on; ``inference_all_reduce``
    uses it whenever a group is attached."""
    from . import shm
    if not (always or shm.route_all_collectives):
        return None
    eng = shm.attached(group)
    if eng is None or any(t.device.type != "cpu" or not t.is_contiguous() for t in tensors):
        return None
    if op is not None and not eng.supports(tensors[0], op):
        return None
    return eng


@timed_op
def all_reduce(tensor, op=ReduceOp.SUM, group=None, async_op=False):
    eng = _host_engine(group, tensor, op=op)
    if eng is not None:
        eng.all_reduce(tensor, op)
        return _Done() if async_op else None
    return tdist.all_reduce(tensor, op=_op(op), group=group, async_op=async_op)


@timed_op
def inference_all_reduce(tensor, op=ReduceOp.SUM, group=None):
    """Latency-oriented all-reduce of the inference path (ref: comm.py:506). On CUDA this is NCCL (NVLS when
    the switch offers it); the engine's own hot path bypasses it with fused GEMM epilogues. CPU fp32 / bf16 tensors
    go through the shared-memory engine once ``comm.shm.attach(group)`` has been called (ref: torch.py:160-166)."""
    eng = _host_engine(group, tensor, op=op, always=True)
    if eng is not None:
        return eng.all_reduce(tensor, op)
    return tdist.all_reduce(tensor, op=_op(op), group=group, async_op=False)


@timed_op
def all_reduce_coalesced(tensors, op=ReduceOp.SUM, group=None, async_op=False):
    handles = [tdist.all_reduce(t, op=_op(op), group=group, async_op=async_op) for t in tensors]
    return handles[-1] if handles else None


def initialize_mesh_device(mesh_shape, mesh_dim_names):
    from torch.distributed.device_mesh import init_device_mesh
    dev = "cuda" if torch.cuda.is_available() else "cpu"
    return init_device_mesh(dev, mesh_shape, mesh_dim_names=mesh_dim_names)


def enable_symm_mem_for_group(group_name: str):
    from torch.distributed._symmetric_memory import enable_symm_mem_for_group as _en
    return _en(group_name)


# ---- bootstrap ---------------------------------------------------------------------------------------
def mpi_discovery(distributed_port: int = 29500, verbose: bool = True):
    """Fill RANK / WORLD_SIZE / MASTER_* from an MPI launch (ref: comm.py:705)."""
    from mpi4py import MPI  # noqa: import error propagates: mpi4py is optional
    import subprocess
    comm = MPI.COMM_WORLD
    rank, world = comm.Get_rank(), comm.Get_size()
    master = None
    if rank == 0:
        try:
            master = subprocess.check_output(["hostname", "-I"]).decode().split()[0]
        except (subprocess.CalledProcessError, OSError, IndexError):
            import socket
            master = socket.gethostbyname(socket.gethostname())
    master = comm.bcast(master, root=0)
    proc_name = MPI.Get_processor_name()
    all_procs = comm.allgather(proc_name)
    local_rank = sum(1 for p in all_procs[:rank] if p == proc_name)
    os.environ.update({"RANK": str(rank), "WORLD_SIZE": str(world), "LOCAL_RANK": str(local_rank),
                       "MASTER_ADDR": master, "MASTER_PORT": str(distributed_port)})
    if verbose:
        logger.info(f"Discovered MPI settings of world_rank={rank}, local_rank={local_rank}, world_size={world}")


_RANK_ENV_SOURCES = (
    # (rank, local rank, world size, local size) as exported by each process manager
    ("OMPI_COMM_WORLD_RANK", "OMPI_COMM_WORLD_LOCAL_RANK", "OMPI_COMM_WORLD_SIZE", "OMPI_COMM_WORLD_LOCAL_SIZE"),
    ("PMI_RANK", "MPI_LOCALRANKID", "PMI_SIZE", "MPI_LOCALNRANKS"),                       # MPICH / Intel MPI (hydra)
    ("PMIX_RANK", "PMIX_LOCAL_RANK", "PMIX_SIZE", "PMIX_LOCAL_SIZE"),
    ("MV2_COMM_WORLD_RANK", "MV2_COMM_WORLD_LOCAL_RANK", "MV2_COMM_WORLD_SIZE", "MV2_COMM_WORLD_LOCAL_SIZE"),
    ("SLURM_PROCID", "SLURM_LOCALID", "SLURM_NTASKS", "SLURM_NTASKS_PER_NODE"),
)
AML_MASTER_PORT = 54965
AML_NCCL_SOCKET_IFNAME = "^docker0,lo"


def process_manager_env(env=None):
    """(rank, local_rank, world, local_size) from the first process manager whose variables are complete, or None.
    No MPI library is needed for this: mpirun / srun export the numbers to every rank."""
    env = os.environ if env is None else env
    for rank_k, local_k, world_k, lsize_k in _RANK_ENV_SOURCES:
        if rank_k in env and world_k in env:
            rank, world = int(env[rank_k]), int(env[world_k])
            local = int(env[local_k]) if local_k in env else rank
            try:
                local_size = int(str(env.get(lsize_k, world)).split("(")[0])     # SLURM writes "8(x2)"
            except ValueError:
                local_size = world
            return rank, local, world, local_size
    return None


def _export_ranks(found):
    rank, local, world, _ = found
    os.environ.update({"RANK": str(rank), "LOCAL_RANK": str(local), "WORLD_SIZE": str(world)})


def patch_aml_env_for_torch_nccl_backend(master_port: int = 6105, verbose: bool = True):
    """AzureML launches ranks with OpenMPI and names the first node in ``AZ_BATCH_MASTER_NODE`` (host:port; several
    nodes) or ``AZ_BATCHAI_MPI_MASTER_NODE`` (one node). (ref: comm.py:765)"""
    found = process_manager_env()
    if found is None:
        raise RuntimeError("AzureML job without OpenMPI rank variables: cannot derive RANK / WORLD_SIZE")
    _export_ranks(found)
    if found[3] < found[2]:      # more than one node
        os.environ["MASTER_ADDR"] = os.environ["AZ_BATCH_MASTER_NODE"].split(":")[0]
        os.environ.setdefault("MASTER_PORT", str(master_port))
    else:
        os.environ["MASTER_ADDR"] = os.environ.get("AZ_BATCHAI_MPI_MASTER_NODE", "127.0.0.1")
        os.environ["MASTER_PORT"] = str(AML_MASTER_PORT)
    os.environ["NCCL_SOCKET_IFNAME"] = AML_NCCL_SOCKET_IFNAME
    if verbose:
        logger.info("Discovered AzureML settings of world_rank={RANK}, local_rank={LOCAL_RANK}, world_size={WORLD_SIZE}, "
                    "master_addr={MASTER_ADDR}, master_port={MASTER_PORT}".format(**os.environ))


def patch_aws_sm_env_for_torch_nccl_backend(verbose: bool = True):
    """SageMaker: ranks come from OpenMPI; the first host of ``SM_TRAINING_ENV`` is the rendezvous address when the
    job did not export one. (ref: comm.py:797)"""
    found = process_manager_env()
    if found is None:
        raise RuntimeError("SageMaker job without OpenMPI rank variables: cannot derive RANK / WORLD_SIZE")
    _export_ranks(found)
    if "MASTER_ADDR" not in os.environ:
        import json
        try:
            hosts = json.loads(os.environ["SM_TRAINING_ENV"]).get("hosts") or []
        except (ValueError, AttributeError):
            hosts = []
        os.environ["MASTER_ADDR"] = sorted(hosts)[0] if hosts else "127.0.0.1"
    os.environ.setdefault("MASTER_PORT", "29500")
    if verbose:
        logger.info("Discovered AWS SageMaker settings of world_rank={RANK}, local_rank={LOCAL_RANK}, "
                    "world_size={WORLD_SIZE}, master_addr={MASTER_ADDR}, master_port={MASTER_PORT}".format(**os.environ))


def env_discovery(distributed_port: int = 29500, verbose: bool = True) -> bool:
    """Rank variables from mpirun / srun without an MPI library. The rendezvous address must already be known
    (``MASTER_ADDR``, or a single-node job, or SLURM's node list); returns False when it is not, so that the caller
    can fall back to ``mpi_discovery`` (which broadcasts it)."""
    found = process_manager_env()
    if found is None:
        return False
    rank, local, world, local_size = found
    if "MASTER_ADDR" not in os.environ:
        if local_size >= world:
            os.environ["MASTER_ADDR"] = "127.0.0.1"
        elif "SLURM_JOB_NODELIST" in os.environ or "SLURM_NODELIST" in os.environ:
            from ..launcher.multinode_runner import first_slurm_host
            os.environ["MASTER_ADDR"] = first_slurm_host(os.environ.get("SLURM_JOB_NODELIST") or
                                                         os.environ["SLURM_NODELIST"])
        else:
            return False
    _export_ranks(found)
    os.environ.setdefault("MASTER_PORT", str(distributed_port))
    if verbose:
        logger.info(f"Discovered launcher settings of world_rank={rank}, local_rank={local}, world_size={world}, "
                    f"master_addr={os.environ['MASTER_ADDR']}")
    return True


def in_aml() -> bool:
    return "AZUREML_EXPERIMENT_ID" in os.environ


def in_aws_sm() -> bool:
    return "SM_TRAINING_ENV" in os.environ


def in_dlts() -> bool:
    return "DLTS_JOB_ID" in os.environ


def init_distributed(dist_backend: Optional[str] = None, auto_mpi_discovery: bool = True,
                     distributed_port: int = 29500, verbose: bool = True, timeout: timedelta = default_pg_timeout,
                     init_method: Optional[str] = None, dist_init_required: Optional[bool] = None, config: Any = None,
                     rank: int = -1, world_size: int = -1) -> None:
    """Initialise ``torch.distributed`` once (NCCL on CUDA, gloo otherwise). A single process without launcher
    environment becomes a world of one on 127.0.0.1."""
    global cdb
    configure(deepspeed_config=config)
    if dist_init_required is None:
        dist_init_required = not is_initialized()
    if dist_init_required is False:
        if not is_initialized():
            raise AssertionError("Distributed backend is not initialized. Please set dist_init_required to True "
                                 "or initialize before calling.")
        cdb = True
        return
    if is_initialized():
        cdb = True
        return
    required = ("RANK", "WORLD_SIZE", "MASTER_ADDR", "MASTER_PORT", "LOCAL_RANK")
    if auto_mpi_discovery and not all(k in os.environ for k in required):
        if in_aml() and not in_dlts():
            patch_aml_env_for_torch_nccl_backend(verbose=verbose)
        elif in_aws_sm():
            patch_aws_sm_env_for_torch_nccl_backend(verbose=verbose)
        elif process_manager_env() is not None and not env_discovery(distributed_port, verbose):
            # several nodes and nobody exported the rendezvous address: rank 0 has to broadcast it through MPI
            mpi_discovery(distributed_port=distributed_port, verbose=verbose)
    os.environ.setdefault("RANK", str(rank if rank >= 0 else 0))
    os.environ.setdefault("WORLD_SIZE", str(world_size if world_size > 0 else 1))
    os.environ.setdefault("LOCAL_RANK", os.environ["RANK"])
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    os.environ.setdefault("MASTER_PORT", str(distributed_port))
    if dist_backend is None:
        dist_backend = "nccl" if torch.cuda.is_available() else "gloo"
    if verbose and int(os.environ["RANK"]) == 0:
        logger.info(f"Initializing torch.distributed with backend: {dist_backend}")
    kwargs = {}
    if dist_backend == "nccl" and torch.cuda.is_available():
        local = int(os.environ["LOCAL_RANK"]) % max(1, torch.cuda.device_count())
        torch.cuda.set_device(local)
        kwargs["device_id"] = torch.device("cuda", local)
    if dist_backend == "ccl":
        # host collectives: gloo underneath, this framework's shared-memory engine for what it covers (comm/ccl.py)
        from .ccl import CCLBackend
        cdb = CCLBackend(rank=int(os.environ["RANK"]), world_size=int(os.environ["WORLD_SIZE"]), timeout=timeout,
                         init_method=init_method)
        return
    tdist.init_process_group(backend=dist_backend, init_method=init_method, timeout=timeout,
                             rank=int(os.environ["RANK"]), world_size=int(os.environ["WORLD_SIZE"]), **kwargs)
    cdb = True


# ---- remaining names of the reference's comm module (comm/comm.py) ---------------------------------------------------------
ProcessGroup = torch.distributed.ProcessGroup


def has_all_gather_into_tensor() -> bool:
    return hasattr(torch.distributed, "all_gather_into_tensor")


def has_reduce_scatter_tensor() -> bool:
    return hasattr(torch.distributed, "reduce_scatter_tensor")


def broadcast_object_list(object_list, src, group=None, device=None):
    """Pickled Python objects from ``src`` to every rank of ``group`` (in place in ``object_list``)."""
    return torch.distributed.broadcast_object_list(object_list, src=src, group=group, device=device)


def init_deepspeed_backend(ds_backend=None, timeout=None, init_method=None):
    """The reference selects one of its own (non-torch) backends here; this framework has one GPU backend (NCCL through
    torch.distributed) plus the host engine, both reached through ``init_distributed`` - so this is that call."""
    kw = {} if timeout is None else {"timeout": timeout}
    return init_distributed(dist_backend=ds_backend or ("nccl" if torch.cuda.is_available() else "gloo"),
                            init_method=init_method, **kw)
