"""Point-to-point transfers between adjacent pipeline stages.
NCCL: ``send`` / ``recv`` on the world group (NVLink peer copies inside a node). gloo: the same calls.
Reference: ``deepspeed/runtime/pipe/p2p.py`` (init_process_groups :31, send :50, recv :72, send_obj / recv_obj :100-165)."""
from __future__ import annotations

import pickle

import torch

from ... import comm as dist

_grid = None


def init_process_groups(grid):
    global _grid
    _grid = grid
    assert _grid.pipe_parallel_size > 1, "There is no pipeline parallelism"


def _check(src_stage: int, dest_stage: int):
    n = _grid.pipe_parallel_size
    first, last = 0, n - 1
    assert abs(src_stage - dest_stage) == 1 or (src_stage == first and dest_stage == last) or \
        (src_stage == last and dest_stage == first), "Functionality currently limited to send and receive between adjacent ranks only"


def send(tensor: torch.Tensor, dest_stage: int, async_op: bool = False):
    _check(_grid.get_stage_id(), dest_stage)
    dst = _grid.stage_to_global(stage_id=dest_stage)
    if async_op:
        return dist.isend(tensor, dst)
    return dist.send(tensor, dst)


def recv(tensor: torch.Tensor, src_stage: int, async_op: bool = False):
    _check(src_stage, _grid.get_stage_id())
    src = _grid.stage_to_global(stage_id=src_stage)
    if async_op:
        return dist.irecv(tensor, src)
    return dist.recv(tensor, src)


def wait():
    if torch.cuda.is_available():
        torch.cuda.synchronize()


def send_obj(obj, dest_stage: int, device="cpu"):
    """Pickle ``obj`` and send (length first)."""
    data = torch.frombuffer(bytearray(pickle.dumps(obj)), dtype=torch.uint8).to(device)
    send(torch.tensor([data.numel()], dtype=torch.long, device=device), dest_stage)
    send(data, dest_stage)


def recv_obj(src_stage: int, device="cpu"):
    n = torch.zeros(1, dtype=torch.long, device=device)
    recv(n, src_stage)
    data = torch.empty(int(n[0]), dtype=torch.uint8, device=device)
    recv(data, src_stage)
    return pickle.loads(bytes(data.cpu().numpy().tobytes()))


def can_send_recv() -> bool:
    """Point-to-point ``send`` / ``recv`` (rather than two-rank broadcasts) is available in every torch this runs on."""
    return hasattr(dist, "send") and hasattr(dist, "recv")
