"""Run a function in N local ranks (gloo on CPU, NCCL on GPUs) — the role of the reference's DistributedTest
harness (tests/unit/common.py:129-485): N processes on one host, MASTER_ADDR=127.0.0.1, a free port, results
and exceptions funnelled back to the test process."""
import os
import socket
import traceback

import torch.multiprocessing as mp


def _free_port() -> int:
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _by_value(x):
    """Tensors are sent as numpy arrays: torch's fd-sharing of tensors breaks once the child has exited."""
    import torch
    if isinstance(x, torch.Tensor):
        return ("__tensor__", x.detach().cpu().float().numpy() if x.dtype in (torch.bfloat16, torch.float16) else x.detach().cpu().numpy())
    if isinstance(x, (list, tuple)):
        return type(x)(_by_value(v) for v in x)
    if isinstance(x, dict):
        return {k: _by_value(v) for k, v in x.items()}
    return x


def _restore(x):
    import torch
    if isinstance(x, tuple) and len(x) == 2 and isinstance(x[0], str) and x[0] == "__tensor__":
        return torch.from_numpy(x[1])
    if isinstance(x, (list, tuple)):
        return type(x)(_restore(v) for v in x)
    if isinstance(x, dict):
        return {k: _restore(v) for k, v in x.items()}
    return x


def _worker(rank, world, port, fn, args, q):
    os.environ.update({"RANK": str(rank), "LOCAL_RANK": str(rank), "WORLD_SIZE": str(world), "MASTER_ADDR": "127.0.0.1",
                       "MASTER_PORT": str(port)})
    try:
        q.put((rank, "ok", _by_value(fn(rank, world, *args))))
    except Exception:
        q.put((rank, "err", traceback.format_exc()))


def run_distributed(fn, world_size: int, *args, timeout: float = 300.0):
    """Returns the list of per-rank return values (index = rank); raises if any rank failed."""
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world_size, port, fn, args, q)) for r in range(world_size)]
    for p in procs:
        p.start()
    results = {}
    try:
        for _ in range(world_size):
            rank, status, payload = q.get(timeout=timeout)
            if status == "err":
                raise RuntimeError(f"rank {rank} failed:\n{payload}")
            results[rank] = _restore(payload)
    finally:
        for p in procs:
            p.join(timeout=10)
            if p.is_alive():
                p.kill()
    return [results[r] for r in range(world_size)]
