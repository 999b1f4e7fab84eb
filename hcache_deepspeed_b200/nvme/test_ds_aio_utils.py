"""Helpers of the aio benchmark (``deepspeed/nvme/test_ds_aio_utils.py``)."""
BYTES_PER_GB = 1024 ** 3
LOG_TIDS = [0]


def task_log(tid, msg, force=False):
    if force or tid in LOG_TIDS:
        print(f"tid {tid}: {msg}")


def task_barrier(barrier, num_parties):
    assert barrier.parties == num_parties
    barrier.wait()
    assert not barrier.broken


def report_results(args, read_op, pool_results):
    io = "Read" if read_op else "Write"
    if None in pool_results:
        print(f"Failure in one of {len(pool_results)} {io} processes")
        return
    total_bytes = sum(r[2] for r in pool_results)
    task_latency = max(r[0] for r in pool_results)
    e2e_latency = max(r[1] for r in pool_results)
    print(f"Task {io} Latency = {task_latency} sec")
    print(f"E2E {io} Speed = {total_bytes / e2e_latency / BYTES_PER_GB} GB/sec")


def refine_integer_value(value):
    mul = {"K": 1024, "M": 1024 ** 2, "G": 1024 ** 3}
    return int(value[:-1]) * mul[value[-1]] if value[-1] in mul else int(value)


def refine_args(args):
    for name in ("io_size", "block_size"):
        v = getattr(args, name, None)
        if isinstance(v, str):
            setattr(args, name, refine_integer_value(v))
