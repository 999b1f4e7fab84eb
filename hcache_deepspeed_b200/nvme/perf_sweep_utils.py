"""Names shared by the sweep scripts (``deepspeed/nvme/perf_sweep_utils.py``)."""
from .tune import BENCH_LOG_DIR, READ_SPEED, WRITE_SPEED  # noqa: F401

READ_OP_DESC, WRITE_OP_DESC = "read", "write"
READ_IO_DIR, WRITE_IO_DIR = f"{BENCH_LOG_DIR}/aio_perf_sweep/read", f"{BENCH_LOG_DIR}/aio_perf_sweep/write"
READ_LOG_DIR, WRITE_LOG_DIR = READ_IO_DIR, WRITE_IO_DIR
