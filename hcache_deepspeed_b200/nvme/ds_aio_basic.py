"""The reference also benchmarks a handle-less API here; this build has the handle API only."""
from .ds_aio_handle import aio_handle_multiprocessing as aio_basic_multiprocessing  # noqa: F401
