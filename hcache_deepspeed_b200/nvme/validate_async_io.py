"""Is the asynchronous I/O engine usable here? (``deepspeed/nvme/validate_async_io.py``)"""
from ..ops.op_builder import AsyncIOBuilder

assert AsyncIOBuilder().is_compatible(), "the aio engine (libhds_host.so) is not built: run __graft_entry__.build()"
assert AsyncIOBuilder().load() is not None
