"""``deepspeed.ops.aio``: the asynchronous file I/O operator (C++ engine in ``ops/csrc_host/aio.cpp``)."""
from ...io.aio import AsyncIOHandle, aio_handle  # noqa: F401
from ..op_builder import AsyncIOBuilder  # noqa: F401
