from .aio import AsyncIOHandle, aio_handle  # noqa: F401
from .gds import GDSHandle, gds_available  # noqa: F401
