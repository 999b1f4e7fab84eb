"""``deepspeed.ops.gds``: GPUDirect Storage operator (cuFile binding, present only where libcufile is)."""
from ...io.gds import GDSHandle, gds_available  # noqa: F401
from ..op_builder import GDSBuilder  # noqa: F401
