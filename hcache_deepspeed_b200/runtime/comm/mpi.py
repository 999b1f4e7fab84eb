"""Transport-specific name kept for API parity; see :mod:`.compressed` (one implementation over hcache_deepspeed_b200.comm)."""
from .compressed import CompressedBackend

MpiBackend = CompressedBackend
