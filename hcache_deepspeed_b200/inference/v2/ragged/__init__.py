from .blocked_allocator import BlockedAllocator  # noqa: F401
from .kv_cache import BlockedKVCache, split_kv  # noqa: F401
from .manager_configs import (AllocationMode, DSStateManagerConfig, KVCacheConfig, KVCacheType,  # noqa: F401
                              MemoryConfig)
from .ragged_manager import DSStateManager  # noqa: F401
from .ragged_wrapper import RaggedBatchWrapper  # noqa: F401
from .sequence_descriptor import (BaseSequenceDescriptor, DSSequenceDescriptor,  # noqa: F401
                                  PlaceholderSequenceDescriptor)
