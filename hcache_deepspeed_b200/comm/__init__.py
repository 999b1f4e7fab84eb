from .comm import *  # noqa: F401,F403
from .comm import (all_gather, all_gather_into_tensor, all_reduce, all_to_all_single, barrier, broadcast,  # noqa: F401
                   comms_logger, configure, get_local_rank, get_rank, get_world_group, get_world_size,
                   inference_all_reduce, init_distributed, is_initialized, log_summary, new_group, reduce_scatter_tensor,
                   timed_op)
from .reduce_op import ReduceOp  # noqa: F401
