"""``aio`` key names (``deepspeed/runtime/swap_tensor/constants.py``)."""
AIO = "aio"
AIO_BLOCK_SIZE, AIO_QUEUE_DEPTH, AIO_INTRA_OP_PARALLELISM, AIO_SINGLE_SUBMIT, AIO_OVERLAP_EVENTS, AIO_USE_GDS = (
    "block_size", "queue_depth", "intra_op_parallelism", "single_submit", "overlap_events", "use_gds")
AIO_BLOCK_SIZE_DEFAULT, AIO_QUEUE_DEPTH_DEFAULT, AIO_INTRA_OP_PARALLELISM_DEFAULT = 1048576, 8, 1
AIO_SINGLE_SUBMIT_DEFAULT, AIO_OVERLAP_EVENTS_DEFAULT, AIO_USE_GDS_DEFAULT = False, True, False
