"""ctypes declarations for ``libhds_host.so`` (``include/hds_host.h``)."""
from __future__ import annotations

import ctypes
from ctypes import POINTER, Structure, c_char_p, c_float, c_int, c_int32, c_int64, c_void_p


class SeqIn(Structure):
    """Mirror of ``hds_seq_in``."""
    _fields_ = [("n_new", c_int32), ("seen", c_int32), ("n_blocks", c_int32), ("_pad", c_int32),
                ("block_ids", c_void_p)]


def declare(lib: ctypes.CDLL) -> None:
    P, I, I32, L = c_void_p, c_int, c_int32, c_int64
    lib.hds_alloc_create.argtypes = [I32]
    lib.hds_alloc_create.restype = P
    lib.hds_alloc_destroy.argtypes = [P]
    lib.hds_alloc_destroy.restype = None
    lib.hds_alloc_allocate.argtypes = [P, I32, P]
    lib.hds_alloc_allocate.restype = I
    lib.hds_alloc_free.argtypes = [P, P, I32]
    lib.hds_alloc_free.restype = I
    lib.hds_alloc_free_blocks.argtypes = [P]
    lib.hds_alloc_free_blocks.restype = I32
    lib.hds_alloc_total_blocks.argtypes = [P]
    lib.hds_alloc_total_blocks.restype = I32

    lib.hds_build_batch.argtypes = [POINTER(SeqIn), I32, I32, I32, I32, I32, I32, I32, I32, P, P, P, P, P, P, P, P]
    lib.hds_build_batch.restype = I

    lib.hds_aio_create.argtypes = [L, I32, I32, I32]
    lib.hds_aio_create.restype = P
    lib.hds_aio_destroy.argtypes = [P]
    lib.hds_aio_destroy.restype = None
    lib.hds_aio_pread.argtypes = [P, P, c_char_p, L, L]
    lib.hds_aio_pread.restype = L
    lib.hds_aio_pwrite.argtypes = [P, P, c_char_p, L, L]
    lib.hds_aio_pwrite.restype = L
    lib.hds_aio_wait.argtypes = [P, L]
    lib.hds_aio_wait.restype = L
    lib.hds_aio_poll.argtypes = [P, L]
    lib.hds_aio_poll.restype = I
    lib.hds_aio_pending.argtypes = [P]
    lib.hds_aio_pending.restype = I32
    lib.hds_aio_stats.argtypes = [P, P]
    lib.hds_aio_stats.restype = None
    lib.hds_mem_lock.argtypes = [P, L]
    lib.hds_mem_lock.restype = I
    lib.hds_mem_unlock.argtypes = [P, L]
    lib.hds_mem_unlock.restype = I

    F = c_float
    lib.hds_cpu_adam_step.argtypes = [P, P, P, P, L, F, F, F, F, F, I, I, I, P, I, I]
    lib.hds_cpu_adam_step.restype = I
    lib.hds_cpu_adagrad_step.argtypes = [P, P, P, L, F, F, F, P, I, I]
    lib.hds_cpu_adagrad_step.restype = I
    lib.hds_cpu_lion_step.argtypes = [P, P, P, L, F, F, F, F, P, I, I]
    lib.hds_cpu_lion_step.restype = I
    lib.hds_shm_open.argtypes = [ctypes.c_char_p, I, I, L]
    lib.hds_shm_open.restype = P
    lib.hds_shm_allreduce.argtypes = [P, P, L, I]
    lib.hds_shm_allreduce.restype = I
    lib.hds_shm_allreduce_op.argtypes = [P, P, L, I, I]
    lib.hds_shm_allreduce_op.restype = I
    lib.hds_shm_broadcast.argtypes = [P, P, L, I]
    lib.hds_shm_broadcast.restype = I
    lib.hds_shm_allgather.argtypes = [P, P, P, L]
    lib.hds_shm_allgather.restype = I
    lib.hds_shm_reduce_scatter.argtypes = [P, P, P, L, I, I]
    lib.hds_shm_reduce_scatter.restype = I
    lib.hds_shm_barrier.argtypes = [P]
    lib.hds_shm_barrier.restype = I
    lib.hds_shm_info.argtypes = [P, ctypes.POINTER(ctypes.c_int64)]
    lib.hds_shm_info.restype = None
    lib.hds_shm_close.argtypes = [P, I]
    lib.hds_shm_close.restype = None
