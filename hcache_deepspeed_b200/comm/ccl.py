"""oneCCL transport (``deepspeed/comm/ccl.py``) — a CPU/XPU back end. This framework targets B200 GPUs over NCCL; host
collectives go through gloo or the shared-memory all-reduce (:mod:`.shm`). The class exists so that configuration that
names it fails with a clear message."""
from .backend import Backend


class CCLBackend(Backend):

    def __init__(self, *args, **kwargs):
        raise RuntimeError("the oneCCL back end is not part of this build: use 'nccl' (GPU) or 'gloo' / shared-memory all-reduce (host)")
