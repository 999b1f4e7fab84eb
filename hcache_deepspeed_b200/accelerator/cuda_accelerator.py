"""CUDA accelerator: thin, explicit forwarding to ``torch.cuda`` (``deepspeed/accelerator/cuda_accelerator.py``)."""
from __future__ import annotations

import functools
import os

import torch

from .abstract_accelerator import DeepSpeedAccelerator


class CUDA_Accelerator(DeepSpeedAccelerator):

    def __init__(self):
        super().__init__()
        self._name = "cuda"
        self._communication_backend_name = "nccl"
        self._compile_backend = "inductor"

    def is_synchronized_device(self):
        return False

    def use_host_timers(self):
        return False

    def resolves_data_dependency(self):
        return False

    def handles_memory_backpressure(self):
        return False

    # ---- device ------------------------------------------------------------------------------------------
    def device_name(self, device_index=None):
        return "cuda" if device_index is None else f"cuda:{device_index}"

    def communication_backend_version(self):
        return torch.cuda.nccl.version()

    def device(self, device_index=None):
        return torch.device("cuda", device_index)

    def set_device(self, device_index):
        torch.cuda.set_device(device_index)

    def current_device(self):
        return torch.cuda.current_device()

    def current_device_name(self):
        return f"cuda:{torch.cuda.current_device()}"

    def device_count(self):
        return torch.cuda.device_count()

    def synchronize(self, device_index=None):
        return torch.cuda.synchronize(device_index)

    # ---- RNG ------------------------------------------------------------------------------------------------
    def random(self):
        return torch.random

    def set_rng_state(self, new_state, device_index=None):
        return torch.cuda.set_rng_state(new_state) if device_index is None else torch.cuda.set_rng_state(new_state, device_index)

    def get_rng_state(self, device_index=None):
        return torch.cuda.get_rng_state() if device_index is None else torch.cuda.get_rng_state(device_index)

    def manual_seed(self, seed):
        return torch.cuda.manual_seed(seed)

    def manual_seed_all(self, seed):
        return torch.cuda.manual_seed_all(seed)

    def initial_seed(self):
        return torch.cuda.initial_seed()

    def default_generator(self, device_index):
        return torch.cuda.default_generators[device_index]

    # ---- streams / events -------------------------------------------------------------------------------------
    @property
    def Stream(self):
        return torch.cuda.Stream

    def stream(self, stream):
        return torch.cuda.stream(stream)

    def current_stream(self, device_index=None):
        return torch.cuda.current_stream(device_index)

    def default_stream(self, device_index=None):
        return torch.cuda.default_stream(device_index)

    @property
    def Event(self):
        return torch.cuda.Event

    # ---- memory ---------------------------------------------------------------------------------------------
    def empty_cache(self):
        return torch.cuda.empty_cache()

    def memory_allocated(self, device_index=None):
        return torch.cuda.memory_allocated(device_index)

    def max_memory_allocated(self, device_index=None):
        return torch.cuda.max_memory_allocated(device_index)

    def reset_max_memory_allocated(self, device_index=None):
        return torch.cuda.reset_max_memory_allocated(device_index)

    def memory_cached(self, device_index=None):
        return torch.cuda.memory_reserved(device_index)

    def max_memory_cached(self, device_index=None):
        return torch.cuda.max_memory_reserved(device_index)

    def reset_max_memory_cached(self, device_index=None):
        return torch.cuda.reset_peak_memory_stats(device_index)

    def memory_stats(self, device_index=None):
        return torch.cuda.memory_stats(device_index)

    def reset_peak_memory_stats(self, device_index=None):
        return torch.cuda.reset_peak_memory_stats(device_index)

    def memory_reserved(self, device_index=None):
        return torch.cuda.memory_reserved(device_index)

    def max_memory_reserved(self, device_index=None):
        return torch.cuda.max_memory_reserved(device_index)

    def total_memory(self, device_index=None):
        return torch.cuda.get_device_properties(device_index or 0).total_memory

    def available_memory(self, device_index=None):
        free, _ = torch.cuda.mem_get_info(device_index)
        return free

    # ---- dtypes / capability -------------------------------------------------------------------------------
    def is_bf16_supported(self):
        return True

    def is_fp16_supported(self):
        return True

    def supported_dtypes(self):
        return [torch.float, torch.half, torch.bfloat16]

    def is_triton_supported(self):
        return False  # no Triton path in this framework by design

    def amp(self):
        return torch.amp

    def is_available(self):
        return torch.cuda.is_available()

    def range_push(self, msg):
        return torch.cuda.nvtx.range_push(msg)

    def range_pop(self):
        return torch.cuda.nvtx.range_pop()

    def lazy_call(self, callback):
        return torch.cuda._lazy_call(callback)

    def communication_backend_name(self):
        return self._communication_backend_name

    def create_graph(self):
        return torch.cuda.CUDAGraph()

    def capture_to_graph(self, graph, pool=None, stream=None):
        return torch.cuda.graph(graph, pool, stream)

    def replay_graph(self, graph):
        graph.replay()

    # ---- tensors ---------------------------------------------------------------------------------------------
    @property
    def BFloat16Tensor(self):
        return functools.partial(torch.tensor, dtype=torch.bfloat16, device="cuda")

    @property
    def HalfTensor(self):
        return functools.partial(torch.tensor, dtype=torch.half, device="cuda")

    @property
    def FloatTensor(self):
        return functools.partial(torch.tensor, dtype=torch.float, device="cuda")

    @property
    def IntTensor(self):
        return functools.partial(torch.tensor, dtype=torch.int, device="cuda")

    @property
    def LongTensor(self):
        return functools.partial(torch.tensor, dtype=torch.long, device="cuda")

    def pin_memory(self, tensor, align_bytes=1):
        return tensor.pin_memory()

    def is_pinned(self, tensor):
        return tensor.is_pinned()

    def on_accelerator(self, tensor):
        return str(tensor.device).startswith("cuda")

    def op_builder_dir(self):
        return "hcache_deepspeed_b200.ops.op_builder"

    def create_op_builder(self, class_name):
        from ..ops.op_builder import get_builder
        return get_builder(class_name)

    def get_op_builder(self, class_name):
        from ..ops.op_builder import get_builder_class
        return get_builder_class(class_name)

    def build_extension(self):
        from ..ops import build
        return build

    def export_envs(self):
        return ["NCCL"]

    def visible_devices_envs(self):
        return ["CUDA_VISIBLE_DEVICES"]

    def set_visible_devices_envs(self, current_env, local_accelerator_ids):
        for env in self.visible_devices_envs():
            current_env[env] = ",".join(map(str, local_accelerator_ids))

    def get_compile_backend(self):
        return self._compile_backend

    def set_compile_backend(self, backend):
        self._compile_backend = backend


class CPU_Accelerator(CUDA_Accelerator):
    """Host stand-in (tests, the CPU configuration of ``init_inference``): gloo, no streams."""

    def __init__(self):
        super().__init__()
        self._name = "cpu"
        self._communication_backend_name = "gloo"

    def is_synchronized_device(self):
        return True

    def use_host_timers(self):
        return True

    def device_name(self, device_index=None):
        return "cpu"

    def device(self, device_index=None):
        return torch.device("cpu")

    def set_device(self, device_index):
        pass

    def current_device(self):
        return 0

    def current_device_name(self):
        return "cpu"

    def device_count(self):
        return int(os.environ.get("LOCAL_SIZE", 1))

    def synchronize(self, device_index=None):
        pass

    def set_rng_state(self, new_state, device_index=None):
        return torch.set_rng_state(new_state)

    def get_rng_state(self, device_index=None):
        return torch.get_rng_state()

    def manual_seed(self, seed):
        return torch.manual_seed(seed)

    manual_seed_all = manual_seed

    def is_available(self):
        return True

    def memory_allocated(self, device_index=None):
        import psutil
        return psutil.Process().memory_info().rss

    max_memory_allocated = memory_allocated
    memory_reserved = memory_allocated

    def total_memory(self, device_index=None):
        import psutil
        return psutil.virtual_memory().total

    def available_memory(self, device_index=None):
        import psutil
        return psutil.virtual_memory().available

    def empty_cache(self):
        pass

    def on_accelerator(self, tensor):
        return str(tensor.device) == "cpu"

    def pin_memory(self, tensor, align_bytes=1):
        return tensor

    def is_pinned(self, tensor):
        return tensor.is_pinned()

    def range_push(self, msg):
        pass

    def range_pop(self):
        pass
