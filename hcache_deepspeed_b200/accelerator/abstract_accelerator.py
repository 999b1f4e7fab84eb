"""Accelerator interface (``deepspeed/accelerator/abstract_accelerator.py``). This framework targets ONE device family
(NVIDIA Blackwell through CUDA) plus a CPU stand-in for tests; there is no multi-vendor dispatch."""
import abc


class DeepSpeedAccelerator(abc.ABC):
    def __init__(self):
        self._name = None
        self._communication_backend_name = None
        self._compile_backend = None

    @abc.abstractmethod
    def device_name(self, device_index=None):
        ...

    @abc.abstractmethod
    def is_available(self):
        ...
