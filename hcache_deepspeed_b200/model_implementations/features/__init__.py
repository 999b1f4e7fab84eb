from .cuda_graph import CUDAGraph, GraphedCallable  # noqa: F401
