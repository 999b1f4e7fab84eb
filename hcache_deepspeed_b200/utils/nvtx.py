"""NVTX range decorator (ref: deepspeed/utils/nvtx.py:6-25)."""
import torch

enable_nvtx = True


def instrument_w_nvtx(func):
    """Wrap ``func`` in an NVTX range named after it (no-op without CUDA or when ``enable_nvtx`` is False)."""

    def wrapped(*args, **kwargs):
        if enable_nvtx and torch.cuda.is_available():
            torch.cuda.nvtx.range_push(func.__qualname__)
            try:
                return func(*args, **kwargs)
            finally:
                torch.cuda.nvtx.range_pop()
        return func(*args, **kwargs)

    wrapped.__name__ = getattr(func, "__name__", "wrapped")
    return wrapped
