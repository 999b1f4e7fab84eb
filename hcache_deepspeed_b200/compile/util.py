from __future__ import annotations


def is_deepcompile_supported() -> bool:
    """The schedule compiler has no dependency beyond torch (the reference needs torch>=2.6 + a CUDA accelerator)."""
    return True


def get_deepcompile_handle():
    from . import backend
    return backend
