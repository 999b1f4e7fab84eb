"""``torch.compile`` guards (``deepspeed/runtime/compiler.py:19-33``). The framework's hot path is hand-written CUDA plus
CUDA graphs, never a tracing compiler; these helpers only keep hook code OUT of a user's own compiled regions."""
from __future__ import annotations

import torch


def is_compile_supported() -> bool:
    return hasattr(torch, "compiler") and hasattr(torch.nn.Module, "compile")


def disable(func):
    """Exclude ``func`` (a module hook, a collective wrapper) from tracing when the user compiles their model."""
    if is_compile_supported():
        return torch.compiler.disable(func)
    return func


def is_compiling() -> bool:
    return is_compile_supported() and torch.compiler.is_compiling()
