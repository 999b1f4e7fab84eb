"""One place where the framework records CUDA graphs.

A stream capture is invalidated by ANY capture-illegal CUDA call made while it is open — including calls the
capturing code never wrote. The one that bit this code base: Python's cyclic garbage collector can run at any
allocation, and if it finalises an old ``torch.cuda.CUDAGraph`` (engines and graphed modules sit in reference
cycles) its ``cudaGraphExecDestroy`` lands inside the open capture ("operation not permitted when stream is
capturing (function reset)") and ``capture_end`` then fails with ``cudaErrorStreamCaptureInvalidated``. Whether the
collector fires inside the window depends on allocation counts, i.e. on which tests ran before — a flaky failure.

:func:`capture` therefore collects garbage *before* opening the capture, keeps the collector off while it is
open, and confines the capture-mode checks to this thread (``thread_local``) so helper threads (pinned-memory
allocator, data loaders) cannot invalidate it either.
"""
from __future__ import annotations

import gc
from typing import Any, Callable, Optional, Tuple

import torch


def capture(fn: Callable[[], Any], device: Optional[torch.device] = None,
            stream: Optional["torch.cuda.Stream"] = None, pool=None) -> Tuple["torch.cuda.CUDAGraph", Any]:
    """Record ``fn()`` into a new graph on ``stream`` (a fresh side stream by default). Returns ``(graph, fn's result)``.
    The caller must have run ``fn`` eagerly before (lazy kernel-attribute setup and allocator warm-up are not capturable).
    On failure the partially built graph is dropped, the device is synchronised and the exception re-raised."""
    if device is None:
        device = torch.device("cuda", torch.cuda.current_device())
    torch.cuda.synchronize(device)
    gc.collect()
    was_enabled = gc.isenabled()
    gc.disable()
    side = stream if stream is not None else torch.cuda.Stream(device=device)
    side.wait_stream(torch.cuda.current_stream(device))
    graph = torch.cuda.CUDAGraph()
    try:
        with torch.cuda.stream(side):
            kw = {"pool": pool} if pool is not None else {}
            with torch.cuda.graph(graph, stream=side, capture_error_mode="thread_local", **kw):
                out = fn()
    except Exception:
        del graph
        torch.cuda.synchronize(device)
        raise
    finally:
        if was_enabled:
            gc.enable()
    torch.cuda.current_stream(device).wait_stream(side)
    return graph, out
