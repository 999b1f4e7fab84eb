"""Optimizer-state offload schedule (reference ``deepspeed/compile/passes/offload_adam_states.py`` ``move_opt_states``):
the Adam moments are only touched inside ``optimizer.step()``, so they can live in pinned host memory for the whole
forward/backward and come back just in time. With ``double_buffer`` the reload is issued on a side stream when the
backward STARTS, so the copy hides under the backward pass; the step waits on that stream."""
from __future__ import annotations

import torch

from ...runtime.zero.offload_states import offload_adam_states, reload_adam_states

NAME = "offload_adam_states"


class AdamStateMover:

    def __init__(self, optimizer, device, double_buffer: bool = True, pin_memory: bool = True) -> None:
        self.base = getattr(optimizer, "optimizer", optimizer)
        self.device = torch.device(device)
        self.double_buffer = double_buffer
        self.pin_memory = pin_memory
        self.stream = torch.cuda.Stream(self.device) if self.device.type == "cuda" else None
        self.offloaded = False
        self.bytes_moved = 0

    def _nbytes(self) -> int:
        return sum(t.numel() * t.element_size() for st in self.base.state.values() for k, t in st.items()
                   if k in ("exp_avg", "exp_avg_sq") and torch.is_tensor(t))

    def offload(self) -> None:
        if self.offloaded or not self.base.state:
            return
        offload_adam_states(self.base, "cpu", pin_memory=self.pin_memory, non_blocking=self.stream is not None)
        self.offloaded = True
        self.bytes_moved += self._nbytes()

    def start_reload(self) -> None:
        if not self.offloaded:
            return
        if self.stream is not None:
            self.stream.wait_stream(torch.cuda.current_stream(self.device))
            with torch.cuda.stream(self.stream):
                reload_adam_states(self.base, self.device, non_blocking=True)
        else:
            reload_adam_states(self.base, self.device)
        self.offloaded = False
        self.bytes_moved += self._nbytes()

    def wait(self) -> None:
        if self.stream is not None:
            torch.cuda.current_stream(self.device).wait_stream(self.stream)


def move_opt_states(optimizer, device, double_buffer: bool = True) -> AdamStateMover:
    """Wrap ``optimizer.step`` so the moments are on ``device`` only while it runs. Returns the mover (call
    ``mover.start_reload()`` at the start of the backward pass to overlap the copy-in)."""
    mover = AdamStateMover(optimizer, device, double_buffer)
    inner = optimizer.step

    def step(*a, **k):
        mover.start_reload()
        mover.wait()
        out = inner(*a, **k)
        mover.offload()
        return out

    optimizer.step = step
    optimizer._adam_state_mover = mover
    return mover
