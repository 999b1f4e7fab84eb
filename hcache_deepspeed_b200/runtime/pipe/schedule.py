"""Pipeline schedules as generators of instruction lists, one list per step, executed by the PipelineEngine.
``TrainSchedule`` is 1F1B (one-forward-one-backward: at most ``stages - stage_id`` activations alive),
``InferenceSchedule`` is a forward-only wavefront with two alternating buffers.
Reference: ``deepspeed/runtime/pipe/schedule.py`` (PipeSchedule :10, InferenceSchedule :135, TrainSchedule :189,
DataParallelSchedule :301, instruction classes :327-494)."""
from __future__ import annotations

from abc import ABC, abstractmethod

from ..utils import call_to_str


class PipeSchedule(ABC):
    def __init__(self, micro_batches: int, stages: int, stage_id: int):
        self.micro_batches, self.stages, self.stage_id = micro_batches, stages, stage_id
        self.prev_stage, self.next_stage = stage_id - 1, stage_id + 1

    @abstractmethod
    def steps(self):
        """Yield a list of :class:`PipeInstruction` per schedule step."""

    def num_pipe_buffers(self) -> int:
        return self.micro_batches

    def _valid_micro_batch(self, mb: int) -> bool:
        return 0 <= mb < self.micro_batches

    def _valid_stage(self, s: int) -> bool:
        return 0 <= s < self.stages

    @property
    def stage(self):
        return self.stage_id

    @property
    def num_stages(self):
        return self.stages

    @property
    def num_micro_batches(self):
        return self.micro_batches

    @property
    def is_first_stage(self):
        return self.stage_id == 0

    @property
    def is_last_stage(self):
        return self.stage_id == self.stages - 1

    def _buffer_idx(self, micro_batch_id: int) -> int:
        assert self._valid_micro_batch(micro_batch_id)
        return micro_batch_id % self.num_pipe_buffers()

    def __iter__(self):
        self.it = None
        return self

    def __next__(self):
        if self.it is None:
            self.it = self.steps()
        return next(self.it)


class InferenceSchedule(PipeSchedule):
    def steps(self):
        total = self.micro_batches + self.stages - 1
        for step in range(total):
            cmds = []
            mb = step - self.stage_id
            # even stages compute into buffer step%2 and receive into the other; odd stages the reverse, so a send
            # always pairs with a recv on the neighbour
            if self.stage_id % 2 == 0:
                recv_buf, send_buf = step % 2, (step + 1) % 2
            else:
                recv_buf, send_buf = (step + 1) % 2, step % 2
            if self.is_first_stage or self.is_last_stage:
                if self._valid_micro_batch(mb):
                    cmds.append(LoadMicroBatch(recv_buf))
            if self.stage_id % 2 == 0:
                if self._valid_stage(self.next_stage) and self._valid_micro_batch(mb - 1):
                    cmds.append(SendActivation(send_buf))
                if self._valid_stage(self.prev_stage) and self._valid_micro_batch(mb):
                    cmds.append(RecvActivation(recv_buf))
            else:
                if self._valid_stage(self.prev_stage) and self._valid_micro_batch(mb):
                    cmds.append(RecvActivation(recv_buf))
                if self._valid_stage(self.next_stage) and self._valid_micro_batch(mb - 1):
                    cmds.append(SendActivation(send_buf))
            if self._valid_micro_batch(mb):
                cmds.append(ForwardPass(recv_buf))
            yield cmds

    def num_pipe_buffers(self):
        return 2


class TrainSchedule(PipeSchedule):
    """1F1B. Even steps on even stages (odd on odd) are forward slots, the others backward slots; a stage's first
    backward happens ``stages - stage_id`` slots after its first forward."""

    def steps(self):
        prev_mb = -1
        total = 2 * (self.micro_batches + self.stages - 1)
        for step in range(total):
            mb, is_forward = self._step_to_micro_batch(step)
            prev_buf = self._buffer_idx(prev_mb) if self._valid_micro_batch(prev_mb) else None
            cur_buf = self._buffer_idx(mb) if self._valid_micro_batch(mb) else None
            cmds = []
            if is_forward:
                # exchange: gradients of the previous (backward) slot go upstream, this slot's activation comes in
                if self._valid_micro_batch(prev_mb) and self._valid_stage(self.prev_stage):
                    cmds.append(SendGrad(prev_buf))
                if self._valid_micro_batch(mb) and self._valid_stage(self.prev_stage):
                    cmds.append(RecvActivation(cur_buf))
            else:
                if self._valid_micro_batch(mb) and self._valid_stage(self.next_stage):
                    cmds.append(RecvGrad(cur_buf))
                if self._valid_micro_batch(prev_mb) and self._valid_stage(self.next_stage):
                    cmds.append(SendActivation(prev_buf))
            if (self.is_first_stage or self.is_last_stage) and is_forward and self._valid_micro_batch(mb):
                cmds.append(LoadMicroBatch(cur_buf))
            if self._valid_micro_batch(mb):
                cmds.append(ForwardPass(cur_buf) if is_forward else BackwardPass(cur_buf))
            if step == total - 1:
                cmds.append(ReduceTiedGrads())
                cmds.append(ReduceGrads())
                cmds.append(OptimizerStep())
            prev_mb = mb
            yield cmds

    def num_pipe_buffers(self):
        return max(2, min(self.stages - self.stage_id, self.micro_batches))

    def _step_to_micro_batch(self, step: int):
        even_step, even_stage = step % 2 == 0, self.stage_id % 2 == 0
        if even_step == even_stage:  # forward slot
            base = step // 2 if even_step else (step - 1) // 2
            return base - self.stage_id // 2, True
        base = step // 2 if even_step else (step - 1) // 2  # backward slot
        if even_step:
            return base - self.stages + (self.stage_id + 1) // 2, False
        return base - self.stages + 1 + self.stage_id // 2, False


class DataParallelSchedule(PipeSchedule):
    """Plain data parallelism expressed as a schedule (single stage)."""

    def steps(self):
        for mb in range(self.micro_batches):
            cmds = [LoadMicroBatch(0), ForwardPass(0), BackwardPass(0)]
            if mb == self.micro_batches - 1:
                cmds += [ReduceGrads(), OptimizerStep()]
            yield cmds

    def num_pipe_buffers(self):
        return 1


class PipeInstruction:
    def __init__(self, **kwargs):
        self.name = self.__class__.__name__
        self.kwargs = kwargs
        for k, v in kwargs.items():
            setattr(self, k, v)

    def __repr__(self):
        return call_to_str(self.name, **self.kwargs)

    def __eq__(self, other):
        return type(self) is type(other) and self.kwargs == other.kwargs

    def __hash__(self):
        return hash((self.name, tuple(sorted(self.kwargs.items()))))


class OptimizerStep(PipeInstruction):
    pass


class ReduceGrads(PipeInstruction):
    pass


class ReduceTiedGrads(PipeInstruction):
    pass


class BufferOpInstruction(PipeInstruction):
    def __init__(self, buffer_id, **kwargs):
        super().__init__(buffer_id=buffer_id, **kwargs)


class LoadMicroBatch(BufferOpInstruction):
    pass


class ForwardPass(BufferOpInstruction):
    pass


class BackwardPass(BufferOpInstruction):
    pass


class SendActivation(BufferOpInstruction):
    pass


class RecvActivation(BufferOpInstruction):
    pass


class SendGrad(BufferOpInstruction):
    pass


class RecvGrad(BufferOpInstruction):
    pass


def _is_even(x):
    return x % 2 == 0


def _is_odd(x):
    return x % 2 != 0
