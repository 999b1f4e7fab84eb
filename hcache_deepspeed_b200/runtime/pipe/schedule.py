"""Pipeline schedules as generators of instruction lists, one list per step, executed by the PipelineEngine.
``TrainSchedule`` is 1F1B (one-forward-one-backward: at most ``stages - stage_id`` activations alive), generated from
the warm-up / steady / drain definition of the schedule; ``InferenceSchedule`` is a forward-only wavefront with two
alternating buffers. Exchanges between neighbouring stages are ordered by stage parity so blocking sends pair up.
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


def _by_parity(stage_id: int, sends, recvs):
    """Transfers that fall on the same tick boundary: even stages send first, odd stages receive first. Neighbours have
    opposite parity, so every blocking send meets its receive and the chain of pairings unwinds from stage 0."""
    return list(sends) + list(recvs) if stage_id % 2 == 0 else list(recvs) + list(sends)


class InferenceSchedule(PipeSchedule):
    """Forward-only wavefront. Stage ``s`` runs micro-batch ``m`` at tick ``s + m``; two buffers alternate by micro-batch
    parity. One yielded step = one tick: hand the previous output downstream and take the next input from upstream
    (parity-ordered), then compute."""

    def steps(self):
        s, S, M = self.stage_id, self.stages, self.micro_batches
        for tick in range(M + S - 1):
            m = tick - s                        # micro-batch this stage computes at this tick (if valid)
            cmds = []
            sends = [SendActivation((m - 1) % 2)] if s + 1 < S and 0 <= m - 1 < M else []
            recvs = [RecvActivation(m % 2)] if s > 0 and 0 <= m < M else []
            if 0 <= m < M and (s == 0 or s == S - 1):
                cmds.append(LoadMicroBatch(m % 2))
            cmds += _by_parity(s, sends, recvs)
            if 0 <= m < M:
                cmds.append(ForwardPass(m % 2))
            yield cmds

    def num_pipe_buffers(self):
        return 2


def _one_f_one_b_order(stages: int, stage_id: int, micro_batches: int):
    """Compute order of one stage under 1F1B, from the definition: ``w = min(stages - stage_id - 1, M)`` warm-up
    forwards, then one forward + one backward per remaining micro-batch, then the ``w`` outstanding backwards."""
    warm = min(stages - stage_id - 1, micro_batches)
    ops = [("F", m) for m in range(warm)]
    for i in range(micro_batches - warm):
        ops.append(("F", warm + i))
        ops.append(("B", i))
    ops += [("B", m) for m in range(micro_batches - warm, micro_batches)]
    return ops


def _timeline(stages: int, micro_batches: int):
    """Tick at which every compute op of every stage runs when all ops take one tick and a tensor is available the tick
    after its producer ran: ``{(stage, kind, micro_batch): tick}``. Every rank evaluates the same pure function, so all
    stages agree on the global clock without exchanging anything."""
    orders = [_one_f_one_b_order(stages, s, micro_batches) for s in range(stages)]
    when, nxt, free_at = {}, [0] * stages, [0] * stages
    remaining = sum(len(o) for o in orders)
    while remaining:
        progressed = False
        for s in range(stages):
            if nxt[s] == len(orders[s]):
                continue
            kind, m = orders[s][nxt[s]]
            dep = None
            if kind == "F" and s > 0:
                dep = (s - 1, "F", m)
            elif kind == "B" and s + 1 < stages:
                dep = (s + 1, "B", m)
            if dep is not None and dep not in when:
                continue
            t = max(free_at[s], when[dep] + 1 if dep is not None else 0)
            when[(s, kind, m)] = t
            free_at[s] = t + 1
            nxt[s] += 1
            remaining -= 1
            progressed = True
        assert progressed, "1F1B orders of the stages are inconsistent"
    return when


class TrainSchedule(PipeSchedule):
    """One-forward-one-backward (1F1B). Each stage's compute order comes from the definition (warm-up forwards, steady
    forward/backward pairs, drain: :func:`_one_f_one_b_order`); a unit-time simulation of all stages (:func:`_timeline`)
    places every op on a global tick. A tensor crosses a stage boundary on the tick boundary right before its consumer
    runs — both neighbours derive that tick from the same timeline, so the blocking send and its receive are issued in
    the same step — and the transfers of one tick boundary are ordered by stage parity. At most ``stages - stage_id``
    activations are alive on a stage, which is its buffer count. One yielded step = one tick of this stage (possibly
    empty while the pipeline fills or drains); the last step also reduces gradients and steps the optimizer."""

    def steps(self):
        s, S, M = self.stage_id, self.stages, self.micro_batches
        when = _timeline(S, M)
        last_tick = max(when.values())
        mine = {t: (kind, m) for (st, kind, m), t in when.items() if st == s}
        # transfers keyed by the tick of their CONSUMER: (tick) -> instructions of this stage
        sends, recvs = {}, {}
        for (st, kind, m), t in when.items():
            buf = self._buffer_idx(m)
            if kind == "F" and st == s + 1:            # my forward output feeds the next stage's forward at tick t
                sends.setdefault(t, []).append(SendActivation(buf))
            elif kind == "B" and st == s - 1:          # my input gradient feeds the previous stage's backward at tick t
                sends.setdefault(t, []).append(SendGrad(buf))
            elif st == s and kind == "F" and s > 0:
                recvs.setdefault(t, []).append(RecvActivation(buf))
            elif st == s and kind == "B" and s + 1 < S:
                recvs.setdefault(t, []).append(RecvGrad(buf))
        for t in range(last_tick + 1):
            cmds = _by_parity(s, sends.get(t, []), recvs.get(t, []))
            if t in mine:
                kind, m = mine[t]
                buf = self._buffer_idx(m)
                if kind == "F":
                    if s == 0 or s == S - 1:
                        cmds.append(LoadMicroBatch(buf))
                    cmds.append(ForwardPass(buf))
                else:
                    cmds.append(BackwardPass(buf))
            if t == last_tick:
                cmds += [ReduceTiedGrads(), ReduceGrads(), OptimizerStep()]
            yield cmds

    def num_pipe_buffers(self):
        return max(2, min(self.stages - self.stage_id, self.micro_batches))


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
