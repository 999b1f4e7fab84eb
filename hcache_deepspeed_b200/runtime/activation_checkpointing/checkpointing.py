"""Activation checkpointing with the Megatron-style extras: partition the saved activations across model-parallel
ranks, move them to host memory, keep an RNG-state tracker so dropout replays identically.

``checkpoint(function, *args)`` is a ``torch.autograd.Function``: forward runs without grad and stores the inputs
(whole, partitioned 1/mp each, or on CPU); backward restores RNG state, re-gathers the inputs, recomputes with grad and
backpropagates. Reference: ``deepspeed/runtime/activation_checkpointing/checkpointing.py`` (CheckpointFunction :488,
partition_activations :377, get_full_inputs :284, CudaRNGStatesTracker :122, configure :1071)."""
from __future__ import annotations

import contextlib
from typing import Optional

import torch

from ... import comm as dist

# module-level configuration (set by configure())
PARTITION_ACTIVATIONS = False
CPU_CHECKPOINT = False
CONTIGUOUS_CHECKPOINTING = False
SYNCHRONIZE = False
PROFILE_TIME = False
num_layers = None
mpu = None
_configured = False
_MODEL_PARALLEL_RNG_TRACKER_NAME = "model-parallel-rng"


class CudaRNGStatesTracker:
    """Named RNG states; ``fork(name)`` runs a block under that state and saves it back (so model-parallel regions use a
    different stream of randomness from data-parallel ones). Works for the CPU generator when CUDA is absent."""

    def __init__(self):
        self.states_ = {}
        self.seeds_ = set()

    def reset(self):
        self.states_, self.seeds_ = {}, set()

    def get_states(self):
        return dict(self.states_)

    def set_states(self, states):
        self.states_ = states

    @staticmethod
    def _get():
        return torch.cuda.get_rng_state() if torch.cuda.is_available() else torch.get_rng_state()

    @staticmethod
    def _set(s):
        torch.cuda.set_rng_state(s) if torch.cuda.is_available() else torch.set_rng_state(s)

    def add(self, name, seed):
        if seed in self.seeds_:
            raise Exception(f"seed {seed} already exists")
        if name in self.states_:
            raise Exception(f"cuda rng state {name} already exists")
        self.seeds_.add(seed)
        orig = self._get()
        torch.cuda.manual_seed(seed) if torch.cuda.is_available() else torch.manual_seed(seed)
        self.states_[name] = self._get()
        self._set(orig)

    @contextlib.contextmanager
    def fork(self, name=_MODEL_PARALLEL_RNG_TRACKER_NAME):
        if name not in self.states_:
            raise Exception(f"cuda rng state {name} is not added")
        orig = self._get()
        self._set(self.states_[name])
        try:
            yield
        finally:
            self.states_[name] = self._get()
            self._set(orig)


_CUDA_RNG_STATE_TRACKER = CudaRNGStatesTracker()


def get_cuda_rng_tracker():
    return _CUDA_RNG_STATE_TRACKER


def model_parallel_cuda_manual_seed(seed: int):
    """Default state: same seed on all model-parallel ranks; tracked state: ``seed + 2718 + mp_rank``."""
    mp_rank = mpu.get_model_parallel_rank() if mpu is not None else 0
    _CUDA_RNG_STATE_TRACKER.reset()
    torch.cuda.manual_seed(seed) if torch.cuda.is_available() else torch.manual_seed(seed)
    _CUDA_RNG_STATE_TRACKER.add(_MODEL_PARALLEL_RNG_TRACKER_NAME, seed + 2718 + mp_rank)


model_parallel_reconfigure_tp_seed = model_parallel_cuda_manual_seed


def _mp_info():
    if mpu is None or not dist.is_initialized():
        return 1, 0, None
    return mpu.get_model_parallel_world_size(), mpu.get_model_parallel_rank(), mpu.get_model_parallel_group()


def partition_activation(t: torch.Tensor):
    """This rank's 1/mp slice (flattened, padded) of an activation."""
    ws, rank, _ = _mp_info()
    if ws == 1:
        return t.detach()
    flat = t.detach().contiguous().view(-1)
    psize = (flat.numel() + ws - 1) // ws
    out = torch.zeros(psize, dtype=flat.dtype, device=flat.device)
    lo, hi = rank * psize, min(flat.numel(), (rank + 1) * psize)
    if lo < hi:
        out[:hi - lo].copy_(flat[lo:hi])
    return out


def gather_partitioned_activation(part: torch.Tensor, shape, device) -> torch.Tensor:
    ws, _, group = _mp_info()
    part = part.to(device)
    if ws == 1:
        return part.view(shape)
    full = torch.empty(part.numel() * ws, dtype=part.dtype, device=device)
    dist.all_gather_into_tensor(full, part.contiguous(), group=group)
    n = 1
    for s in shape:
        n *= s
    return full[:n].view(shape)


class CheckpointFunction(torch.autograd.Function):

    @staticmethod
    def forward(ctx, run_function, all_outputs, *args):
        ctx.run_function = run_function
        ctx.cpu_rng = torch.get_rng_state()
        ctx.cuda_rng = torch.cuda.get_rng_state() if torch.cuda.is_available() else None
        ctx.tracker_states = get_cuda_rng_tracker().get_states()
        ctx.partitioned = PARTITION_ACTIVATIONS and _mp_info()[0] > 1
        ctx.cpu = CPU_CHECKPOINT
        ctx.meta = []
        saved = []
        for a in args:
            if torch.is_tensor(a):
                ctx.meta.append(("t", tuple(a.shape), a.device, a.requires_grad, a.is_floating_point()))
                s = partition_activation(a) if (ctx.partitioned and a.is_floating_point()) else a.detach()
                saved.append(s.cpu() if ctx.cpu else s)
            else:
                ctx.meta.append(("o", a))
        ctx.save_for_backward(*saved)
        with torch.no_grad():
            outputs = run_function(*args)
        if SYNCHRONIZE and torch.cuda.is_available():
            torch.cuda.synchronize()
        outs = outputs if isinstance(outputs, tuple) else (outputs, )
        all_outputs += list(outs)
        non_grad = [o for o in outs if torch.is_tensor(o) and not o.is_floating_point()]
        ctx.mark_non_differentiable(*non_grad)
        return outputs if isinstance(outputs, tuple) else outputs

    @staticmethod
    def backward(ctx, *grads):
        saved = list(ctx.saved_tensors)
        inputs, ti = [], 0
        for m in ctx.meta:
            if m[0] == "o":
                inputs.append(m[1])
                continue
            _, shape, device, rg, is_fp = m
            t = saved[ti]
            ti += 1
            if ctx.partitioned and is_fp:
                t = gather_partitioned_activation(t, shape, device)
            else:
                t = t.to(device)
            t = t.detach()
            t.requires_grad = rg
            inputs.append(t)
        cpu_rng, cuda_rng = torch.get_rng_state(), (torch.cuda.get_rng_state() if torch.cuda.is_available() else None)
        tracker = get_cuda_rng_tracker().get_states()
        torch.set_rng_state(ctx.cpu_rng)
        if ctx.cuda_rng is not None:
            torch.cuda.set_rng_state(ctx.cuda_rng)
        get_cuda_rng_tracker().set_states(ctx.tracker_states)
        with torch.enable_grad():
            outputs = ctx.run_function(*inputs)
        torch.set_rng_state(cpu_rng)
        if cuda_rng is not None:
            torch.cuda.set_rng_state(cuda_rng)
        get_cuda_rng_tracker().set_states(tracker)
        outs = outputs if isinstance(outputs, tuple) else (outputs, )
        pairs = [(o, g) for o, g in zip(outs, grads) if torch.is_tensor(o) and o.requires_grad and g is not None]
        if pairs:
            torch.autograd.backward([o for o, _ in pairs], [g for _, g in pairs])
        in_grads = tuple(i.grad if torch.is_tensor(i) else None for i in inputs)
        return (None, None) + in_grads


def checkpoint(function, *args):
    """Checkpoint ``function(*args)``: activations inside are recomputed during backward."""
    all_outputs = []
    CheckpointFunction.apply(function, all_outputs, *args)
    return all_outputs[0] if len(all_outputs) == 1 else tuple(all_outputs)


def non_reentrant_checkpoint(function, *args):
    from torch.utils.checkpoint import checkpoint as tc
    return tc(function, *args, use_reentrant=False)


def partition_activations_in_checkpoint(partition_activation_: bool):
    global PARTITION_ACTIVATIONS
    PARTITION_ACTIVATIONS = partition_activation_


def set_num_layers(nlayers):
    global num_layers
    num_layers = nlayers


def reset():
    """Free buffers kept between iterations (contiguous checkpointing); nothing is cached in this implementation."""


def _configure_defaults():
    global mpu, num_layers, PARTITION_ACTIVATIONS, CONTIGUOUS_CHECKPOINTING, CPU_CHECKPOINT, SYNCHRONIZE, PROFILE_TIME, _configured
    PARTITION_ACTIVATIONS = CONTIGUOUS_CHECKPOINTING = CPU_CHECKPOINT = SYNCHRONIZE = PROFILE_TIME = False
    num_layers = None
    _configured = True


def configure(mpu_, deepspeed_config=None, partition_activations: Optional[bool] = None,
              contiguous_checkpointing: Optional[bool] = None, num_checkpoints: Optional[int] = None,
              checkpoint_in_cpu: Optional[bool] = None, synchronize: Optional[bool] = None, profile: Optional[bool] = None):
    """Configure from a DeepSpeed config (dict / path / DeepSpeedConfig) and/or explicit overrides."""
    global mpu, num_layers, PARTITION_ACTIVATIONS, CONTIGUOUS_CHECKPOINTING, CPU_CHECKPOINT, SYNCHRONIZE, PROFILE_TIME, _configured
    _configure_defaults()
    mpu = mpu_
    if deepspeed_config is not None:
        from ..config import DeepSpeedConfig
        cfg = deepspeed_config if isinstance(deepspeed_config, DeepSpeedConfig) else DeepSpeedConfig(deepspeed_config, mpu=mpu_)
        ac = cfg.activation_checkpointing_config
        PARTITION_ACTIVATIONS, CONTIGUOUS_CHECKPOINTING = ac.partition_activations, ac.contiguous_memory_optimization
        num_layers, CPU_CHECKPOINT = ac.number_checkpoints, ac.cpu_checkpointing
        SYNCHRONIZE, PROFILE_TIME = ac.synchronize_checkpoint_boundary, ac.profile
    for name, val in (("PARTITION_ACTIVATIONS", partition_activations), ("CONTIGUOUS_CHECKPOINTING", contiguous_checkpointing),
                      ("num_layers", num_checkpoints), ("CPU_CHECKPOINT", checkpoint_in_cpu), ("SYNCHRONIZE", synchronize),
                      ("PROFILE_TIME", profile)):
        if val is not None:
            globals()[name] = val
    if CONTIGUOUS_CHECKPOINTING:
        assert PARTITION_ACTIVATIONS, "Contiguous Checkpointing is only available with partitioned activations"
        assert num_layers is not None, "Must specify the number of layers with contiguous memory checkpointing"
    _configured = True


def is_configured() -> bool:
    return _configured
