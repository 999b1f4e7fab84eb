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


# ---- the reference's helper names around CheckpointFunction (checkpointing.py:68-480), over the functions above -----------
def detach_variable(inputs, device=None):
    """Detached copies of the tensors of a tuple (moved to ``device`` when given), keeping ``requires_grad``; non-tensors
    pass through."""
    if not isinstance(inputs, tuple):
        raise RuntimeError(f"Only tuple of tensors is supported. Got Unsupported input type: {type(inputs).__name__}")
    out = []
    for inp in inputs:
        if not torch.is_tensor(inp):
            out.append(inp)
            continue
        req = inp.requires_grad
        x = inp.to(device=device) if device is not None else inp
        x = x.detach()
        x.requires_grad = req
        out.append(x)
    return tuple(out)


def get_partition_size(item: torch.Tensor) -> int:
    """Elements of ``item`` one model-parallel rank keeps (the padded 1/mp share used by ``partition_activation``)."""
    ws, _, _ = _mp_info()
    return (item.numel() + ws - 1) // ws


def get_partition_start(item: torch.Tensor) -> int:
    _, rank, _ = _mp_info()
    return rank * get_partition_size(item)


def is_activation_to_checkpoint(item) -> bool:
    """Floating-point tensors large enough to split over the model-parallel group, unless flagged ``no_checkpointing``."""
    ws, _, _ = _mp_info()
    return torch.is_tensor(item) and item.is_floating_point() and item.numel() >= ws and \
        not getattr(item, "no_checkpointing", False)


def extract_tensors(all_objects):
    """Split a list / tuple into (tensors, non-tensors, flags) - ``flags[i]`` says whether position ``i`` was a tensor."""
    tensors = [v for v in all_objects if torch.is_tensor(v)]
    others = [v for v in all_objects if not torch.is_tensor(v)]
    flags = [torch.is_tensor(v) for v in all_objects]
    return (tuple(tensors), tuple(others), tuple(flags)) if isinstance(all_objects, tuple) else (tensors, others, flags)


def merge_tensors(tensor_objects, non_tensor_objects, tensor_flags):
    """Inverse of ``extract_tensors``. ``tensor_objects`` may also be the (partition, size) pair list a partitioned
    checkpoint saves: then every tensor position takes two entries and the second is dropped."""
    t_it, o_it = iter(tensor_objects), iter(non_tensor_objects)
    paired = len(tensor_objects) == 2 * sum(bool(f) for f in tensor_flags) and len(tensor_objects) > 0
    out = []
    for is_tensor in tensor_flags:
        if is_tensor:
            out.append(next(t_it))
            if paired:
                next(t_it)
        else:
            out.append(next(o_it))
    return tuple(out)


def partition_activations(args, cpu_checkpoint: bool = False, contiguous_checkpoint: bool = False):
    """What the forward of a partitioned checkpoint saves for ``args``: this rank's slice of each checkpointable tensor
    (on the host under ``cpu_checkpoint``), other items as they are. ``contiguous_checkpoint`` is accepted for signature
    parity - slices here are allocated by the caching allocator, which already places equal-sized blocks together."""
    out = []
    for item in args:
        if not is_activation_to_checkpoint(item):
            out.append(item)
            continue
        part = partition_activation(item)
        out.append(part.cpu() if cpu_checkpoint else part)
    return out


def get_partitioned_activations_for_backward(args, inputs, contiguous_checkpoint: bool = False):
    """The flat list ``[partition0, size0, partition1, size1, ...]`` kept for backward: the saved slice of every
    checkpointable argument followed by its original shape as a tensor; other arguments followed by ``None``."""
    out = []
    for arg, inp in zip(args, inputs):
        if not is_activation_to_checkpoint(arg):
            out += [arg, None]
            continue
        out += [inp, torch.tensor(list(arg.shape), dtype=torch.long)]
    return out


def gather_partitioned_activations(tensors, device=None):
    """Inverse of the above: every (slice, shape) pair is all-gathered over the model-parallel group and reshaped."""
    assert len(tensors) % 2 == 0, f"Expected even count of tensors, instead got {len(tensors)}"
    out = []
    for item, size in zip(tensors[0::2], tensors[1::2]):
        if size is None or not torch.is_tensor(item):
            out.append(item)
            continue
        dev = device if device is not None else (item.device if item.device.type != "cpu" or not torch.cuda.is_available()
                                                 else torch.device("cuda", torch.cuda.current_device()))
        full = gather_partitioned_activation(item, [int(s) for s in size.tolist()], dev)
        out.append(full)
    return tuple(out)


def get_cpu_activations_for_backward(args, inputs):
    """Under CPU checkpointing without partitioning: the host copies replace the checkpointable arguments."""
    return [inp if is_activation_to_checkpoint(arg) else arg for arg, inp in zip(args, inputs)]
