"""ZeRO-3 parameter partitioning: ``zero.Init`` converts every parameter of a module into a *partitioned parameter*
whose full tensor exists only while it is needed; ``GatheredParameters`` temporarily reassembles parameters.

Each converted ``torch.nn.Parameter`` carries:
  ``ds_id``, ``ds_shape``, ``ds_numel``, ``ds_tensor`` (this rank's padded 1-D slice, optionally in pinned host
  memory with ``remote_device='cpu'``), ``ds_status`` (:class:`ZeroParamStatus`), ``ds_persist`` (small parameters
  stay gathered), ``ds_process_group`` and the methods ``all_gather()``, ``partition()``, ``partition_numel()``.
While partitioned, ``param.data`` is an empty tensor. Gathering is one ``all_gather_into_tensor`` per parameter
(or per coalesced list via :func:`all_gather_params`).
Reference: ``deepspeed/runtime/zero/partition_parameters.py`` (Init :816, _convert_to_deepspeed_param :1107,
GatheredParameters :2165, ZeroParamStatus :215).
"""
from __future__ import annotations

import itertools
from enum import Enum
from typing import Iterable, List, Optional

import torch

from ... import comm as dist

_param_id = itertools.count()


class ZeroParamStatus(Enum):
    AVAILABLE = 1      # full tensor present
    NOT_AVAILABLE = 2  # only the local slice exists
    INFLIGHT = 3       # gather issued


class ZeroParamType(Enum):
    NORMAL = 1
    PARTITIONED = 2
    REMOTE = 3


def is_zero_param(p) -> bool:
    return hasattr(p, "ds_id")


# ---- NVMe tier for parameter partitions (ZeRO-Infinity; ref: partition_parameters.py + swap_tensor/partitioned_param_swapper.py) ----
# With ``offload_param.device = "nvme"`` a partition big enough for direct I/O lives in a file between uses:
# ``read_partition`` brings it into a pinned staging buffer, ``release_partition`` gives the buffer back (the file stays
# valid), ``write_partition`` stores new values and writes them out again. Small partitions stay in host memory.
_param_swapper = None


def set_param_swapper(swapper) -> None:
    global _param_swapper
    _param_swapper = swapper


def get_param_swapper():
    return _param_swapper


def _on_nvme(p) -> bool:
    return getattr(p, "ds_nvme", False) and _param_swapper is not None


def read_partition(p) -> torch.Tensor:
    """This rank's slice of ``p`` as a tensor that is valid until ``release_partition(p)``."""
    if _on_nvme(p):
        from ..swap_tensor.partitioned_param_swapper import PartitionedParamStatus
        if getattr(p.ds_tensor, "status", None) != PartitionedParamStatus.AVAILABLE:
            _param_swapper.swap_in([p], async_op=False)
    return p.ds_tensor


def release_partition(p) -> None:
    if _on_nvme(p):
        _param_swapper.remove_partition_and_release_buffers([p])


def write_partition(p, values: torch.Tensor) -> None:
    """Replace this rank's slice of ``p`` (``values``: the slice, or a prefix of it) and persist it."""
    if _on_nvme(p):
        read_partition(p)
        p.ds_tensor[:values.numel()].copy_(values.reshape(-1).to(p.ds_tensor.dtype))
        _param_swapper.swap_out_and_release([p])
    else:
        p.ds_tensor[:values.numel()].copy_(values.reshape(-1).to(p.ds_tensor.dtype))


def _group_info(group):
    if dist.is_initialized():
        return dist.get_world_size(group=group), dist.get_rank(group=group)
    return 1, 0


def _convert(param: torch.nn.Parameter, group, remote_device: Optional[str], pin_memory: bool, persistence_threshold: int,
             dtype: Optional[torch.dtype]):
    if is_zero_param(param):
        return
    ws, rank = _group_info(group)
    if dtype is not None and param.dtype.is_floating_point and param.dtype != dtype:
        param.data = param.data.to(dtype)
    if ws > 1 and param.data.device.type != "meta":
        # every rank initialised its own copy (possibly from a different RNG state): rank 0's values are THE values, or the
        # gathered weight would be a patchwork of differently seeded slices and persistent parameters would differ per rank
        dist.broadcast(param.data, src=dist.get_global_rank(group, 0) if group is not None else 0, group=group)
    param.ds_id = next(_param_id)
    param.ds_shape = param.data.shape
    param.ds_numel = param.data.numel()
    param.ds_process_group = group
    param.ds_persist = param.ds_numel <= persistence_threshold
    param.ds_active_sub_modules = set()
    param.ds_param_type = ZeroParamType.PARTITIONED
    psize = (param.ds_numel + ws - 1) // ws
    param.ds_partition_numel = psize
    flat = param.data.contiguous().view(-1)
    lo, hi = rank * psize, min(param.ds_numel, (rank + 1) * psize)
    part = torch.zeros(psize, dtype=flat.dtype, device=flat.device)
    if lo < hi:
        part[:hi - lo].copy_(flat[lo:hi])
    if remote_device in ("cpu", "nvme"):
        part = part.cpu()
        if pin_memory and torch.cuda.is_available():
            part = part.pin_memory()
    param.ds_tensor = part
    param.ds_tensor.ds_numel = psize
    param.ds_device = flat.device
    param.ds_status = ZeroParamStatus.AVAILABLE
    param.ds_nvme = False
    if remote_device == "nvme" and _param_swapper is not None and _param_swapper.swappable_tensor(numel=psize) and \
            _param_swapper._aligned(psize) <= _param_swapper.elements_per_buffer and part.dtype == _param_swapper.dtype:
        param.ds_nvme = True
        _param_swapper.swap_out_and_release([param])       # the slice now lives in its file

    def all_gather(param_list: Optional[List] = None, async_op: bool = False, hierarchy: int = 0):
        """Gather ``param_list`` (default: this parameter). With ``async_op`` the collectives are only launched and a
        handle is returned whose ``wait()`` completes them."""
        ps = param_list if param_list is not None else [param]
        all_gather_params(ps, async_op=async_op)
        return AllGatherCoalescedHandle(ps) if async_op else None

    def all_gather_coalesced(params, safe_mode: bool = False, quantize: bool = False):
        """Launch the gathers of ``params`` and return the handle to wait on (ref: partition_parameters.py:1230)."""
        ps = [q for q in params if q.ds_status == ZeroParamStatus.NOT_AVAILABLE]
        if not ps:
            return NoGatherCoalescedHandle(list(params))
        all_gather_params(ps, async_op=True)
        return AllGatherCoalescedHandle(list(params))

    def partition(param_list: Optional[List] = None, hierarchy: int = 0, has_been_updated: bool = False):
        for p in (param_list if param_list is not None else [param]):
            _partition_param(p, has_been_updated)

    def partition_numel():
        return param.ds_tensor.ds_numel

    def padding_size():
        return psize * ws - param.ds_numel

    def ds_summary():
        return {"id": param.ds_id, "status": param.ds_status.name, "numel": param.numel(), "ds_numel": param.ds_numel,
                "shape": tuple(param.shape), "ds_shape": tuple(param.ds_shape), "requires_grad": param.requires_grad,
                "persist": param.ds_persist, "device": str(param.device)}

    param.all_gather, param.partition, param.partition_numel = all_gather, partition, partition_numel
    param.all_gather_coalesced = all_gather_coalesced
    param.padding_size, param.ds_summary = padding_size, ds_summary
    if not param.ds_persist:
        _partition_param(param, has_been_updated=False)


@torch.no_grad()
def _partition_param(p, has_been_updated: bool = False):
    """Release the full tensor (first writing this rank's slice back if the full tensor was modified)."""
    if p.ds_status == ZeroParamStatus.INFLIGHT and getattr(p, "_ds_inflight", None) is not None:
        _wait_inflight(p)
    if p.ds_status != ZeroParamStatus.AVAILABLE:
        return
    if has_been_updated:
        ws, rank = _group_info(p.ds_process_group)
        psize = p.ds_partition_numel
        flat = p.data.contiguous().view(-1)
        lo, hi = rank * psize, min(p.ds_numel, (rank + 1) * psize)
        if lo < hi:
            write_partition(p, flat[lo:hi])
    if p.ds_persist:
        return
    p.data = torch.empty(0, dtype=p.dtype, device=p.ds_device)
    p.ds_status = ZeroParamStatus.NOT_AVAILABLE


@torch.no_grad()
def all_gather_params(params: Iterable, async_op: bool = False) -> None:
    """Make the full tensors of ``params`` available (no-op for those that already are). ``async_op``: only LAUNCH the
    collectives (status ``INFLIGHT``); a later blocking call on the same parameters waits for them. This is what the
    trace-driven prefetcher uses to overlap the next modules' gathers with the current module's compute."""
    for p in params:
        if not is_zero_param(p) or p.ds_status == ZeroParamStatus.AVAILABLE:
            continue
        if p.ds_status == ZeroParamStatus.INFLIGHT:
            if not async_op:
                _wait_inflight(p)
            continue
        ws, _ = _group_info(p.ds_process_group)
        psize = p.ds_partition_numel
        local = read_partition(p).to(p.ds_device, non_blocking=True)
        if ws == 1:
            full = local.clone()
            release_partition(p)
        elif getattr(p, "ds_quantized_gather", False) and local.dtype.is_floating_point and psize >= QWZ_GROUP:
            # ZeRO++ qwZ: the slice travels as int8 codes + one (scale) per 256 values, a quarter / half of the bytes
            # (ref: partition_parameters.py all_gather with quantization, csrc/quantization/swizzled_quantize.cu)
            full = _quantized_all_gather(local, ws, p.ds_process_group)
            release_partition(p)
        else:
            full = torch.empty(psize * ws, dtype=local.dtype, device=p.ds_device)
            p.ds_status = ZeroParamStatus.INFLIGHT
            work = dist.all_gather_into_tensor(full, local.contiguous(), group=p.ds_process_group, async_op=async_op)
            if async_op:
                p._ds_inflight = (work, full)
                continue
            release_partition(p)
        p.data = full.narrow(0, 0, p.ds_numel).view(p.ds_shape)
        p.ds_status = ZeroParamStatus.AVAILABLE


QWZ_GROUP = 256


def _quantized_all_gather(local: torch.Tensor, ws: int, group) -> torch.Tensor:
    from ...ops.quantizer import dequantize_int, quantize_int
    flat = local.reshape(-1)
    pad = (-flat.numel()) % QWZ_GROUP
    if pad:
        flat = torch.cat([flat, flat.new_zeros(pad)])
    codes, params = quantize_int(flat.float(), flat.numel() // QWZ_GROUP, 8, symmetric=True)
    all_codes = torch.empty((ws, ) + tuple(codes.shape), dtype=codes.dtype, device=codes.device)
    all_params = torch.empty((ws, ) + tuple(params.shape), dtype=params.dtype, device=params.device)
    dist.all_gather_into_tensor(all_codes.view(-1), codes.reshape(-1).contiguous(), group=group)
    dist.all_gather_into_tensor(all_params.view(-1), params.reshape(-1).contiguous(), group=group)
    rows = [dequantize_int(all_codes[r], all_params[r], 8, local.dtype).reshape(-1)[:local.numel()] for r in range(ws)]
    return torch.cat(rows)


def _wait_inflight(p) -> None:
    work, full = p._ds_inflight
    if work is not None:
        work.wait()
    release_partition(p)         # the staging buffer was the collective's source until now
    p._ds_inflight = None
    p.data = full.narrow(0, 0, p.ds_numel).view(p.ds_shape)
    p.ds_status = ZeroParamStatus.AVAILABLE


# ---- handles returned by asynchronous gathers (ref: partition_parameters.py:640-780) --------------------------------------
class NoGatherHandle:
    """Nothing is in flight (one rank, or the parameter is already resident)."""

    def __init__(self, param) -> None:
        if param.ds_status == ZeroParamStatus.INFLIGHT:
            _wait_inflight(param)
        self.__param = param

    def wait(self, **kwargs) -> None:
        self.__param.ds_status = ZeroParamStatus.AVAILABLE


class NoGatherCoalescedHandle:

    def __init__(self, params) -> None:
        self.__params = list(params)
        self.__complete = False

    def wait(self, **kwargs) -> None:
        if self.__complete:
            return
        for p in self.__params:
            if p.ds_status == ZeroParamStatus.INFLIGHT:
                _wait_inflight(p)
        self.__complete = True


class AllGatherHandle:
    """One parameter's gather: ``wait()`` blocks on the collective and installs the full tensor."""

    def __init__(self, handle, param, quantization=None) -> None:
        self.__handle, self.__param, self.__quantization = handle, param, quantization

    def wait(self, handle_dependency: bool = True) -> None:
        if self.__param.ds_status == ZeroParamStatus.INFLIGHT:
            _wait_inflight(self.__param)
        elif self.__handle is not None and hasattr(self.__handle, "wait"):
            self.__handle.wait()


class AllGatherCoalescedHandle:
    """Several parameters gathered together; a second ``wait()`` is an error, as in the reference."""

    def __init__(self, params, allgather_handle=None, partitions=None, world_size: Optional[int] = None, use_secondary_tensor=False,
                 quantization=None) -> None:
        self.params = list(params)
        self.complete = False
        for p in self.params:
            if p.ds_status not in (ZeroParamStatus.INFLIGHT, ZeroParamStatus.AVAILABLE):
                raise RuntimeError(f"expected param {p.ds_summary()} to not be available")

    def wait(self, handle_dependency: bool = True) -> None:
        if self.complete:
            return
        for p in self.params:
            if p.ds_status == ZeroParamStatus.INFLIGHT:
                _wait_inflight(p)
        self.complete = True


class MultipleAllGatherHandles:

    def __init__(self, handles) -> None:
        self.handles = list(handles)

    def wait(self, handle_dependency: bool = True) -> None:
        for h in self.handles:
            h.wait(handle_dependency=handle_dependency)


class AllReduceCoalescedHandle:
    """Fetch by all-reduce (``use_all_reduce_for_fetch_params``): every rank contributed its slice in place of a zero
    buffer; after the reduction each parameter's data is a view of that buffer."""

    def __init__(self, handle, params) -> None:
        self.handle, self.params, self.complete = handle, list(params), False

    def wait(self, handle_dependency: bool = True) -> None:
        if self.complete:
            return
        if self.handle is not None and hasattr(self.handle, "wait"):
            self.handle.wait()
        for p in self.params:
            if p.ds_status == ZeroParamStatus.INFLIGHT:
                _wait_inflight(p) if getattr(p, "_ds_inflight", None) is not None else setattr(p, "ds_status", ZeroParamStatus.AVAILABLE)
        self.complete = True


class QuantizationInfo:
    """What travels next to the int8 codes of a quantised gather (ZeRO++ qwZ)."""

    def __init__(self, quantized_param=None, backend=None, quant_handle=None, scale_buffer=None, partition_sz: int = 0,
                 world_size: int = 1) -> None:
        self.quantized_param, self.backend, self.quant_handle = quantized_param, backend, quant_handle
        self.scale_buffer, self.partition_sz, self.world_size = scale_buffer, partition_sz, world_size


class CUDAQuantizer:
    """Block-wise symmetric int8 quantisation of weight slices for qwZ: groups of ``group_size`` values share a scale
    (the fused quantisation kernels of ``ops/quantizer`` on the GPU, their torch formulation on the host)."""
    async_flag = True
    target_group_size = 8000
    group_size_cache: dict = {}

    def __init__(self, group_size: int = QWZ_GROUP) -> None:
        self.group_size = group_size

    def _groups(self, numel: int) -> int:
        return max(1, -(-numel // self.group_size))

    def quantize(self, param: torch.Tensor, groups: Optional[int] = None):
        from ...ops.quantizer import quantize_int
        flat = param.reshape(-1)
        groups = groups or self._groups(flat.numel())
        per = -(-flat.numel() // groups)
        per += per % 2                                        # the packing kernels want an even group length
        pad = groups * per - flat.numel()
        if pad:
            flat = torch.cat([flat, flat.new_zeros(pad)])
        return quantize_int(flat.float(), groups, 8, symmetric=True)

    def dequantize(self, quantized_param: torch.Tensor, scale: torch.Tensor, dtype: torch.dtype = torch.float32) -> torch.Tensor:
        from ...ops.quantizer import dequantize_int
        return dequantize_int(quantized_param, scale, 8, dtype)


def free_param(param) -> None:
    """Drop the full tensor of a parameter nobody uses any more (its partition is untouched)."""
    assert not getattr(param, "ds_active_sub_modules", None), param.ds_summary() if hasattr(param, "ds_summary") else param
    param.data = torch.empty(0, dtype=param.dtype, device=getattr(param, "ds_device", param.device))
    if hasattr(param, "ds_status"):
        param.ds_status = ZeroParamStatus.NOT_AVAILABLE


def get_all_subclasses(cls) -> set:
    return Init._all_subclasses(cls)


def print_rank_0(message, debug: bool = False, force: bool = False) -> None:
    if (debug or force) and (not dist.is_initialized() or dist.get_rank() == 0):
        print(message)


def debug_rank0(msg: str) -> None:
    if not dist.is_initialized() or dist.get_rank() == 0:
        from ...utils.logging import logger
        logger.debug(msg)


def zero_wrapper_for_fp_tensor_constructor(fn, target_fp_dtype: torch.dtype):
    """``fn`` (``torch.empty`` / ``zeros`` / ``ones`` / ``full`` / ``arange`` / ``eye`` / ``randn``) wrapped so that the
    floating-point tensors it creates come out in ``target_fp_dtype`` on the current accelerator by default."""
    import functools

    @functools.wraps(fn)
    def wrapped(*args, **kwargs):
        if kwargs.get("device") is None and torch.cuda.is_available():
            kwargs["device"] = torch.device("cuda", torch.cuda.current_device())
        t = fn(*args, **kwargs)
        return t.to(target_fp_dtype) if t.is_floating_point() and "dtype" not in kwargs else t

    return wrapped


def get_new_tensor_fn_for_dtype(dtype: torch.dtype):
    """A ``Tensor.new_tensor`` replacement whose floating-point results are ``dtype``."""

    def new_tensor(cls, *args, **kwargs):
        t = torch.Tensor.new_empty(cls, []).new_tensor(*args, **kwargs) if not isinstance(cls, torch.Tensor) else \
            torch.Tensor.new_tensor(cls, *args, **kwargs)
        return t.to(dtype) if t.is_floating_point() else t

    return new_tensor


class Init:
    """Context manager: modules constructed inside are partitioned as they are created (``module=...`` converts an
    existing one). ``remote_device`` ('cpu') keeps slices in host memory, ``pin_memory`` pins them."""
    _active = 0
    _suspended = 0      # > 0 while a HuggingFace from_pretrained runs inside the context (see _patch_hf_weight_init)

    def __init__(self, module: Optional[torch.nn.Module] = None, data_parallel_group=None, mem_efficient_linear: bool = True,
                 remote_device: Optional[str] = None, pin_memory: bool = False, config_dict_or_path=None, config=None,
                 enabled: bool = True, dtype: Optional[torch.dtype] = None, mpu=None, zero_param_parallel_group=None,
                 zero_quantized_weights: bool = False, zero_quantized_nontrainable_weights: bool = False,
                 sequence_data_parallel_group=None, param_swapper=None):
        cfg = config_dict_or_path if config_dict_or_path is not None else config
        zc = {}
        if isinstance(cfg, dict):
            zc = cfg.get("zero_optimization", {}) or {}
            if dtype is None:
                dtype = torch.float16 if (cfg.get("fp16") or {}).get("enabled") else \
                    torch.bfloat16 if (cfg.get("bf16") or {}).get("enabled") else None
            off = zc.get("offload_param") or {}
            if remote_device is None and off.get("device") in ("cpu", "nvme"):
                remote_device, pin_memory = off["device"], bool(off.get("pin_memory", pin_memory))
        self.enabled = enabled
        self.group = data_parallel_group if data_parallel_group is not None else sequence_data_parallel_group
        self.remote_device, self.pin_memory, self.dtype = remote_device, pin_memory, dtype
        self.persistence_threshold = int(zc.get("stage3_param_persistence_threshold", zc.get("param_persistence_threshold", 0)))
        self._wrapped = []
        if module is not None and enabled:
            self._convert_module(module)

    # ---- partitioning geometry (ref: partition_parameters.py:2103-2117) ----------------------------------------------------
    def get_partition_dp_group(self, param):
        return param.ds_process_group

    def get_dp_process_group(self):
        """The group parameters are partitioned over (``None`` = the default world group)."""
        return self.group

    def get_partition_rank(self) -> int:
        return _group_info(self.group)[1]

    @property
    def num_partitions(self) -> int:
        return _group_info(self.group)[0]

    def _convert_module(self, module: torch.nn.Module):
        for p in module.parameters(recurse=True):
            _convert(p, self.group, self.remote_device, self.pin_memory, self.persistence_threshold, self.dtype)

    @staticmethod
    def _all_subclasses(cls):
        seen, stack = set(), [cls]
        while stack:
            c = stack.pop()
            for sub in c.__subclasses__():
                if sub not in seen:
                    seen.add(sub)
                    stack.append(sub)
        return seen

    def __enter__(self):
        """Wrap ``__init__`` of every ``nn.Module`` subclass: when a module's constructor returns, the parameters it
        registered directly are partitioned (children were handled by their own constructors), so initialisers that run
        inside ``__init__`` still see full tensors."""
        if not self.enabled:
            return self
        Init._active += 1
        if Init._active > 1:
            return self
        ctx = self
        self._wrapped = []

        def wrap(cls):
            orig = cls.__dict__.get("__init__")
            if orig is None or getattr(orig, "_ds_zero_wrapped", False):
                return
            import functools

            @functools.wraps(orig)
            def init(mod, *a, **k):
                mod._ds_init_depth = getattr(mod, "_ds_init_depth", 0) + 1
                try:
                    orig(mod, *a, **k)
                finally:
                    mod._ds_init_depth -= 1
                if mod._ds_init_depth == 0 and Init._active > 0 and Init._suspended == 0:
                    for p in mod.parameters(recurse=False):
                        _convert(p, ctx.group, ctx.remote_device, ctx.pin_memory, ctx.persistence_threshold, ctx.dtype)

            init._ds_zero_wrapped = True
            cls.__init__ = init
            self._wrapped.append((cls, orig))

        for cls in self._all_subclasses(torch.nn.Module):
            wrap(cls)
        self._patch_hf_weight_init()
        return self

    def _patch_hf_weight_init(self):
        """HuggingFace models re-initialise every sub-module AFTER construction (``post_init`` -> ``_initialize_weights``),
        i.e. after this context has already partitioned the parameters. Their initialisers would then write into empty
        placeholders. While the context is active the per-module initialiser therefore runs with the module's own
        parameters gathered; rank 0's values are broadcast and re-partitioned on the way out. (HF does the same from its side
        for the ``deepspeed`` package; this is that hook for this one.)"""
        import sys
        self._hf_patch = None
        mu = sys.modules.get("transformers.modeling_utils")
        base = getattr(mu, "PreTrainedModel", None) if mu is not None else None
        orig = getattr(base, "_initialize_weights", None) if base is not None else None
        if orig is None or getattr(orig, "_ds_zero_wrapped", False):
            return
        import functools

        @functools.wraps(orig)
        def initialize(model, module, *a, **k):
            own = [p for p in module.parameters(recurse=False) if is_zero_param(p)]
            if not own or getattr(module, "_is_hf_initialized", False):
                return orig(model, module, *a, **k)
            with GatheredParameters(own, modifier_rank=0):
                return orig(model, module, *a, **k)

        initialize._ds_zero_wrapped = True
        base._initialize_weights = initialize
        # ``from_pretrained`` builds the model on the meta device and then loads tensors by shape, which cannot work on
        # partitioned placeholders: it runs with partitioning suspended and the finished model is partitioned in one go
        # (every rank reads the checkpoint; peak memory is one full copy per rank, as without this context)
        orig_fp = base.__dict__["from_pretrained"]
        fp_func = orig_fp.__func__ if isinstance(orig_fp, classmethod) else orig_fp
        ctx = self

        @functools.wraps(fp_func)
        def from_pretrained(cls, *a, **k):
            Init._suspended += 1
            try:
                model = fp_func(cls, *a, **k)
            finally:
                Init._suspended -= 1
            if Init._suspended == 0 and Init._active > 0:
                ctx._convert_module(model)
            return model

        base.from_pretrained = classmethod(from_pretrained)
        self._hf_patch = (base, orig, orig_fp)

    def __exit__(self, *exc):
        if not self.enabled:
            return False
        Init._active -= 1
        if Init._active == 0:
            for cls, orig in self._wrapped:
                cls.__init__ = orig
            self._wrapped = []
            if getattr(self, "_hf_patch", None):
                base, orig, orig_fp = self._hf_patch
                base._initialize_weights = orig
                base.from_pretrained = orig_fp
                self._hf_patch = None
        return False


class GatheredParameters:
    """Temporarily gather ``params``; with ``modifier_rank`` set, that rank's in-place modifications are broadcast and
    written back to every rank's slice on exit."""

    def __init__(self, params, modifier_rank: Optional[int] = None, fwd_module: Optional[torch.nn.Module] = None,
                 enabled: bool = True):
        if isinstance(params, torch.nn.Module):
            params = list(params.parameters())
        elif isinstance(params, torch.Tensor):
            params = [params]
        self.params = [p for p in (params or []) if is_zero_param(p)]
        self.enabled = enabled and len(self.params) > 0
        self.src_rank = modifier_rank
        self._was_available = []

    def __enter__(self):
        if not self.enabled:
            return
        self._was_available = [p.ds_status == ZeroParamStatus.AVAILABLE for p in self.params]
        all_gather_params(self.params)

    def __exit__(self, *exc):
        if not self.enabled:
            return False
        if self.src_rank is not None and dist.is_initialized():
            for p in self.params:
                if dist.get_world_size(group=p.ds_process_group) > 1:
                    dist.broadcast(p.data, dist.get_global_rank(p.ds_process_group, self.src_rank), group=p.ds_process_group)
        for p, was in zip(self.params, self._was_available):
            if self.src_rank is not None:
                _partition_param(p, has_been_updated=True)
                if was and p.ds_status != ZeroParamStatus.AVAILABLE:
                    all_gather_params([p])
            elif not was:
                _partition_param(p, has_been_updated=False)
        return False


_external_params = {}


def register_external_parameter(module: torch.nn.Module, parameter: torch.nn.Parameter):
    """Declare that ``module``'s forward uses ``parameter`` owned by another module, so the stage-3 hooks gather it."""
    if not isinstance(parameter, torch.nn.Parameter):
        raise RuntimeError("Parameter is not a torch.nn.Parameter")
    _external_params.setdefault(id(module), {})[id(parameter)] = parameter
    if not hasattr(module, "_external_params"):
        module._external_params = {}
    module._external_params[id(parameter)] = parameter


def unregister_external_parameter(module: torch.nn.Module, parameter: torch.nn.Parameter):
    if hasattr(module, "_external_params"):
        module._external_params.pop(id(parameter), None)


def shutdown_init_context():
    pass


def restore_init_context():
    pass


# the reference splits Init into a generic "patch every nn.Module subclass" base and the ZeRO part; here it is one class
InsertPostInitMethodToModuleSubClasses = Init
