"""FLOPs / MACs / latency profiler for ``torch.nn.Module`` models (ref: deepspeed/profiling/flops_profiler/profiler.py).

While profiling, the functional ops that dominate transformer cost (linear, matmul family, conv, norms, softmax,
activations, embedding, scaled-dot-product attention) are wrapped by counters; forward pre/post hooks attribute the
counts and the wall time to the module that was running, giving a per-module tree. For the ragged engine (which is
not an ``nn.Module``) :func:`engine_flops` gives the analytical count of one forward from its ArchSpec.
"""
from __future__ import annotations

import time
from functools import partial
from typing import Callable, Dict, List, Optional

import torch
import torch.nn.functional as F

_module_stack: List["torch.nn.Module"] = []
_patched: Dict[object, Dict[str, Callable]] = {}


def _prod(xs) -> int:
    p = 1
    for x in xs:
        p *= int(x)
    return p


def _add(flops: int, macs: int) -> None:
    for m in _module_stack:
        m.__flops__ += flops
        m.__macs__ += macs


def _linear_flops(inp, weight, bias=None):
    macs = _prod(inp.shape) * weight.shape[0]
    return 2 * macs, macs


def _matmul_flops(a, b, *_, **__):
    if a.dim() == 0 or b.dim() == 0:
        return 0, 0
    macs = _prod(a.shape) * (b.shape[-1] if b.dim() > 1 else 1)
    return 2 * macs, macs


def _bmm_flops(a, b, *_, **__):
    macs = a.shape[0] * a.shape[1] * a.shape[2] * b.shape[2]
    return 2 * macs, macs


def _addmm_flops(inp, a, b, *_, **__):
    macs = a.shape[0] * a.shape[1] * b.shape[1]
    return 2 * macs + _prod(inp.shape), macs


def _conv_flops(inp, weight, bias=None, stride=1, padding=0, dilation=1, groups=1):
    batch, out_ch = inp.shape[0], weight.shape[0]
    spatial = list(inp.shape[2:])
    k = list(weight.shape[2:])
    nd = len(spatial)
    st = [stride] * nd if isinstance(stride, int) else list(stride)
    pd = [padding] * nd if isinstance(padding, int) else ([0] * nd if isinstance(padding, str) else list(padding))
    dl = [dilation] * nd if isinstance(dilation, int) else list(dilation)
    out_sp = [(spatial[i] + 2 * pd[i] - dl[i] * (k[i] - 1) - 1) // st[i] + 1 for i in range(nd)]
    macs = batch * out_ch * _prod(out_sp) * (weight.shape[1]) * _prod(k)
    return 2 * macs + (batch * out_ch * _prod(out_sp) if bias is not None else 0), macs


def _elementwise(mult: int):
    def fn(inp, *_, **__):
        return mult * _prod(inp.shape), 0
    return fn


def _sdpa_flops(q, k, v, *_, **__):
    b = _prod(q.shape[:-2])
    macs = b * q.shape[-2] * k.shape[-2] * (q.shape[-1] + v.shape[-1])
    return 2 * macs, macs


def _embedding_flops(inp, weight, *_, **__):
    return 0, 0


FUNCTIONAL_PATCHES = {
    (F, "linear"): _linear_flops, (F, "conv1d"): _conv_flops, (F, "conv2d"): _conv_flops, (F, "conv3d"): _conv_flops,
    (F, "relu"): _elementwise(1), (F, "gelu"): _elementwise(8), (F, "silu"): _elementwise(4), (F, "softmax"): _elementwise(5),
    (F, "layer_norm"): _elementwise(5), (F, "embedding"): _embedding_flops,
    (F, "scaled_dot_product_attention"): _sdpa_flops,
    (torch, "matmul"): _matmul_flops, (torch, "bmm"): _bmm_flops, (torch, "addmm"): _addmm_flops,
    (torch.Tensor, "matmul"): _matmul_flops, (torch.Tensor, "__matmul__"): _matmul_flops,
}


def _wrap(orig, counter):
    def wrapped(*args, **kwargs):
        try:
            fl, mc = counter(*args, **kwargs)
            _add(int(fl), int(mc))
        except Exception:
            pass
        return orig(*args, **kwargs)
    wrapped.__wrapped_by_flops_profiler__ = True
    return wrapped


def _patch_all():
    for (owner, name), counter in FUNCTIONAL_PATCHES.items():
        orig = getattr(owner, name, None)
        if orig is None or getattr(orig, "__wrapped_by_flops_profiler__", False):
            continue
        _patched.setdefault(owner, {})[name] = orig
        setattr(owner, name, _wrap(orig, counter))


def _unpatch_all():
    for owner, names in _patched.items():
        for name, orig in names.items():
            setattr(owner, name, orig)
    _patched.clear()


def number_to_string(num: float, units: Optional[str] = None, precision: int = 2) -> str:
    for scale, suffix in ((1e12, "T"), (1e9, "G"), (1e6, "M"), (1e3, "K")):
        if (units is None and abs(num) >= scale) or units == suffix:
            return f"{num / scale:.{precision}f} {suffix}"
    return f"{num:.{precision}f} "


class FlopsProfiler:
    """``prof = FlopsProfiler(model); prof.start_profile(); model(x); prof.stop_profile();
    prof.get_total_flops(); prof.print_model_profile(); prof.end_profile()``"""

    def __init__(self, model: torch.nn.Module, ds_engine=None, recompute_fwd_factor: float = 0.0):
        self.model = model
        self.ds_engine = ds_engine
        self.recompute_fwd_factor = recompute_fwd_factor
        self.started = False
        self.func_patched = False
        self._handles = []

    def start_profile(self, ignore_list=None):
        self.reset_profile()
        _patch_all()
        self.func_patched = True

        def pre(module, inp):
            _module_stack.append(module)
            if torch.cuda.is_available():
                torch.cuda.synchronize()
            module.__start_time__ = time.time()

        def post(module, inp, out):
            if torch.cuda.is_available():
                torch.cuda.synchronize()
            module.__duration__ += time.time() - module.__start_time__
            _module_stack.pop()

        def register(module):
            if ignore_list and type(module) in ignore_list:
                return
            self._handles.append(module.register_forward_pre_hook(pre))
            self._handles.append(module.register_forward_hook(post))

        self.model.apply(register)
        self.started = True

    def stop_profile(self):
        if self.started and self.func_patched:
            _unpatch_all()
            self.func_patched = False
        for h in self._handles:
            h.remove()
        self._handles.clear()

    def reset_profile(self):
        def reset(module):
            module.__flops__ = 0
            module.__macs__ = 0
            module.__duration__ = 0.0
            module.__params__ = sum(p.numel() for p in module.parameters(recurse=False))
        self.model.apply(reset)

    def end_profile(self):
        if not self.started:
            return
        self.stop_profile()
        self.started = False

        def cleanup(module):
            for attr in ("__flops__", "__macs__", "__duration__", "__params__", "__start_time__"):
                if hasattr(module, attr):
                    delattr(module, attr)
        self.model.apply(cleanup)

    def get_total_flops(self, as_string: bool = False):
        v = self.model.__flops__
        return number_to_string(v) + "FLOPs" if as_string else v

    def get_total_macs(self, as_string: bool = False):
        v = self.model.__macs__
        return number_to_string(v) + "MACs" if as_string else v

    def get_total_duration(self, as_string: bool = False):
        v = self.model.__duration__
        return f"{v * 1e3:.2f} ms" if as_string else v

    def get_total_params(self, as_string: bool = False):
        v = sum(p.numel() for p in self.model.parameters())
        return number_to_string(v) if as_string else v

    def print_model_profile(self, profile_step: int = 1, module_depth: int = -1, top_modules: int = 1, detailed: bool = True,
                            output_file: Optional[str] = None):
        lines = ["-------------------------- Flops Profiler --------------------------",
                 f"Profile Summary at step {profile_step}:",
                 f"params:               {self.get_total_params(True)}",
                 f"fwd MACs:             {self.get_total_macs(True)}",
                 f"fwd flops:            {self.get_total_flops(True)}",
                 f"fwd latency:          {self.get_total_duration(True)}"]
        dur = self.get_total_duration()
        if dur > 0:
            lines.append(f"fwd FLOPS:            {number_to_string(self.get_total_flops() / dur)}FLOPS")
        if detailed:
            lines.append("")
            def walk(module, name, depth):
                if module_depth >= 0 and depth > module_depth:
                    return
                lines.append("  " * depth + f"{name or type(module).__name__}: {number_to_string(module.__params__)}params, "
                             f"{number_to_string(module.__macs__)}MACs, {module.__duration__ * 1e3:.2f} ms")
                for n, c in module.named_children():
                    walk(c, n, depth + 1)
            walk(self.model, type(self.model).__name__, 0)
        text = "\n".join(lines)
        if output_file:
            with open(output_file, "w") as f:
                f.write(text + "\n")
        else:
            print(text)
        return text


def get_model_profile(model: torch.nn.Module, input_shape=None, args=(), kwargs=None, print_profile: bool = True,
                      detailed: bool = True, module_depth: int = -1, top_modules: int = 1, warm_up: int = 1,
                      as_string: bool = True, output_file: Optional[str] = None, ignore_modules=None, mode: str = "forward"):
    """Returns ``(flops, macs, params)`` of one forward (ref: profiler.py ``get_model_profile``)."""
    kwargs = kwargs or {}
    if input_shape is not None:
        args = (torch.ones(()).new_empty(tuple(input_shape), dtype=next(model.parameters()).dtype,
                                          device=next(model.parameters()).device), )
    prof = FlopsProfiler(model)
    model.eval()
    with torch.no_grad():
        for _ in range(warm_up):
            model(*args, **kwargs)
        prof.start_profile(ignore_list=ignore_modules)
        model(*args, **kwargs)
    flops, macs, params = prof.get_total_flops(), prof.get_total_macs(), prof.get_total_params()
    if print_profile:
        prof.print_model_profile(detailed=detailed, module_depth=module_depth, top_modules=top_modules, output_file=output_file)
    prof.end_profile()
    if as_string:
        return number_to_string(flops) + "FLOPs", number_to_string(macs) + "MACs", number_to_string(params)
    return flops, macs, params


def engine_flops(spec, n_tokens: int, context: int = 0) -> int:
    """Analytical FLOPs of one ragged forward over ``n_tokens`` new tokens attending to ``context`` earlier ones."""
    inter = spec.moe_intermediate * spec.top_k if spec.is_moe else spec.intermediate
    per_layer = 2 * n_tokens * (spec.hidden * (spec.q_dim + 2 * spec.kv_dim) + spec.q_dim * spec.hidden +
                                (3 if spec.gated_mlp else 2) * spec.hidden * inter)
    attn = 4 * spec.q_dim * n_tokens * (context + (n_tokens + 1) / 2)
    return int(spec.n_layers * (per_layer + attn) + 2 * spec.hidden * spec.vocab)
