"""FLOPs / MACs / latency profiler for ``torch.nn.Module`` models (ref: deepspeed/profiling/flops_profiler/profiler.py).

While profiling, the functional ops that dominate transformer cost (linear, matmul family, conv, norms, softmax,
activations, embedding, scaled-dot-product attention) are wrapped by counters; forward pre/post hooks attribute the
counts and the wall time to the module that was running, giving a per-module tree. For the ragged engine (which is
not an ``nn.Module``) :func:`engine_flops` gives the analytical count of one forward from its ArchSpec.
"""
from __future__ import annotations

import time
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


def _einsum_flops(equation, *operands):
    """MACs of an einsum = product of the sizes of every index that appears (each output element sums over the
    contracted indices)."""
    if len(operands) == 1 and isinstance(operands[0], (list, tuple)):
        operands = tuple(operands[0])
    lhs = equation.replace(" ", "").split("->")[0].split(",")
    sizes = {}
    for term, op in zip(lhs, operands):
        term = term.replace("...", "")
        for ch, n in zip(term[::-1], tuple(op.shape)[::-1]):
            sizes[ch] = int(n)
    macs = _prod(sizes.values())
    return 2 * macs, macs


def _baddbmm_flops(inp, a, b, *_, **__):
    macs = a.shape[0] * a.shape[1] * a.shape[2] * b.shape[-1]
    return 2 * macs + inp.numel(), macs


def _norm_flops(mult: int):
    def fn(inp, *_, **__):
        return mult * inp.numel(), 0
    return fn


def _pool_flops(inp, *_, **__):
    return inp.numel(), 0


def _interpolate_flops(inp, size=None, scale_factor=None, *_, **__):
    n = inp.numel()
    if size is not None:
        out = _prod(size) if isinstance(size, (tuple, list)) else int(size) ** max(1, inp.dim() - 2)
        n = _prod(inp.shape[:2]) * out
    elif scale_factor is not None:
        sf = scale_factor if isinstance(scale_factor, (tuple, list)) else [scale_factor] * max(1, inp.dim() - 2)
        n = int(inp.numel() * _prod([float(x) for x in sf]) if all(float(x).is_integer() for x in sf) else inp.numel())
    return n, 0


def _conv_transpose_flops(inp, weight, bias=None, stride=1, padding=0, output_padding=0, groups=1, dilation=1):
    # every input element is multiplied with its kernel slice: in_elems * (out_channels / groups) * kernel_volume
    macs = inp.numel() * (weight.shape[1]) * _prod(weight.shape[2:])
    return 2 * macs + (0 if bias is None else macs // max(1, _prod(weight.shape[2:]))), macs


def _tensor_binary(a, b=None, *_, **__):
    n = a.numel() if hasattr(a, "numel") else 0
    if hasattr(b, "numel"):
        n = max(n, b.numel())
    return n, 0


FUNCTIONAL_PATCHES = {
    (F, "linear"): _linear_flops, (F, "conv1d"): _conv_flops, (F, "conv2d"): _conv_flops, (F, "conv3d"): _conv_flops,
    (F, "conv_transpose1d"): _conv_transpose_flops, (F, "conv_transpose2d"): _conv_transpose_flops,
    (F, "conv_transpose3d"): _conv_transpose_flops,
    (F, "relu"): _elementwise(1), (F, "relu6"): _elementwise(1), (F, "leaky_relu"): _elementwise(1), (F, "prelu"): _elementwise(1),
    (F, "elu"): _elementwise(4), (F, "gelu"): _elementwise(8), (F, "silu"): _elementwise(4), (F, "sigmoid"): _elementwise(4),
    (F, "tanh"): _elementwise(4), (F, "softmax"): _elementwise(5), (F, "log_softmax"): _elementwise(6),
    (F, "layer_norm"): _norm_flops(5), (F, "group_norm"): _norm_flops(5), (F, "batch_norm"): _norm_flops(2),
    (F, "instance_norm"): _norm_flops(5), (F, "rms_norm"): _norm_flops(4),
    (F, "avg_pool1d"): _pool_flops, (F, "avg_pool2d"): _pool_flops, (F, "avg_pool3d"): _pool_flops,
    (F, "max_pool1d"): _pool_flops, (F, "max_pool2d"): _pool_flops, (F, "max_pool3d"): _pool_flops,
    (F, "adaptive_avg_pool1d"): _pool_flops, (F, "adaptive_avg_pool2d"): _pool_flops, (F, "adaptive_avg_pool3d"): _pool_flops,
    (F, "adaptive_max_pool1d"): _pool_flops, (F, "adaptive_max_pool2d"): _pool_flops, (F, "adaptive_max_pool3d"): _pool_flops,
    (F, "interpolate"): _interpolate_flops, (F, "upsample"): _interpolate_flops,
    (F, "embedding"): _embedding_flops,
    (F, "scaled_dot_product_attention"): _sdpa_flops,
    (torch, "matmul"): _matmul_flops, (torch, "mm"): _matmul_flops, (torch, "bmm"): _bmm_flops, (torch, "addmm"): _addmm_flops,
    (torch, "baddbmm"): _baddbmm_flops, (torch, "einsum"): _einsum_flops,
    (torch.Tensor, "matmul"): _matmul_flops, (torch.Tensor, "__matmul__"): _matmul_flops, (torch.Tensor, "mm"): _matmul_flops,
    (torch.Tensor, "bmm"): _bmm_flops, (torch.Tensor, "baddbmm"): _baddbmm_flops,
}
# element-wise tensor arithmetic is counted only on request (``FlopsProfiler(count_elementwise=True)``): it is a rounding
# error next to the GEMMs and wrapping ``Tensor.__add__`` slows every profiled step down
ELEMENTWISE_PATCHES = {
    (torch, "add"): _tensor_binary, (torch, "mul"): _tensor_binary, (torch, "sub"): _tensor_binary, (torch, "div"): _tensor_binary,
    (torch.Tensor, "__add__"): _tensor_binary, (torch.Tensor, "__mul__"): _tensor_binary, (torch.Tensor, "__sub__"): _tensor_binary,
    (torch.Tensor, "__truediv__"): _tensor_binary,
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


def _patch_all(elementwise: bool = False):
    table = dict(FUNCTIONAL_PATCHES)
    if elementwise:
        table.update(ELEMENTWISE_PATCHES)
    for (owner, name), counter in table.items():
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


# ---- formatting / per-module readers under the reference's names (profiler.py:1080-1200) ------------------------------------
def macs_to_string(macs, units: Optional[str] = None, precision: int = 2) -> str:
    return f"{number_to_string(macs, units=units, precision=precision)}MACs"


def flops_to_string(flops, units: Optional[str] = None, precision: int = 2) -> str:
    return f"{number_to_string(flops, units=units, precision=precision)}FLOPS"


def bytes_to_string(b, units: Optional[str] = None, precision: int = 2) -> str:
    return f"{number_to_string(b, units=units, precision=precision)}B"


def params_to_string(params_num, units: Optional[str] = None, precision: int = 2) -> str:
    units = units.replace("B", "G") if units else units
    return number_to_string(params_num, units=units, precision=precision).replace("G", "B").strip()


def duration_to_string(duration, units: Optional[str] = None, precision: int = 2) -> str:
    """Seconds -> ``s`` / ``ms`` / ``us`` text."""
    if units is None:
        units = "s" if duration >= 1 else "ms" if duration >= 1e-3 else "us"
    scale = {"s": 1.0, "ms": 1e3, "us": 1e6}[units]
    return f"{round(duration * scale, precision):g} {units}"


def get_module_flops(module: torch.nn.Module) -> int:
    """FLOPs counted in ``module`` and everything below it during the profiled forward (the counters here are already
    cumulative: an operator is credited to every module on the call stack)."""
    return int(getattr(module, "__flops__", 0))


def get_module_macs(module: torch.nn.Module) -> int:
    return int(getattr(module, "__macs__", 0))


def get_module_duration(module: torch.nn.Module) -> float:
    """Wall time of ``module``'s forward; for a container that has no timing of its own, the sum of its children."""
    d = float(getattr(module, "__duration__", 0.0))
    return d if d > 0 else sum(get_module_duration(c) for c in module.children())


def wrapFunc(func, funcFlopCompute):
    """``func`` wrapped so that every call also adds ``funcFlopCompute(*args) -> (flops, macs)`` to the modules being
    profiled; used to count an operator the profiler does not know."""
    counter = lambda *a, **k: funcFlopCompute(*a, **k)     # noqa: E731
    return _wrap(func, counter)


class FlopsProfiler:
    """``prof = FlopsProfiler(model); prof.start_profile(); model(x); prof.stop_profile();
    prof.get_total_flops(); prof.print_model_profile(); prof.end_profile()``"""

    def __init__(self, model: torch.nn.Module, ds_engine=None, recompute_fwd_factor: float = 0.0,
                 count_elementwise: bool = False):
        self.model = model
        self.ds_engine = ds_engine
        self.recompute_fwd_factor = recompute_fwd_factor
        self.count_elementwise = count_elementwise
        self.started = False
        self.func_patched = False
        self._handles = []

    def start_profile(self, ignore_list=None):
        self.reset_profile()
        _patch_all(self.count_elementwise)
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
        v = self._total_params_of(self.model)          # ZeRO-3 placeholders count with their full size
        return number_to_string(v) if as_string else v

    # ---- aggregation ------------------------------------------------------------------------------------------------
    def _total_params_of(self, module) -> int:
        return sum(p.numel() if not hasattr(p, "ds_numel") else p.ds_numel for p in module.parameters())

    def print_model_aggregated_profile(self, module_depth: int = -1, top_modules: int = 1) -> List[str]:
        """Top ``top_modules`` module classes per depth by parameters, MACs and forward latency (ref: profiler.py
        print_model_aggregated_profile)."""
        info: Dict[int, Dict[str, List[float]]] = {}

        def walk(module, depth):
            row = info.setdefault(depth, {}).setdefault(type(module).__name__, [0, 0, 0.0])
            row[0] += self._total_params_of(module)
            row[1] += module.__macs__
            row[2] += module.__duration__
            for c in module.children():
                walk(c, depth + 1)

        walk(self.model, 0)
        depth = max(info) if module_depth < 0 else min(module_depth, max(info))
        lines = [f"Top {top_modules} modules in terms of params, MACs or fwd latency at different model depths:"]
        for d in range(depth + 1):
            n = min(top_modules, len(info[d]))
            top = lambda k: {name: v[k] for name, v in sorted(info[d].items(), key=lambda kv: kv[1][k], reverse=True)[:n]}
            lines.append(f"depth {d}:")
            lines.append("    params      - " + str({k: number_to_string(v).strip() for k, v in top(0).items()}))
            lines.append("    MACs        - " + str({k: number_to_string(v).strip() + "MACs" for k, v in top(1).items()}))
            lines.append("    fwd latency - " + str({k: f"{v * 1e3:.2f} ms" for k, v in top(2).items()}))
        return lines

    def _engine_summary(self, fwd_s: float) -> List[str]:
        """The training-run header of the reference's report: parallel layout, batch size, fwd / bwd / step latency
        (bwd and step need ``wall_clock_breakdown``), achieved FLOPS per GPU and samples per second."""
        e = self.ds_engine
        if e is None:
            return []
        flops = self.get_total_flops()
        lines = [f"world size:                                                   {getattr(e, 'world_size', 1)}",
                 f"data parallel size:                                           {getattr(e, 'dp_world_size', 1)}",
                 f"model parallel size:                                          {getattr(e, 'mp_world_size', 1)}",
                 f"batch size per GPU:                                           {e.train_micro_batch_size_per_gpu()}"]
        bwd_s = step_s = 0.0
        timers = getattr(e, "timers", None)
        try:
            from ...runtime.engine import BACKWARD_GLOBAL_TIMER, STEP_GLOBAL_TIMER
            if timers is not None and hasattr(timers, "has_timer") and timers.has_timer(BACKWARD_GLOBAL_TIMER):
                bwd_s = timers(BACKWARD_GLOBAL_TIMER).elapsed(reset=False) / 1e3
                step_s = timers(STEP_GLOBAL_TIMER).elapsed(reset=False) / 1e3 if timers.has_timer(STEP_GLOBAL_TIMER) else 0.0
        except Exception:
            pass
        factor = 2.0 + self.recompute_fwd_factor           # backward = 2 x forward (+ the recomputed share of the forward)
        lines.append(f"fwd flops per GPU:                                            {number_to_string(flops)}FLOPs")
        lines.append(f"fwd flops of model = fwd flops per GPU * mp_size:             "
                     f"{number_to_string(flops * getattr(e, 'mp_world_size', 1))}FLOPs")
        lines.append(f"fwd latency:                                                  {fwd_s * 1e3:.2f} ms")
        if fwd_s > 0:
            lines.append(f"fwd FLOPS per GPU = fwd flops per GPU / fwd latency:          {number_to_string(flops / fwd_s)}FLOPS")
        if bwd_s > 0:
            lines.append(f"bwd latency:                                                  {bwd_s * 1e3:.2f} ms")
            lines.append(f"bwd FLOPS per GPU = {factor:g} * fwd flops per GPU / bwd latency:    "
                         f"{number_to_string(factor * flops / bwd_s)}FLOPS")
            lines.append(f"fwd+bwd FLOPS per GPU:                                        "
                         f"{number_to_string((1 + factor) * flops / (fwd_s + bwd_s))}FLOPS")
        if step_s > 0:
            lines.append(f"step latency:                                                 {step_s * 1e3:.2f} ms")
        it = fwd_s + bwd_s + step_s
        if it > 0:
            lines.append(f"iter latency:                                                 {it * 1e3:.2f} ms")
            if bwd_s > 0:
                lines.append(f"FLOPS per GPU = {1 + factor:g} * fwd flops per GPU / iter latency:       "
                             f"{number_to_string((1 + factor) * flops / it)}FLOPS")
                lines.append(f"samples/second:                                               "
                             f"{e.train_micro_batch_size_per_gpu() * getattr(e, 'dp_world_size', 1) / it:.2f}")
        return lines

    def is_expert_tensor_parallelism_enabled(self) -> bool:
        from ...moe.layer import MoE
        for m in self.model.modules():
            if isinstance(m, MoE) and hasattr(m, "enable_expert_tensor_parallelism"):
                return bool(m.enable_expert_tensor_parallelism)
        return False

    def print_model_profile(self, profile_step: int = 1, module_depth: int = -1, top_modules: int = 1, detailed: bool = True,
                            output_file: Optional[str] = None):
        """Summary, per-depth top modules and (``detailed``) the module tree with each module's share of parameters, MACs
        and forward latency plus its achieved FLOPS (ref: profiler.py print_model_profile)."""
        dur = self.get_total_duration()
        total_params = self._total_params_of(self.model)
        total_macs = max(1, self.get_total_macs())
        lines = ["-------------------------- Flops Profiler --------------------------",
                 f"Profile Summary at step {profile_step}:",
                 "Notations: data parallel size (dp_size), model parallel size(mp_size), number of parameters (params), "
                 "number of multiply-accumulate operations(MACs), number of floating-point operations (flops), "
                 "floating-point operations per second (FLOPS), fwd latency (forward propagation latency), "
                 "bwd latency (backward propagation latency), step (weights update latency), iter latency (sum of fwd, bwd and step latency)",
                 ""]
        lines += self._engine_summary(dur)
        lines += [f"params:               {number_to_string(total_params)}",
                  f"fwd MACs:             {self.get_total_macs(True)}",
                  f"fwd flops:            {self.get_total_flops(True)}",
                  f"fwd latency:          {self.get_total_duration(True)}"]
        if dur > 0:
            lines.append(f"fwd FLOPS:            {number_to_string(self.get_total_flops() / dur)}FLOPS")
        lines += ["", "----------------------------- Aggregated Profile per GPU -----------------------------"]
        lines += self.print_model_aggregated_profile(module_depth=module_depth, top_modules=top_modules)
        if detailed:
            lines += ["", "------------------------------ Detailed Profile per GPU ------------------------------",
                      "Each module profile is listed after its name in the following order:",
                      "params, percentage of total params, MACs, percentage of total MACs, fwd latency, percentage of total fwd "
                      "latency, fwd FLOPS", ""]

            def walk(module, name, depth):
                if module_depth >= 0 and depth > module_depth:
                    return
                p = self._total_params_of(module)
                d = module.__duration__
                fl = f"{number_to_string(module.__flops__ / d)}FLOPS" if d > 0 else "0 FLOPS"
                lines.append("  " * depth + f"({name}): {type(module).__name__}({number_to_string(p).strip()}, "
                             f"{100.0 * p / max(1, total_params):.2f}% Params, {number_to_string(module.__macs__)}MACs, "
                             f"{100.0 * module.__macs__ / total_macs:.2f}% MACs, {d * 1e3:.2f} ms, "
                             f"{100.0 * d / dur if dur > 0 else 0.0:.2f}% latency, {fl})")
                for n, c in module.named_children():
                    walk(c, n, depth + 1)

            walk(self.model, type(self.model).__name__, 0)
        lines.append("------------------------------------------------------------------------------")
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
