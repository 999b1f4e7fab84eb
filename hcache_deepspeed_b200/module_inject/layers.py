"""Tensor-parallel building blocks used by AutoTP and the policies.

``LinearLayer``            column-parallel: owns a slice of the OUTPUT features, no communication
``LinearAllreduce``        row-parallel: owns a slice of the INPUT features, output all-reduced (bias added after)
``LmHeadLinearAllreduce``  row-parallel LM head whose input is the full hidden state (it slices it itself)
``EmbeddingLayer`` / ``OPTEmbedding`` / ``Normalize`` / ``RMSNormalize``   thin parameter holders for policies
``RowParallel`` / ``ColumnParallel``  autograd functions for training with AutoTP
Reference: ``deepspeed/module_inject/layers.py`` (RowParallel :51, ColumnParallel :80, LinearAllreduce :565,
LinearLayer :419, LmHeadLinearAllreduce :660, EmbeddingLayer :730, OPTEmbedding :750, RMSNormalize :775).
"""
from __future__ import annotations

from typing import Optional

import torch
import torch.nn.functional as F

from .. import comm as dist
from .tp_shard import get_shard_size_list


def _ws(group) -> int:
    return dist.get_world_size(group=group) if (group is not None and dist.is_initialized()) else 1


class RowParallel(torch.autograd.Function):
    """Forward: all-reduce the partial outputs; backward: identity."""

    @staticmethod
    def forward(ctx, group, x: torch.Tensor, is_inference_mode: bool = False):
        ctx.group = group
        if group is None or _ws(group) == 1:
            return x
        out = x.contiguous().clone() if x.requires_grad else x.contiguous()
        if is_inference_mode:
            dist.inference_all_reduce(out, group=group)
        else:
            dist.all_reduce(out, group=group)
        return out

    @staticmethod
    def backward(ctx, grad):
        return None, grad, None


class ColumnParallel(torch.autograd.Function):
    """Forward: identity; backward: all-reduce the input gradient."""

    @staticmethod
    def forward(ctx, group, x: torch.Tensor):
        ctx.group = group
        return x

    @staticmethod
    def backward(ctx, grad):
        if ctx.group is None or _ws(ctx.group) == 1:
            return None, grad
        g = grad.contiguous()
        dist.all_reduce(g, group=ctx.group)
        return None, g


class AsyncColumnParallel(torch.autograd.Function):
    """Column-parallel linear whose backward overlaps the input-gradient all-reduce with the weight-gradient GEMM
    (``tp_overlap_comm``; ref: layers.py:83)."""

    @staticmethod
    def forward(ctx, group, x: torch.Tensor, weight: torch.Tensor, bias):
        ctx.use_bias, ctx.group = bias is not None, group
        out = torch.matmul(x, weight.transpose(-1, -2))
        if bias is not None:
            out = out + bias
        ctx.save_for_backward(x, weight)
        return out

    @staticmethod
    def backward(ctx, grad_output):
        x, weight = ctx.saved_tensors
        grad_input = grad_output.matmul(weight).contiguous()
        handle = dist.all_reduce(grad_input, group=ctx.group, async_op=True) if _ws(ctx.group) > 1 else None
        g2 = grad_output.reshape(-1, grad_output.shape[-1])
        grad_weight = g2.t().matmul(x.reshape(-1, x.shape[-1]))
        grad_bias = g2.sum(0) if ctx.use_bias else None
        if handle is not None:
            handle.wait()
        return None, grad_input, grad_weight, grad_bias


# ---- AutoTP mode (inference: frozen shards; training: shards that get gradients and can be gathered for saving) --------
def get_auto_tp_mode():
    from ..runtime.tensor_parallel import config as tpc
    return getattr(tpc, "_CURRENT_MODE", tpc.AUTOTP_MODE.INFERENCE)


def is_autotp_training_mode() -> bool:
    from ..runtime.tensor_parallel.config import AUTOTP_MODE
    return get_auto_tp_mode() == AUTOTP_MODE.TRAINING


def set_autotp_mode(training: bool = False) -> None:
    from ..runtime.tensor_parallel import config as tpc
    tpc._CURRENT_MODE = tpc.AUTOTP_MODE.TRAINING if training else tpc.AUTOTP_MODE.INFERENCE


def configure_tensor_parallel_runtime(config) -> None:
    """Copy the runtime switches of a tensor-parallel config (``tp_overlap_comm``) onto the layer classes."""
    for key in ("tp_overlap_comm", ):
        if hasattr(config, key):
            setattr(TensorParallel_Layer, key, getattr(config, key))
        elif isinstance(config, dict) and key in config:
            setattr(TensorParallel_Layer, key, config[key])


class TensorParallel_Layer(torch.nn.Module):
    """Base of the layers that shard THEMSELVES from a full module: ``Layer(module, mp_group)`` takes the module's
    parameters, keeps this rank's part (``_tp_partition``) and can reassemble the full tensors for saving
    (``gather_params``). ``name`` feeds the uneven-shard rules of ``tp_shard``; ``keep_module_on_host`` leaves shards on the
    CPU; ``tp_overlap_comm`` selects the overlapped backward (ref: layers.py:142-296)."""
    keep_module_on_host: bool = False
    tp_overlap_comm: bool = False

    def __init__(self, mp_group=None, **kwargs):
        super().__init__()
        self.support_training = False
        self.mp_group = mp_group
        live = mp_group is not None and dist.is_initialized()
        self.tp_world_size = dist.get_world_size(group=mp_group) if live else 1
        self.tp_index = dist.get_rank(group=mp_group) if live else 0
        if kwargs.get("name") is not None:
            self.name = kwargs["name"]
        elif not hasattr(self, "name"):
            self.name = None

    @classmethod
    def set_keep_module_on_host(cls, value: bool) -> None:
        TensorParallel_Layer.keep_module_on_host = bool(value)

    def is_training_mode(self) -> bool:
        return is_autotp_training_mode()

    def gather_params(self, params_list):
        raise NotImplementedError

    def _tp_partition(self, params_list):
        raise NotImplementedError

    def config_tp_params(self, weight) -> None:
        """Mark a shard: trainable or frozen by mode, tagged as tensor-parallel, with the gather / partition hooks that
        checkpoint saving uses."""
        if self.is_training_mode():
            assert self.support_training, "No implementation of backward."
        if weight is None:
            return
        weight.requires_grad = bool(self.is_training_mode())
        weight.tensor_model_parallel = True
        weight.ds_is_replaced_module = True
        weight.gather_params = self.gather_params
        weight._tp_partition = self._tp_partition

    def move(self, tensor: torch.Tensor) -> torch.Tensor:
        """This rank's shard on its final device, as its own storage (so the full tensor can be freed)."""
        if tensor.is_meta:
            return tensor
        if TensorParallel_Layer.keep_module_on_host or not torch.cuda.is_available():
            return tensor.detach().clone() if tensor.device.type == "cpu" else tensor.to("cpu")
        return tensor.to(torch.device("cuda", torch.cuda.current_device()), copy=True)

    def __deepcopy__(self, memo):
        import copy
        new = self.__class__.__new__(self.__class__)
        memo[id(self)] = new
        for k, v in vars(self).items():
            setattr(new, k, v if k == "mp_group" else copy.deepcopy(v, memo))
        return new

    def _split(self, t: torch.Tensor, dim: int) -> torch.Tensor:
        sizes = get_shard_size_list(t.shape[dim], self.tp_world_size, self.name)
        return self.move(t.data.split(sizes, dim=dim)[self.tp_index]).detach()

    def _gather(self, t: torch.Tensor, dim: int) -> torch.Tensor:
        if self.tp_world_size == 1:
            return t.data
        full = t.shape[dim] * self.tp_world_size if not hasattr(t, "_tp_full_size") else t._tp_full_size
        sizes = get_shard_size_list(full, self.tp_world_size, self.name)
        parts = []
        for r, n in enumerate(sizes):
            shape = list(t.shape)
            shape[dim] = n
            buf = t.data.contiguous() if r == self.tp_index else torch.empty(shape, dtype=t.dtype, device=t.device)
            dist.broadcast(buf, src=dist.get_global_rank(self.mp_group, r), group=self.mp_group)
            parts.append(buf)
        return torch.cat(parts, dim=dim)

    def _take(self, module: torch.nn.Module):
        self.weight = module.weight
        self.bias = getattr(module, "bias", None)


class LinearAllreduce(TensorParallel_Layer):
    """Row-parallel. ``LinearAllreduce(weight, bias, mp_group)`` wraps an already sharded weight;
    ``LinearAllreduce(module, mp_group)`` shards the module's weight along its input features itself."""

    def __init__(self, weight, bias: Optional[torch.Tensor] = None, mp_group=None, **kwargs):
        if isinstance(weight, torch.nn.Module):
            module, group = weight, (bias if mp_group is None else mp_group)
            super().__init__(group, **kwargs)
            self._take(module)
            self.support_training = True
            self._tp_partition([self.weight, self.bias])
            self.config_tp_params(self.weight)
            if self.bias is not None:
                self.bias.requires_grad = bool(self.is_training_mode())
            return
        super().__init__(mp_group, **kwargs)
        self.weight = weight if isinstance(weight, torch.nn.Parameter) else torch.nn.Parameter(weight, requires_grad=False)
        self.bias = None if bias is None else (bias if isinstance(bias, torch.nn.Parameter) else
                                               torch.nn.Parameter(bias, requires_grad=False))

    @torch.no_grad()
    def _tp_partition(self, params_list):
        w = params_list[0]
        if w is not None and self.tp_world_size > 1:
            w._tp_full_size = w.shape[1]
            w.data = self._split(w, 1)
        b = params_list[1] if len(params_list) > 1 else None
        if b is not None:
            b.data = self.move(b.data).detach()

    @torch.no_grad()
    def gather_params(self, params_list):
        w = params_list[0]
        if w is not None:
            w.data = self._gather(w, 1)


    def forward(self, x: torch.Tensor) -> torch.Tensor:
        out = torch.matmul(x, self.weight.transpose(-1, -2))
        out = RowParallel.apply(self.mp_group, out, not self.training)
        return out if self.bias is None else out + self.bias

    def extra_repr(self):
        return f"in_features(local)={self.weight.shape[1]}, out_features={self.weight.shape[0]}, bias={self.bias is not None}"


class LmHeadLinearAllreduce(torch.nn.Module):

    def __init__(self, weight, rank: int, world_size: int, bias=None, mp_group=None):
        super().__init__()
        self.weight = weight if isinstance(weight, torch.nn.Parameter) else torch.nn.Parameter(weight, requires_grad=False)
        self.bias = None if bias is None else torch.nn.Parameter(bias, requires_grad=False)
        self.mp_group, self.rank, self.world_size = mp_group, rank, world_size

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        sizes = get_shard_size_list(x.shape[-1], self.world_size, "lm_head")
        start = sum(sizes[:self.rank])
        if self.training and x.requires_grad:
            # each rank back-propagates into ITS slice of x only: the input gradient is complete after a sum over the group
            x = ColumnParallel.apply(self.mp_group, x)
        out = torch.matmul(x[..., start:start + sizes[self.rank]], self.weight.transpose(-1, -2))
        out = RowParallel.apply(self.mp_group, out, not self.training)
        return out if self.bias is None else out + self.bias


class LinearLayer(TensorParallel_Layer):
    """Column-parallel. ``LinearLayer(weight=..., bias=..., mp_group=...)`` wraps sharded tensors;
    ``LinearLayer(module, mp_group, skip_partition=False)`` shards the module's output features itself."""

    def __init__(self, weight_shape=None, dtype=torch.half, weight=None, bias=None, mp_group=None, skip_partition: bool = False,
                 **kwargs):
        if isinstance(weight_shape, torch.nn.Module):
            module = weight_shape
            group = mp_group if mp_group is not None else (dtype if not isinstance(dtype, torch.dtype) else None)
            super().__init__(group, **kwargs)
            self._take(module)
            self.support_training = True
            if not skip_partition:
                self._tp_partition([self.weight, self.bias])
            self.config_tp_params(self.weight)
            if self.bias is not None:
                self.config_tp_params(self.bias)
            return
        super().__init__(mp_group, **kwargs)
        if weight is not None:
            self.weight = weight if isinstance(weight, torch.nn.Parameter) else torch.nn.Parameter(weight, requires_grad=False)
            self.bias = None if bias is None else (bias if isinstance(bias, torch.nn.Parameter) else
                                                   torch.nn.Parameter(bias, requires_grad=False))
        else:
            self.weight = torch.nn.Parameter(torch.empty(weight_shape, dtype=dtype), requires_grad=False)
            self.bias = torch.nn.Parameter(torch.empty(weight_shape[0], dtype=dtype), requires_grad=False) if bias is not None else None
        self.mp_group = mp_group

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        if self.training and x.requires_grad and TensorParallel_Layer.tp_overlap_comm and self.mp_group is not None:
            return AsyncColumnParallel.apply(self.mp_group, x, self.weight, self.bias)
        x = ColumnParallel.apply(self.mp_group, x) if (self.training and x.requires_grad) else x
        out = torch.matmul(x, self.weight.transpose(-1, -2))
        return out if self.bias is None else out + self.bias

    @torch.no_grad()
    def _tp_partition(self, params_list):
        if self.tp_world_size == 1:
            return
        for p in params_list:
            if p is not None:
                p._tp_full_size = p.shape[0]
                p.data = self._split(p, 0)

    @torch.no_grad()
    def gather_params(self, params_list):
        for p in params_list:
            if p is not None:
                p.data = self._gather(p, 0)


class FusedModuleWrapper:
    """Every attribute lookup returns the wrapped block: lets ``prepare_tp_fused_qkvw`` read the block type through
    ``fused_module.module`` whatever the attribute is called."""

    def __init__(self, fused_module: torch.nn.Module):
        self.fused_module = fused_module

    def __getattr__(self, name):
        return self.__dict__["fused_module"]


class fused_LinearLayer(LinearLayer):
    """Column-parallel fused QKV (or fused gate/up) projection: the cut follows the block's fused layout."""

    def __init__(self, module, mp_group, skip_partition: bool = False, **kwargs):
        assert kwargs.get("fused_module") is not None, "'fused_module' is required but not provided"
        self.__dict__["_fused"] = FusedModuleWrapper(kwargs.pop("fused_module"))
        self.__dict__["_fused_kw"] = {k: kwargs.pop(k) for k in ("num_heads", "num_kv_heads") if k in kwargs}
        super().__init__(module, mp_group=mp_group, skip_partition=skip_partition, **kwargs)

    @property
    def fused_module(self):
        return self.__dict__["_fused"]

    @torch.no_grad()
    def _tp_partition(self, params_list):
        from .fusedqkv_utils import prepare_tp_fused_qkvw
        for p in params_list:
            if p is None:
                return
            part = prepare_tp_fused_qkvw(self.fused_module.module, p.data, self.tp_world_size, self.tp_index,
                                         **self.__dict__["_fused_kw"])
            p.data = self.move(part).detach()


class conv_LinearLayer(LinearLayer):
    """Column-parallel layer built from a HuggingFace ``Conv1D`` (weight stored ``[in, out]``): cut along dim 1."""

    @torch.no_grad()
    def _tp_partition(self, params_list):
        w = params_list[0]
        b = params_list[1] if len(params_list) > 1 else None
        sizes = get_shard_size_list(w.shape[1], self.tp_world_size, self.name)
        if b is not None:
            b.data = self.move(b.data.split(sizes, dim=0)[self.tp_index]).detach()
        w.data = self.move(w.data.split(sizes, dim=1)[self.tp_index]).detach()

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        x = ColumnParallel.apply(self.mp_group, x) if (self.training and x.requires_grad) else x
        out = torch.matmul(x, self.weight)
        return out if self.bias is None else out + self.bias


class Conv_LinearALlreduce(LinearAllreduce):
    """Row-parallel layer built from a ``Conv1D``: transposed to ``[out, in]`` first, then cut along the input features."""

    @torch.no_grad()
    def _tp_partition(self, params_list):
        w = params_list[0]
        if w is None:
            return
        wt = w.data.transpose(-1, -2).contiguous()
        sizes = get_shard_size_list(wt.shape[1], self.tp_world_size, self.name)
        w.data = self.move(wt.split(sizes, dim=1)[self.tp_index]).detach()
        b = params_list[1] if len(params_list) > 1 else None
        if b is not None:
            b.data = self.move(b.data).detach()


class Yuan_LinearAllreduce(LinearAllreduce):
    """Output projection of attention with V heads shared by Q/K head pairs (Yuan): see ``shard_value_with_share_qk``."""

    @torch.no_grad()
    def _tp_partition(self, params_list):
        from .fusedqkv_utils import shard_value_with_share_qk
        w, b = shard_value_with_share_qk(params_list[0].data, params_list[1], self.tp_index, self.tp_world_size, False)
        params_list[0].data = self.move(w.data).detach()
        if b is not None:
            params_list[1].data = self.move(b.data).detach()


class Yuan_LinearLayer(LinearLayer):

    @torch.no_grad()
    def _tp_partition(self, params_list):
        from .fusedqkv_utils import shard_value_with_share_qk
        w, b = shard_value_with_share_qk(params_list[0].data, params_list[1], self.tp_index, self.tp_world_size, True)
        params_list[0].data = self.move(w.data).detach()
        if b is not None:
            params_list[1].data = self.move(b.data).detach()


class GateUpPack_LinearLayer(LinearLayer):
    """Column-parallel packed gate+up projection (Phi-3): each half is cut separately so a rank keeps matching slices."""

    @torch.no_grad()
    def _tp_partition(self, params_list):
        from .fusedqkv_utils import shard_chunk_mlp
        w, b = shard_chunk_mlp(params_list[0].data, params_list[1], self.tp_index, self.tp_world_size)
        params_list[0].data = self.move(w.data if isinstance(w, torch.nn.Parameter) else w).detach()
        if b is not None:
            params_list[1].data = self.move(b.data if isinstance(b, torch.nn.Parameter) else b).detach()


class TensorParallelConv2d(torch.nn.Module):
    """A ``Conv2d`` sharded over output channels (``shard_by_oc``) or input channels (then outputs are partial sums);
    the later ranks take the larger shares when channels do not divide evenly, as in the reference (layers.py:656)."""

    def __init__(self, conv: torch.nn.Conv2d, rank: int, world_size: int, shard_by_oc: bool):
        super().__init__()
        self.rank, self.world_size, self.shard_by_oc = rank, world_size, shard_by_oc
        total = conv.weight.shape[0 if shard_by_oc else 1]
        bounds = [0]
        for i in range(world_size - 1, -1, -1):
            bounds.append(bounds[-1] + total // world_size + (1 if i < total % world_size else 0))
        lo, hi = bounds[rank], bounds[rank + 1]
        w = conv.weight.data[lo:hi] if shard_by_oc else conv.weight.data[:, lo:hi]
        b = None
        if conv.bias is not None:
            b = conv.bias.data[lo:hi] if shard_by_oc else conv.bias.data / float(world_size)
        self.conv = torch.nn.Conv2d(w.shape[1], w.shape[0], conv.kernel_size, conv.stride, conv.padding, conv.dilation,
                                    conv.groups, conv.bias is not None, conv.padding_mode)
        self.conv.weight = torch.nn.Parameter(w.clone())
        if b is not None:
            self.conv.bias = torch.nn.Parameter(b.clone())

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        return self.conv(x)


class TensorParallelOcShardConv2d(TensorParallelConv2d):

    def __init__(self, conv, rank, world_size):
        super().__init__(conv, rank, world_size, True)


class TensorParallelIcShardConv2d(TensorParallelConv2d):

    def __init__(self, conv, rank, world_size):
        super().__init__(conv, rank, world_size, False)

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        out = self.conv(x)
        if self.world_size > 1 and dist.is_initialized():
            dist.inference_all_reduce(out)
        return out


class Normalize(torch.nn.Module):

    def __init__(self, dim=None, dtype=torch.float, eps: float = 1e-5, weight=None, bias=None):
        super().__init__()
        if weight is not None:
            self.weight, self.bias = weight, bias
        else:
            self.norm = torch.nn.LayerNorm(dim, eps=eps).to(dtype)
            self.weight, self.bias = self.norm.weight, self.norm.bias
        self.eps = eps

    def forward(self, x):
        return F.layer_norm(x, x.shape[-1:], self.weight, self.bias, eps=self.eps)


class EmbeddingLayer(torch.nn.Module):

    def __init__(self, weight_shape=None, dtype=torch.half, weight=None, bias=None):
        super().__init__()
        self.weight = weight if weight is not None else torch.nn.Parameter(torch.empty(weight_shape, dtype=dtype))

    def forward(self, ids):
        return F.embedding(ids, self.weight)


class OPTEmbedding(EmbeddingLayer):
    """OPT's learned positions are offset by 2 and derived from the attention mask."""

    def __init__(self, weight_shape=None, weight=None, bias=None):
        self.offset = 2
        super().__init__(weight_shape, weight=weight)

    def forward(self, attention_mask: torch.LongTensor, past_key_values_length: int = 0, position_ids=None):
        attention_mask = attention_mask.long()
        positions = (torch.cumsum(attention_mask, dim=1).type_as(attention_mask) * attention_mask).long() - 1
        positions = positions[:, past_key_values_length:]
        return super().forward(positions + self.offset)


class RMSNormalize(torch.nn.Module):

    def __init__(self, dim=None, dtype=torch.float, eps: float = 1e-5, weight=None):
        super().__init__()
        self.weight = weight if weight is not None else torch.nn.Parameter(torch.ones(dim, dtype=dtype))
        self.eps = eps

    def forward(self, x):
        var = x.to(torch.float32).pow(2).mean(-1, keepdim=True)
        x = x * torch.rsqrt(var + self.eps)
        if self.weight.dtype in (torch.float16, torch.bfloat16):
            x = x.to(self.weight.dtype)
        return x * self.weight


class GatherReplacedLayerParams:
    """Context manager that reassembles the full weight of TP-sharded layers (e.g. for saving) and re-shards on exit."""

    def __init__(self, params, module, enabled: bool = True):
        self.enabled = enabled
        self.module = module
        self.layers = [m for m in module.modules() if isinstance(m, (LinearLayer, LinearAllreduce))] if params is None else params

    def __enter__(self):
        if not self.enabled:
            return
        self._saved = []
        for l in self.layers:
            g = l.mp_group
            ws = _ws(g)
            if ws == 1:
                continue
            dim = 1 if isinstance(l, LinearAllreduce) else 0
            parts = [torch.empty_like(l.weight.data) for _ in range(ws)]
            dist.all_gather(parts, l.weight.data.contiguous(), group=g)
            self._saved.append((l, l.weight.data, dim))
            l.weight.data = torch.cat(parts, dim=dim)

    def __exit__(self, *exc):
        if not self.enabled:
            return False
        for l, local, _ in self._saved:
            l.weight.data = local
        return False
