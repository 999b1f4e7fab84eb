"""Compression-aware layers. ``LinearLayer_Compress`` is an ``nn.Linear`` that can, independently and on schedule:
quantise its weight (symmetric / asymmetric, straight-through estimator), quantise its input activations (dynamic or
static range), and prune by element (l1 / topk masks), by row, or by attention head.
Reference: ``deepspeed/compression/basic_layer.py`` (QuantAct :16, Embedding_Compress :61, LinearLayer_Compress :121,
Conv2dLayer_Compress :444, BNLayer_Compress :713, ColumnParallelLinear_Compress :835)."""
from __future__ import annotations

import math
from typing import Optional

import torch
import torch.nn.functional as F
from torch import nn


class _STEQuant(torch.autograd.Function):
    """Group-wise fake quantisation with a straight-through gradient."""

    @staticmethod
    def forward(ctx, x, bits: int, symmetric: bool, groups: int):
        shape = x.shape
        v = x.reshape(groups, -1).float()
        if symmetric:
            qmax = 2 ** (bits - 1) - 1
            scale = v.abs().amax(dim=1, keepdim=True).clamp_min(1e-8) / qmax
            q = torch.clamp(torch.round(v / scale), -qmax - 1, qmax) * scale
        else:
            mn, mx = v.amin(dim=1, keepdim=True), v.amax(dim=1, keepdim=True)
            scale = (mx - mn).clamp_min(1e-8) / (2 ** bits - 1)
            q = torch.round((v - mn) / scale) * scale + mn
        return q.reshape(shape).to(x.dtype)

    @staticmethod
    def backward(ctx, g):
        return g, None, None, None


class _STEBinary(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, groups: int):
        v = x.reshape(groups, -1)
        m = v.abs().mean(dim=1, keepdim=True)
        return (v.sign() * m).reshape(x.shape)

    @staticmethod
    def backward(ctx, g):
        return g, None


class _STETernary(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, groups: int):
        v = x.reshape(groups, -1)
        thres = 0.7 * v.abs().mean(dim=1, keepdim=True)
        mask = (v.abs() > thres).float()
        alpha = (v.abs() * mask).sum(dim=1, keepdim=True) / mask.sum(dim=1, keepdim=True).clamp_min(1)
        return (v.sign() * mask * alpha).reshape(x.shape)

    @staticmethod
    def backward(ctx, g):
        return g, None


def fake_quantize(x, bits, symmetric=True, groups=1):
    if bits >= 16:
        return x
    if bits == 1:
        return _STEBinary.apply(x, groups)
    if bits == 2:
        return _STETernary.apply(x, groups)
    return _STEQuant.apply(x, bits, symmetric, groups)


class QuantAct(nn.Module):
    """Activation quantiser with a momentum-tracked static range or a per-call dynamic range."""

    def __init__(self, act_range_momentum: float = 0.95, quant_mode: str = "symmetric"):
        super().__init__()
        self.act_range_momentum, self.quant_mode = act_range_momentum, quant_mode
        self.register_buffer("x_min_max", torch.zeros(2))

    def forward(self, x, num_bits: int, *args):
        if self.training:
            mn, mx = x.detach().min(), x.detach().max()
            if float(self.x_min_max[0]) == float(self.x_min_max[1]):
                self.x_min_max[0], self.x_min_max[1] = mn, mx
            self.x_min_max[0] = self.x_min_max[0] * self.act_range_momentum + mn * (1 - self.act_range_momentum)
            self.x_min_max[1] = self.x_min_max[1] * self.act_range_momentum + mx * (1 - self.act_range_momentum)
        lo, hi = float(self.x_min_max[0]), float(self.x_min_max[1])
        xc = torch.clamp(x, lo, hi) if hi > lo else x
        return fake_quantize(xc, num_bits, self.quant_mode == "symmetric", 1)


class LinearLayer_Compress(nn.Linear):

    def __init__(self, *kargs, bias: bool = True):
        super().__init__(*kargs, bias=bias)
        self.sparse_pruning_method = self.row_pruning_method = self.head_pruning_method = None
        self.activation_quantization_method = None
        self.weight.start_bits = self.weight.target_bits = None
        self.weight.q_period = 0
        self.weight_quantization_enabled_in_forward = False
        self.weight_quantization_enabled = False
        self.sparse_pruning_enabled = self.row_pruning_enabled = self.head_pruning_enabled = False
        self.activation_quantization_enabled = False
        self.weight_quantize_num_groups = 1
        self.weight_symmetric = True

    def __repr__(self):
        return (f"LinearLayer_Compress(in={self.in_features}, out={self.out_features}, bias={self.bias is not None}, "
                f"sparse={self.sparse_pruning_method is not None}, row={self.row_pruning_method is not None}, "
                f"head={self.head_pruning_method is not None}, wq={self.weight.target_bits}, aq={self.activation_quantization_method})")

    # ---- enabling (called once from the config) ---------------------------------------------------------
    def enable_sparse_pruning(self, ratio: float, method: str):
        self.sparse_pruning_ratio, self.sparse_pruning_method = ratio, method
        if method == "l1":
            self.register_buffer("sparse_pruning_mask", self._topk_mask(self.weight.abs(), ratio))
        elif method == "topk":
            self.sparse_mask_scores = nn.Parameter(torch.empty_like(self.weight))
            nn.init.kaiming_uniform_(self.sparse_mask_scores, a=math.sqrt(5))
        else:
            raise NotImplementedError(f"sparse pruning method {method}")

    def enable_row_pruning(self, ratio: float, method: str):
        self.row_pruning_ratio, self.row_pruning_method = ratio, method
        if method == "l1":
            norms = self.weight.abs().sum(dim=1)
            self.register_buffer("row_pruning_mask", self._topk_mask(norms, ratio).reshape(-1, 1))
        elif method == "topk":
            self.row_mask_scores = nn.Parameter(torch.empty(self.out_features, 1))
            nn.init.kaiming_uniform_(self.row_mask_scores, a=math.sqrt(5))
        else:
            raise NotImplementedError(f"row pruning method {method}")

    def enable_head_pruning(self, ratio: float, method: str, num_heads: int):
        assert method in ("topk", "l1"), "head pruning supports topk / l1"
        self.num_heads, self.head_pruning_ratio, self.head_pruning_method = num_heads, ratio, method
        if method == "topk":
            self.head_pruning_scores = nn.Parameter(torch.empty(1, num_heads))
            nn.init.kaiming_uniform_(self.head_pruning_scores, a=math.sqrt(5))
        else:
            per = self.weight.abs().reshape(self.out_features, num_heads, -1).sum(dim=(0, 2))
            self.register_buffer("head_pruning_mask", self._topk_mask(per, ratio).reshape(1, -1))

    def enable_weight_quantization(self, start_bits, target_bits, quantization_period, weight_quantization_enabled_in_forward,
                                   quantization_type, num_groups):
        self.weight.start_bits, self.weight.target_bits, self.weight.q_period = start_bits, target_bits, quantization_period
        self.weight_quantization_enabled_in_forward = weight_quantization_enabled_in_forward
        self.weight_symmetric = quantization_type == "symmetric"
        self.weight_quantize_num_groups = num_groups

    def enable_activation_quantization(self, bits, quantization_type, range_calibration):
        assert bits in (4, 8), "Only 4/8 bits activation quantization are supported for now"
        self.activation_quantization_bits = bits
        self.activation_quantization_method = f"{quantization_type}_{range_calibration}"
        if range_calibration == "static":
            self.activation_quantizer = QuantAct(quant_mode=quantization_type)

    @staticmethod
    def _topk_mask(scores: torch.Tensor, dense_ratio: float) -> torch.Tensor:
        """Keep the largest ``dense_ratio`` fraction."""
        k = max(1, int(round(scores.numel() * dense_ratio)))
        flat = scores.reshape(-1)
        thresh = torch.topk(flat, k).values[-1]
        return (scores >= thresh).to(scores.dtype if scores.dtype.is_floating_point else torch.float32)

    # ---- masks -------------------------------------------------------------------------------------------
    def get_mask(self, pruning_type: str = "row"):
        if pruning_type == "sparse":
            if self.sparse_pruning_method == "l1":
                return self.sparse_pruning_mask.to(self.weight.device)
            return _TopKBinarizer.apply(self.sparse_mask_scores, self.sparse_pruning_ratio)
        if pruning_type == "row":
            if self.row_pruning_method == "l1":
                return self.row_pruning_mask.to(self.weight.device)
            return _TopKBinarizer.apply(self.row_mask_scores, self.row_pruning_ratio)
        if pruning_type == "head":
            if self.head_pruning_method == "topk":
                return _TopKBinarizer.apply(self.head_pruning_scores, self.head_pruning_ratio)
            return self.head_pruning_mask.to(self.weight.device)
        raise NotImplementedError(pruning_type)

    # ---- permanent application (redundancy_clean) ----------------------------------------------------------
    def fix_sparse_pruning_helper(self):
        self.weight.data = self.weight.data * self.get_mask("sparse").detach()
        self.sparse_pruning_enabled = False
        if self.sparse_pruning_method == "topk":
            del self.sparse_mask_scores
        self.sparse_pruning_method = None

    def fix_row_col_pruning_helper(self, mask=None, dim_reduction: bool = False):
        if mask is None:
            mask = self.get_mask("row").detach().bool().reshape(-1)
            if dim_reduction:
                self.weight = nn.Parameter(self.weight.data[mask])
                if self.bias is not None:
                    self.bias = nn.Parameter(self.bias.data[mask])
                self.out_features = self.weight.shape[0]
            else:
                self.weight.data = self.weight.data * mask.reshape(-1, 1)
                if self.bias is not None:
                    self.bias.data = self.bias.data * mask
            self.row_pruning_enabled, self.row_pruning_method = False, None
            return mask
        # a mask from the PREVIOUS layer: prune our input columns
        if dim_reduction:
            self.weight = nn.Parameter(self.weight.data[:, mask])
            self.in_features = self.weight.shape[1]
        else:
            self.weight.data = self.weight.data * mask.reshape(1, -1)
        return None

    def fix_head_pruning_helper(self, mask=None, num_heads=None, dim_reduction: bool = False):
        num_heads = num_heads or self.num_heads
        if mask is None:
            mask = self.get_mask("head").detach().bool().reshape(-1)
        col_mask = mask.repeat_interleave(self.in_features // num_heads)
        if dim_reduction:
            self.weight = nn.Parameter(self.weight.data[:, col_mask])
            self.in_features = self.weight.shape[1]
        else:
            self.weight.data = self.weight.data * col_mask.reshape(1, -1)
        self.head_pruning_enabled, self.head_pruning_method = False, None
        return mask

    def fix_weight_quantization(self):
        self.weight.data = fake_quantize(self.weight.data, self.weight.target_bits, self.weight_symmetric,
                                         self.weight_quantize_num_groups).data
        self.weight_quantization_enabled_in_forward = False
        return None

    # ---- forward ---------------------------------------------------------------------------------------------
    def forward(self, input, skip_bias_add: bool = False):
        w, b = self.weight, self.bias
        if self.weight_quantization_enabled_in_forward and self.weight_quantization_enabled:
            w = fake_quantize(w, self.weight.start_bits, self.weight_symmetric, self.weight_quantize_num_groups)
        if self.sparse_pruning_enabled and self.sparse_pruning_method:
            w = w * self.get_mask("sparse")
        if self.row_pruning_enabled and self.row_pruning_method:
            m = self.get_mask("row")
            w = w * m
            if b is not None:
                b = b * m.reshape(-1)
        if self.head_pruning_enabled and self.head_pruning_method:
            m = self.get_mask("head")
            w = (w.reshape(self.out_features, self.num_heads, -1) * m.reshape(1, -1, 1)).reshape(w.shape)
        if self.activation_quantization_enabled and self.activation_quantization_method:
            if "dynamic" in self.activation_quantization_method:
                groups = input.numel() // input.size(-1)
                input = fake_quantize(input, self.activation_quantization_bits,
                                      "symmetric" in self.activation_quantization_method, groups)
            else:
                input = self.activation_quantizer(input, self.activation_quantization_bits)
        if skip_bias_add:
            return F.linear(input, w, None), b
        return F.linear(input, w, b)


class _TopKBinarizer(torch.autograd.Function):
    """mask = 1 for the top ``ratio`` fraction of scores; gradient passes straight through to the scores."""

    @staticmethod
    def forward(ctx, scores, ratio):
        k = max(1, int(round(scores.numel() * ratio)))
        thresh = torch.topk(scores.reshape(-1), k).values[-1]
        return (scores >= thresh).to(scores.dtype)

    @staticmethod
    def backward(ctx, g):
        return g, None


class Embedding_Compress(nn.Embedding):
    def __init__(self, *kargs):
        super().__init__(*kargs)
        self.weight.start_bits = self.weight.target_bits = None
        self.weight.q_period = 0
        self.weight_quantization_enabled_in_forward = self.weight_quantization_enabled = False
        self.weight_quantize_num_groups, self.weight_symmetric = 1, True

    def enable_weight_quantization(self, start_bits, target_bits, quantization_period, weight_quantization_enabled_in_forward,
                                   quantization_type, num_groups):
        self.weight.start_bits, self.weight.target_bits, self.weight.q_period = start_bits, target_bits, quantization_period
        self.weight_quantization_enabled_in_forward = weight_quantization_enabled_in_forward
        self.weight_symmetric = quantization_type == "symmetric"
        self.weight_quantize_num_groups = self.weight.size(0)  # one group per token row

    def fix_weight_quantization(self):
        self.weight.data = fake_quantize(self.weight.data, self.weight.target_bits, self.weight_symmetric,
                                         self.weight_quantize_num_groups).data
        self.weight_quantization_enabled_in_forward = False

    def forward(self, input):
        w = self.weight
        if self.weight_quantization_enabled_in_forward and self.weight_quantization_enabled:
            w = fake_quantize(w, self.weight.start_bits, self.weight_symmetric, self.weight_quantize_num_groups)
        return F.embedding(input, w, self.padding_idx, self.max_norm, self.norm_type, self.scale_grad_by_freq, self.sparse)


class Conv2dLayer_Compress(nn.Conv2d):
    """Conv2d with sparse / channel pruning and weight quantisation (same switches as the linear layer)."""

    def __init__(self, *kargs):
        super().__init__(*kargs)
        self.sparse_pruning_method = self.channel_pruning_method = None
        self.sparse_pruning_enabled = self.channel_pruning_enabled = False
        self.weight.start_bits = self.weight.target_bits = None
        self.weight_quantization_enabled_in_forward = self.weight_quantization_enabled = False
        self.weight_quantize_num_groups, self.weight_symmetric = 1, True

    def enable_sparse_pruning(self, ratio, method):
        self.sparse_pruning_ratio, self.sparse_pruning_method = ratio, method
        self.register_buffer("sparse_pruning_mask", LinearLayer_Compress._topk_mask(self.weight.abs(), ratio))

    def enable_channel_pruning(self, ratio, method):
        self.channel_pruning_ratio, self.channel_pruning_method = ratio, method
        norms = self.weight.abs().sum(dim=(1, 2, 3))
        self.register_buffer("channel_pruning_mask", LinearLayer_Compress._topk_mask(norms, ratio).reshape(-1, 1, 1, 1))

    def enable_weight_quantization(self, start_bits, target_bits, quantization_period, in_forward, quantization_type, num_groups):
        self.weight.start_bits, self.weight.target_bits, self.weight.q_period = start_bits, target_bits, quantization_period
        self.weight_quantization_enabled_in_forward, self.weight_symmetric = in_forward, quantization_type == "symmetric"
        self.weight_quantize_num_groups = num_groups

    def fix_sparse_pruning_helper(self):
        self.weight.data = self.weight.data * self.sparse_pruning_mask
        self.sparse_pruning_enabled, self.sparse_pruning_method = False, None

    def fix_channel_pruning_helper(self, mask=None, dim_reduction=False):
        if mask is None:
            mask = self.channel_pruning_mask.bool().reshape(-1)
            if dim_reduction:
                self.weight = nn.Parameter(self.weight.data[mask])
                if self.bias is not None:
                    self.bias = nn.Parameter(self.bias.data[mask])
                self.out_channels = self.weight.shape[0]
            else:
                self.weight.data = self.weight.data * mask.reshape(-1, 1, 1, 1)
            self.channel_pruning_enabled, self.channel_pruning_method = False, None
            return mask
        if dim_reduction:
            self.weight = nn.Parameter(self.weight.data[:, mask])
            self.in_channels = self.weight.shape[1]
        else:
            self.weight.data = self.weight.data * mask.reshape(1, -1, 1, 1)
        return None

    def forward(self, input):
        w, b = self.weight, self.bias
        if self.weight_quantization_enabled_in_forward and self.weight_quantization_enabled:
            w = fake_quantize(w, self.weight.start_bits, self.weight_symmetric, self.weight_quantize_num_groups)
        if self.sparse_pruning_enabled and self.sparse_pruning_method:
            w = w * self.sparse_pruning_mask
        if self.channel_pruning_enabled and self.channel_pruning_method:
            w = w * self.channel_pruning_mask
            if b is not None:
                b = b * self.channel_pruning_mask.reshape(-1)
        return F.conv2d(input, w, b, self.stride, self.padding, self.dilation, self.groups)


class BNLayer_Compress(nn.BatchNorm2d):
    def fix_channel_pruning_helper(self, mask, dim_reduction: bool = True):
        self.weight = nn.Parameter(self.weight.data[mask])
        self.bias = nn.Parameter(self.bias.data[mask])
        self.running_mean, self.running_var = self.running_mean[mask], self.running_var[mask]
        self.num_features = int(mask.sum())


# ---- Megatron-style model-parallel variants (ref: basic_layer.py:680-840) -----------------------------------------------------
g_mpu = None


def split_tensor_along_last_dim(tensor, num_partitions: int, contiguous_split_chunks: bool = False):
    assert tensor.shape[-1] % num_partitions == 0, f"{tensor.shape[-1]} is not divisible by {num_partitions}"
    parts = torch.split(tensor, tensor.shape[-1] // num_partitions, dim=-1)
    return tuple(p.contiguous() for p in parts) if contiguous_split_chunks else parts


def _mp():
    from .. import comm as dist
    group = g_mpu.get_model_parallel_group()
    return dist, group, dist.get_world_size(group=group), dist.get_rank(group=group)


class _CopyToMP(torch.autograd.Function):
    """Identity forward, all-reduce of the gradient backward (input of a column-parallel layer)."""

    @staticmethod
    def forward(ctx, x):
        return x

    @staticmethod
    def backward(ctx, g):
        dist, group, ws, _ = _mp()
        if ws > 1:
            g = g.contiguous()
            dist.all_reduce(g, group=group)
        return g


class _ReduceFromMP(torch.autograd.Function):
    """All-reduce forward, identity backward (output of a row-parallel layer)."""

    @staticmethod
    def forward(ctx, x):
        dist, group, ws, _ = _mp()
        if ws > 1:
            x = x.contiguous().clone()
            dist.all_reduce(x, group=group)
        return x

    @staticmethod
    def backward(ctx, g):
        return g


class _ScatterToMP(torch.autograd.Function):
    """Keep this rank's slice of the last dimension forward, all-gather the gradient backward."""

    @staticmethod
    def forward(ctx, x):
        _, _, ws, rank = _mp()
        return x if ws == 1 else split_tensor_along_last_dim(x, ws)[rank].contiguous()

    @staticmethod
    def backward(ctx, g):
        return _gather_last(g)


class _GatherFromMP(torch.autograd.Function):
    """All-gather along the last dimension forward, keep this rank's slice of the gradient backward."""

    @staticmethod
    def forward(ctx, x):
        return _gather_last(x)

    @staticmethod
    def backward(ctx, g):
        _, _, ws, rank = _mp()
        return g if ws == 1 else split_tensor_along_last_dim(g, ws)[rank].contiguous()


def _gather_last(x):
    dist, group, ws, rank = _mp()
    if ws == 1:
        return x
    parts = [torch.empty_like(x) for _ in range(ws)]
    parts[rank] = x.contiguous()
    dist.all_gather(parts, parts[rank], group=group)
    return torch.cat(parts, dim=-1).contiguous()


def copy_to_model_parallel_region(input_):
    return _CopyToMP.apply(input_)


def reduce_from_model_parallel_region(input_):
    return _ReduceFromMP.apply(input_)


def scatter_to_model_parallel_region(input_):
    return _ScatterToMP.apply(input_)


def gather_from_model_parallel_region(input_):
    return _GatherFromMP.apply(input_)


class ColumnParallelLinear_Compress(LinearLayer_Compress):
    """Compressible linear whose OUTPUT features are split over the model-parallel group of ``mpu``; returns
    ``(output, bias or None)`` like Megatron's layer (``skip_bias_add`` hands the bias back unapplied)."""

    def __init__(self, mpu, input_size: int, output_size: int, bias: bool = True, gather_output: bool = True,
                 skip_bias_add: bool = False):
        global g_mpu
        g_mpu = mpu
        ws = mpu.get_model_parallel_world_size()
        assert output_size % ws == 0
        self.input_size, self.output_size, self.gather_output, self.skip_bias_add = input_size, output_size, gather_output, skip_bias_add
        self.output_size_per_partition = output_size // ws
        super().__init__(input_size, self.output_size_per_partition, bias=bias)

    def forward(self, input_):
        x = copy_to_model_parallel_region(input_)
        if self.skip_bias_add:
            out, bias = super().forward(x, True)
        else:
            out, bias = super().forward(x), None
        return (gather_from_model_parallel_region(out) if self.gather_output else out), bias


class RowParallelLinear_Compress(LinearLayer_Compress):
    """Compressible linear whose INPUT features are split over the model-parallel group; partial products are summed."""

    def __init__(self, mpu, input_size: int, output_size: int, bias: bool = True, input_is_parallel: bool = False,
                 skip_bias_add: bool = False):
        global g_mpu
        g_mpu = mpu
        ws = mpu.get_model_parallel_world_size()
        assert input_size % ws == 0
        self.input_size, self.output_size, self.input_is_parallel, self.skip_bias_add = input_size, output_size, input_is_parallel, skip_bias_add
        self.input_size_per_partition = input_size // ws
        super().__init__(self.input_size_per_partition, output_size, bias=bias)

    def forward(self, input_):
        x = input_ if self.input_is_parallel else scatter_to_model_parallel_region(input_)
        partial, bias = super().forward(x, True)
        out = reduce_from_model_parallel_region(partial)
        if not self.skip_bias_add:
            return (out + bias if bias is not None else out), None
        return out, bias
