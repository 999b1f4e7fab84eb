"""TiledLinear: a Linear split into ``in_splits x out_splits`` smaller Linears so ZeRO-3 only ever has one tile's
parameters gathered at a time. Reference: ``deepspeed/runtime/zero/tiling.py:26-296``."""
from __future__ import annotations

import torch

from ..utils import partition_uniform


def split_tensor_along_last_dim(tensor, partitions, contiguous_split_chunks=False):
    sizes = [partitions[i + 1] - partitions[i] for i in range(len(partitions) - 1)]
    chunks = torch.split(tensor, sizes, dim=tensor.dim() - 1)
    return tuple(c.contiguous() for c in chunks) if contiguous_split_chunks else chunks


class TiledLinear(torch.nn.Module):

    def __init__(self, in_features, out_features, bias=True, in_splits=1, out_splits=1, input_is_already_split=False,
                 combine_out_splits=True, linear_cls=torch.nn.Linear, init_linear=None, **kwargs):
        super().__init__()
        if not 1 <= in_splits <= in_features or not 1 <= out_splits <= out_features:
            raise RuntimeError("splits must be in [1, features]")
        self.in_features, self.out_features, self.use_bias = in_features, out_features, bias
        self.in_splits, self.out_splits = in_splits, out_splits
        self.input_is_already_split, self.combine_out_splits = input_is_already_split, combine_out_splits
        self.in_parts = partition_uniform(in_features, in_splits)
        self.out_parts = partition_uniform(out_features, out_splits)
        self.linears = torch.nn.ModuleList()
        for o in range(out_splits):
            row = torch.nn.ModuleList()
            for i in range(in_splits):
                # only the last column tile carries the bias so it is added exactly once
                row.append(linear_cls(self.in_parts[i + 1] - self.in_parts[i], self.out_parts[o + 1] - self.out_parts[o],
                                      bias=bias and i == in_splits - 1, **kwargs))
            self.linears.append(row)
        if init_linear is not None:
            self.copy_params_from(init_linear)

    def forward(self, input_):
        if self.in_splits > 1 and not self.input_is_already_split:
            inputs = split_tensor_along_last_dim(input_, self.in_parts)
        elif self.in_splits > 1:
            inputs = input_
        else:
            inputs = [input_]
        outs = []
        for o in range(self.out_splits):
            acc = None
            for i in range(self.in_splits):
                y = self.linears[o][i](inputs[i])
                acc = y if acc is None else self._reduce_local_output(i, o, acc, y)
            outs.append(acc)
        return self._combine_output_splits(outs) if self.combine_out_splits else outs

    def _reduce_local_output(self, in_id, out_id, current_out, new_out):
        return current_out + new_out

    def _combine_output_splits(self, outputs):
        return torch.cat(outputs, dim=-1)

    @torch.no_grad()
    def copy_params_from(self, other: torch.nn.Linear):
        assert other.in_features == self.in_features and other.out_features == self.out_features
        for o in range(self.out_splits):
            rs, re = self.out_parts[o], self.out_parts[o + 1]
            for i in range(self.in_splits):
                cs, ce = self.in_parts[i], self.in_parts[i + 1]
                self.linears[o][i].weight.copy_(other.weight[rs:re, cs:ce])
                if self.linears[o][i].bias is not None:
                    self.linears[o][i].bias.copy_(other.bias[rs:re])


class TiledLinearReturnBias(TiledLinear):
    """For Megatron-style linears whose forward returns ``(output, bias)``."""

    def _reduce_local_output(self, in_id, out_id, current_out, new_out):
        t, b = current_out if isinstance(current_out, tuple) else (current_out, None)
        nt, nb = new_out if isinstance(new_out, tuple) else (new_out, None)
        return t + nt, (nb if nb is not None else b)

    def _combine_output_splits(self, outputs):
        ts = [o[0] if isinstance(o, tuple) else o for o in outputs]
        bs = [o[1] for o in outputs if isinstance(o, tuple) and o[1] is not None]
        return torch.cat(ts, dim=-1), (torch.cat(bs) if bs else None)
