"""NHWC bias-add variants used by the diffusers injection (``csrc/spatial/csrc/opt_bias_add.cu``): bias add, bias add +
residual add, bias add + (other + other_bias)."""
import torch


def nhwc_bias_add(activation: torch.Tensor, bias: torch.Tensor, other: torch.Tensor = None, other_bias: torch.Tensor = None):
    """``activation``: channels-last [N, C, H, W] (memory NHWC); ``bias``: [C]."""
    out = activation + bias.view(1, -1, 1, 1)
    if other is not None:
        out = out + (other if other_bias is None else other + other_bias.view(1, -1, 1, 1))
    return out.contiguous(memory_format=torch.channels_last)
