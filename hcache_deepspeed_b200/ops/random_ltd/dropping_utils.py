"""Token sampling / gather / scatter for random layer-wise token dropping with explicit autograd
(``deepspeed/ops/random_ltd/dropping_utils.py:18-135``; kernels ``csrc/random_ltd/*.cu``). Index math is torch:
these ops run once per layer on ``[batch, kept]`` indices and are not on any hot path of this framework."""
from __future__ import annotations

from typing import List, Optional, Tuple

import torch

from ...runtime.data_pipeline.random_ltd import gather_tokens, scatter_tokens, token_sort_


def gpt_sample_tokens(reserved_length: int, seq_length: int, batch_size: int, layers: int = 1, device="cpu",
                      attn_mask: Optional[torch.Tensor] = None) -> Tuple[torch.Tensor, Optional[torch.Tensor]]:
    """Sorted random token subsets ``[layers, batch, reserved]`` and the causal mask cropped to ``reserved``."""
    idx = token_sort_(reserved_length, seq_length, batch_size, layers, device=device).to(torch.int32)
    mask = attn_mask[:, :, :reserved_length, :reserved_length] if attn_mask is not None else None
    return idx, mask


def bert_sample_tokens(reserved_length: int, seq_length: int, batch_size: int, layers: int = 1, device="cpu",
                       attn_mask: Optional[torch.Tensor] = None) -> Tuple[torch.Tensor, List[torch.Tensor]]:
    """As above, but a bidirectional mask must be gathered at the sampled rows AND columns, per layer."""
    assert attn_mask is not None
    idx = token_sort_(reserved_length, seq_length, batch_size, layers, device=device)
    masks = []
    for l in range(layers):
        rows = idx[l].long()                                            # [b, r]
        m = attn_mask.gather(2, rows[:, None, :, None].expand(-1, attn_mask.shape[1], -1, attn_mask.shape[3]))
        masks.append(m.gather(3, rows[:, None, None, :].expand(-1, attn_mask.shape[1], reserved_length, -1)))
    return idx.to(torch.int32), masks


class GatherTokens(torch.autograd.Function):
    """``(activations, gathered)``; the gradient of the gathered part is scattered back into the full gradient."""

    @staticmethod
    def forward(ctx, activations: torch.Tensor, sorted_indices: torch.Tensor, batch_first: bool):
        ctx.save_for_backward(sorted_indices)
        ctx.batch_first = batch_first
        return activations, gather_tokens(activations, sorted_indices.long(), batch_first)

    @staticmethod
    def backward(ctx, a_gradients: torch.Tensor, g_gradients: torch.Tensor):
        (idx, ) = ctx.saved_tensors
        return scatter_tokens(a_gradients, g_gradients.contiguous(), idx.long(), ctx.batch_first), None, None


class ScatterTokens(torch.autograd.Function):
    """Write the layer's outputs for the kept tokens back into the full sequence."""

    @staticmethod
    def forward(ctx, all_activations: torch.Tensor, layer_activations: torch.Tensor, sorted_indices: torch.Tensor,
                batch_first: bool):
        ctx.save_for_backward(sorted_indices)
        ctx.batch_first = batch_first
        return scatter_tokens(all_activations, layer_activations, sorted_indices.long(), batch_first)

    @staticmethod
    def backward(ctx, out_gradients: torch.Tensor):
        (idx, ) = ctx.saved_tensors
        out_gradients = out_gradients.contiguous()
        part = gather_tokens(out_gradients, idx.long(), ctx.batch_first)
        # the kept positions of ``all_activations`` were overwritten, so no gradient reaches them through this input
        # (the reference passes ``out_gradients`` through unmasked, which double-counts the kept tokens)
        return scatter_tokens(out_gradients, torch.zeros_like(part), idx.long(), ctx.batch_first), part, None, None
