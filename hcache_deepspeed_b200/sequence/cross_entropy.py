"""Cross entropy over logits sharded along the SEQUENCE dimension: every rank computes the loss of its own tokens and
the per-token losses are all-gathered (autograd-aware). Reference: ``deepspeed/sequence/cross_entropy.py:11-60``."""
from __future__ import annotations

import torch

from .. import comm as dist


class _VocabSequenceParallelCrossEntropy(torch.autograd.Function):

    @staticmethod
    def forward(ctx, vocab_seq_parallel_logits, target, sp_group):
        # logits: [S/P, B, V], target: [S/P, B]
        logits = vocab_seq_parallel_logits.float()
        m = logits.max(dim=-1, keepdim=True).values
        e = (logits - m).exp()
        s = e.sum(dim=-1, keepdim=True)
        softmax = e / s
        tgt = target.unsqueeze(-1)
        loss = (torch.log(s) + m - logits.gather(-1, tgt)).squeeze(-1)
        ctx.save_for_backward(softmax, target)
        ctx.sp_group = sp_group
        ws = dist.get_world_size(group=sp_group) if dist.is_initialized() else 1
        ctx.ws, ctx.rank = ws, (dist.get_rank(group=sp_group) if dist.is_initialized() else 0)
        ctx.seqlen = loss.shape[0]
        if ws == 1:
            return loss
        out = torch.empty((ws * loss.shape[0], ) + tuple(loss.shape[1:]), dtype=loss.dtype, device=loss.device)
        dist.all_gather_into_tensor(out, loss.contiguous(), group=sp_group)
        return out

    @staticmethod
    def backward(ctx, grad_output):
        softmax, target = ctx.saved_tensors
        g = grad_output[ctx.rank * ctx.seqlen:(ctx.rank + 1) * ctx.seqlen] if ctx.ws > 1 else grad_output
        grad = softmax.clone()
        grad.scatter_add_(-1, target.unsqueeze(-1), -torch.ones_like(target, dtype=grad.dtype).unsqueeze(-1))
        return grad * g.unsqueeze(-1), None, None


def vocab_sequence_parallel_cross_entropy(vocab_parallel_logits, target, sp_group):
    return _VocabSequenceParallelCrossEntropy.apply(vocab_parallel_logits, target, sp_group)
