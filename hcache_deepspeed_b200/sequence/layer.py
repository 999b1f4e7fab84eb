"""DeepSpeed-Ulysses. Activations are sharded along the sequence between attention blocks; around the local attention
an all-to-all re-shards them along heads: each rank then sees the FULL sequence for ``heads / sp`` heads, runs any
attention implementation unchanged, and a second all-to-all restores sequence sharding. 4 ``all_to_all_single`` per
attention (q, k, v in; context out), each moving ``1/sp`` of the activation — constant per-rank volume as ``sp`` grows.
Heads that do not divide by ``sp`` are handled by an uneven-split all-to-all.
Reference: ``deepspeed/sequence/layer.py`` (single_all_to_all :221, _SeqAllToAll :257, DistributedAttention :311,
uneven heads :111-218).
"""
from __future__ import annotations

from typing import Any, Optional, Tuple

import torch
from torch import Tensor

from .. import comm as dist


def _generate_layout_params(scatter_idx: int, batch_dim_idx: int, seq_world_size: int, x: Tensor):
    if batch_dim_idx == 0:
        if scatter_idx < 2:
            bs, global_seq, nh_local, hd = x.shape
            pre_shape = [bs, seq_world_size, global_seq // seq_world_size, nh_local, hd]
            pre_perm = (1, 0, 2, 3, 4)
            post_perm = (1, 2, 0, 3, 4)
            post_shape = [bs, global_seq // seq_world_size, seq_world_size * nh_local, hd]
        else:
            bs, local_seq, nh, hd = x.shape
            pre_shape = [bs, local_seq, seq_world_size, nh // seq_world_size, hd]
            pre_perm = (2, 0, 1, 3, 4)
            post_perm = (1, 0, 2, 3, 4)
            post_shape = [bs, seq_world_size * local_seq, nh // seq_world_size, hd]
    else:
        if scatter_idx < 2:
            global_seq, bs, nh_local, hd = x.shape
            pre_shape = [seq_world_size, global_seq // seq_world_size, bs, nh_local, hd]
            pre_perm = None
            post_perm = (1, 2, 0, 3, 4)
            post_shape = [global_seq // seq_world_size, bs, seq_world_size * nh_local, hd]
        else:
            local_seq, bs, nh, hd = x.shape
            pre_shape = [local_seq, bs, seq_world_size, nh // seq_world_size, hd]
            pre_perm = (2, 0, 1, 3, 4)
            post_perm = None
            post_shape = [seq_world_size * local_seq, bs, nh // seq_world_size, hd]
    return pre_shape, pre_perm, post_perm, post_shape


def single_all_to_all(x: Tensor, scatter_idx: int, gather_idx: int, batch_dim_idx: int, group, async_op: bool = False,
                      handle=None, type=None, total_heads: Optional[int] = None) -> Tensor:
    """[b, s/p, h, d] -> [b, s, h/p, d] (scatter_idx=2) or the inverse (scatter_idx<2); sequence-first layouts with
    ``batch_dim_idx=1``. ``total_heads`` (needed on the inverse path) selects the uneven-split exchange when the head
    count does not divide by the group size."""
    ws = dist.get_world_size(group)
    if ws == 1:
        return x
    if scatter_idx >= 2 and total_heads is None:
        total_heads = x.shape[2]
    if total_heads is not None and total_heads % ws != 0:
        return uneven_heads_all2all(x, scatter_idx, gather_idx, batch_dim_idx, group, total_heads)
    pre_shape, pre_perm, post_perm, post_shape = _generate_layout_params(scatter_idx, batch_dim_idx, ws, x)
    t = x.reshape(pre_shape)
    if pre_perm is not None:
        t = t.permute(*pre_perm)
    t = t.contiguous()
    out = torch.empty_like(t)
    dist.all_to_all_single(out, t, group=group)
    if post_perm is not None:
        out = out.permute(*post_perm)
    return out.contiguous().reshape(post_shape)


def _head_splits(num_heads: int, ws: int):
    base, rem = divmod(num_heads, ws)
    return [base + (1 if r < rem else 0) for r in range(ws)]


def uneven_heads_all2all(x: Tensor, scatter_idx: int, gather_idx: int, batch_dim_idx: int, group,
                         total_heads: Optional[int] = None) -> Tensor:
    """All-to-all when ``heads % sp != 0``: rank r receives ``base + (r < rem)`` heads. Batch-first layout."""
    ws, rank = dist.get_world_size(group), dist.get_rank(group)
    assert batch_dim_idx == 0, "uneven-head all-to-all supports the [batch, seq, heads, dim] layout"
    if scatter_idx >= 2:  # [b, s/p, H, d] -> [b, s, h_r, d]
        b, s_local, H, d = x.shape
        splits = _head_splits(H, ws)
        send = list(torch.split(x, splits, dim=2))
        send = [t.contiguous() for t in send]
        recv = [torch.empty(b, s_local, splits[rank], d, dtype=x.dtype, device=x.device) for _ in range(ws)]
        dist.all_to_all(recv, send, group=group)
        return torch.cat(recv, dim=1).contiguous()
    b, s, h_local, d = x.shape  # inverse: [b, s, h_r, d] -> [b, s/p, H, d]
    splits = _head_splits(total_heads, ws)
    assert splits[rank] == h_local
    send = [t.contiguous() for t in torch.chunk(x, ws, dim=1)]
    recv = [torch.empty(b, s // ws, splits[r], d, dtype=x.dtype, device=x.device) for r in range(ws)]
    dist.all_to_all(recv, send, group=group)
    return torch.cat(recv, dim=2).contiguous()


class _SeqAllToAll(torch.autograd.Function):
    @staticmethod
    def forward(ctx: Any, group, x: Tensor, scatter_idx: int, gather_idx: int, batch_dim_idx: int = 0,
                total_heads: Optional[int] = None) -> Tensor:
        if scatter_idx >= 2 and total_heads is None:
            total_heads = x.shape[2]
        ctx.group, ctx.scatter_idx, ctx.gather_idx, ctx.batch_dim_idx = group, scatter_idx, gather_idx, batch_dim_idx
        ctx.total_heads = total_heads
        return single_all_to_all(x, scatter_idx, gather_idx, batch_dim_idx, group, total_heads=total_heads)

    @staticmethod
    def backward(ctx: Any, *grad_output: Tensor) -> Tuple:
        return (None, _SeqAllToAll.apply(ctx.group, grad_output[0], ctx.gather_idx, ctx.scatter_idx, ctx.batch_dim_idx,
                                         ctx.total_heads), None, None, None, None)


class DistributedAttention(torch.nn.Module):
    """Wraps a local attention callable ``attn(q, k, v, *args)`` that expects full-sequence tensors."""

    def __init__(self, local_attention, sequence_process_group, scatter_idx: int = 2, gather_idx: int = 0, sp_stream=None):
        super().__init__()
        self.local_attn = local_attention
        self.spg = sequence_process_group
        self.scatter_idx, self.gather_idx = scatter_idx, gather_idx
        self.sp_overlap_comm = sp_stream is not None
        self.sp_stream = sp_stream

    def layer_sync(self, layer):
        if self.sp_overlap_comm and hasattr(layer, "done_event"):
            torch.cuda.current_stream().wait_event(layer.done_event)

    def forward(self, query: Tensor, key: Tensor, value: Tensor, batch_dim_idx: int = 0, *args: Any, **kwargs) -> Tensor:
        q = _SeqAllToAll.apply(self.spg, query, self.scatter_idx, self.gather_idx, batch_dim_idx)
        k = _SeqAllToAll.apply(self.spg, key, self.scatter_idx, self.gather_idx, batch_dim_idx)
        v = _SeqAllToAll.apply(self.spg, value, self.scatter_idx, self.gather_idx, batch_dim_idx)
        ctx = self.local_attn(q, k, v, *args, **kwargs)
        return _SeqAllToAll.apply(self.spg, ctx, self.gather_idx, self.scatter_idx, batch_dim_idx, query.shape[2])


# ---- small helpers the reference exports from this module (sequence/layer.py:59-108) ---------------------------------------
def pre_all2all_fun(permute_idx, inp_shape, input: Tensor) -> Tensor:
    """View ``input`` as ``inp_shape`` and (optionally) permute it into the layout the all-to-all sends."""
    x = input.reshape(inp_shape).contiguous()
    return x.permute(permute_idx).contiguous() if permute_idx is not None else x


def post_all2all(permute_idx, res_shape):
    """The matching receive-side step: returns a function that permutes (optionally) and reshapes to ``res_shape``."""

    def post_func(input: Tensor) -> Tensor:
        x = input.permute(permute_idx).contiguous() if permute_idx is not None else input
        return x.reshape(res_shape).contiguous()

    return post_func


def _rotate_half(x: Tensor) -> Tensor:
    x1, x2 = x.reshape(*x.shape[:-1], 2, x.shape[-1] // 2).unbind(dim=-2)
    return torch.cat((-x2, x1), dim=-1)


def apply_rotary_pos_emb(t: Tensor, freqs_cos: Tensor, freqs_sin: Tensor) -> Tensor:
    """Half-split rotary embedding on ``t`` of shape ``[seq, ..., dim]`` with per-position ``cos`` / ``sin`` tables of
    shape ``[seq, ..., rot_dim]``; features beyond ``rot_dim`` pass through."""
    rot = freqs_cos.shape[-1]
    head, rest = t[..., :rot], t[..., rot:]
    head = head * freqs_cos + _rotate_half(head) * freqs_sin
    return head if rest.shape[-1] == 0 else torch.cat((head, rest), dim=-1)
