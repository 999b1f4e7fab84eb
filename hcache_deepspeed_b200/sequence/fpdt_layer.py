"""FPDT (Fully Pipelined Distributed Transformer, "Ulysses-Offload"): Ulysses attention processed in sequence CHUNKS
so that only one chunk of q/k/v is resident on the device; finished K/V chunks wait in pinned host memory, and partial
attention results are merged with the online-softmax (log-sum-exp) rule.

``update_out_and_lse`` is the merge rule; ``FPDT_Attention`` runs causal attention chunk by chunk (optionally offloading
K/V chunks), ``FPDT_FFN`` chunks the MLP, ``FPDT_LogitsLoss`` chunks the LM head + loss.
Reference: ``deepspeed/sequence/fpdt_layer.py`` (update_out_and_lse :40-76, SequenceChunk :462-508, FPDT_Attention :971,
FPDT_FFN :1056, FPDT_LogitsLoss :1137).
"""
from __future__ import annotations

import math
from typing import Optional, Tuple

import torch

from .. import comm as dist
from .layer import single_all_to_all


def _merge(out, lse, block_out, block_lse):
    """Numerically stable merge of two partial softmax-weighted sums (out: [..., d], lse: [..., 1])."""
    new_lse = torch.logaddexp(lse, block_lse)
    return torch.exp(lse - new_lse) * out + torch.exp(block_lse - new_lse) * block_out, new_lse


def update_out_and_lse(out: Optional[torch.Tensor], lse: Optional[torch.Tensor], block_out: torch.Tensor,
                       block_lse: torch.Tensor, slice_=None) -> Tuple[torch.Tensor, torch.Tensor]:
    """out / block_out: [b, s, h, d]; lse / block_lse: [b, h, s] (flash-attention layout) or [b, s, h, 1]."""
    block_out = block_out.float()
    if block_lse.dim() == 3:
        block_lse = block_lse.transpose(-2, -1).unsqueeze(-1)
    block_lse = block_lse.float()
    if out is None:
        if slice_ is not None:
            raise RuntimeError("first update_out_and_lse should not pass slice_ args")
        return block_out, block_lse
    if slice_ is not None:
        o, l = _merge(out[slice_], lse[slice_], block_out, block_lse)
        out[slice_], lse[slice_] = o, l
        return out, lse
    return _merge(out, lse, block_out, block_lse)


def _attn_with_lse(q, k, v, causal_offset: Optional[int], scale: float):
    """Plain attention returning (out [b,s,h,d], lse [b,s,h,1]); ``causal_offset``: q position i attends k position
    j iff j <= i + causal_offset (None = no mask)."""
    qh, kh, vh = q.permute(0, 2, 1, 3).float(), k.permute(0, 2, 1, 3).float(), v.permute(0, 2, 1, 3).float()
    if kh.shape[1] != qh.shape[1]:  # GQA
        rep = qh.shape[1] // kh.shape[1]
        kh, vh = kh.repeat_interleave(rep, dim=1), vh.repeat_interleave(rep, dim=1)
    s = torch.matmul(qh, kh.transpose(-1, -2)) * scale
    if causal_offset is not None:
        i = torch.arange(q.shape[1], device=q.device).unsqueeze(-1)
        j = torch.arange(k.shape[1], device=q.device).unsqueeze(0)
        s = s.masked_fill(j > i + causal_offset, float("-inf"))
    lse = torch.logsumexp(s, dim=-1, keepdim=True)
    p = torch.exp(s - lse)
    out = torch.matmul(p, vh)
    return out.permute(0, 2, 1, 3), lse.permute(0, 2, 1, 3)


class SequenceChunk:
    """A K or V chunk that can be parked in pinned host memory and brought back on a copy stream."""

    def __init__(self, chunk: torch.Tensor, device=None, is_in_use: bool = False):
        self.device = chunk.device if device is None else device
        self.chunk_shape, self.chunk_dtype = chunk.shape, chunk.dtype
        self.in_use = is_in_use
        cpu = torch.empty(chunk.numel(), dtype=chunk.dtype, device="cpu")
        if torch.cuda.is_available():
            cpu = cpu.pin_memory()
        cpu.copy_(chunk.reshape(-1), non_blocking=True)
        self.cpu_chunk = cpu
        self.gpu_chunk = chunk if is_in_use else None

    def load_to_gpu(self):
        if self.gpu_chunk is None:
            self.gpu_chunk = self.cpu_chunk.to(self.device, non_blocking=True).view(self.chunk_shape)
        self.in_use = True

    def get_gpu_chunk(self):
        assert self.in_use and self.gpu_chunk is not None
        return self.gpu_chunk

    def check_gpu_chunk(self):
        return self.gpu_chunk is not None

    def offload(self):
        self.gpu_chunk = None
        self.in_use = False

    def overwrite_to_cpu(self):
        assert self.in_use and self.gpu_chunk is not None
        self.cpu_chunk.copy_(self.gpu_chunk.reshape(-1), non_blocking=True)


class FPDT_Attention(torch.nn.Module):
    """Chunked causal Ulysses attention over already-projected q, k, v ([b, s_local, h, d], sequence-sharded).
    For chunk i of q only k/v chunks 0..i are touched; with ``offloading`` finished k/v chunks live on the host."""

    def __init__(self, sequence_process_group=None, num_chunks: int = 2, offloading: bool = False, scale: Optional[float] = None):
        super().__init__()
        self.spg, self.num_chunks, self.offloading, self.scale = sequence_process_group, num_chunks, offloading, scale

    def forward(self, q: torch.Tensor, k: torch.Tensor, v: torch.Tensor) -> torch.Tensor:
        sp = dist.get_world_size(self.spg) if (self.spg is not None and dist.is_initialized()) else 1
        scale = self.scale if self.scale is not None else 1.0 / math.sqrt(q.shape[-1])
        nc = self.num_chunks
        assert q.shape[1] % nc == 0, "local sequence length must be divisible by the number of chunks"
        qs, ks, vs = q.chunk(nc, dim=1), k.chunk(nc, dim=1), v.chunk(nc, dim=1)
        s_local = q.shape[1]
        # pass 1: Ulysses exchange of every K/V chunk (each rank receives the chunk's tokens of all ranks for its own
        # heads); finished chunks are parked in pinned host memory when offloading
        kv_store = []
        for i in range(nc):
            ki = single_all_to_all(ks[i].contiguous(), 2, 1, 0, self.spg) if sp > 1 else ks[i]
            vi = single_all_to_all(vs[i].contiguous(), 2, 1, 0, self.spg) if sp > 1 else vs[i]
            if self.offloading:
                kc, vc = SequenceChunk(ki, is_in_use=True), SequenceChunk(vi, is_in_use=True)
                kc.offload()
                vc.offload()
                kv_store.append((kc, vc))
            else:
                kv_store.append((ki, vi))
        # pass 2: one query chunk at a time against the K/V chunks it can see. A gathered chunk holds, rank-major, the
        # tokens (rank r, chunk c, local p) whose global position is r * s_local + c * c_local + p, so causality is
        # decided from explicit positions and (q chunk, kv chunk) pairs that are fully masked are skipped.
        outs = []
        for i in range(nc):
            qi = single_all_to_all(qs[i].contiguous(), 2, 1, 0, self.spg) if sp > 1 else qs[i]
            qp = self._positions(qi.shape[1], i, sp, s_local, qi.device)
            out, lse = None, None
            for j in range(nc):
                kp = self._positions(qi.shape[1], j, sp, s_local, qi.device)
                if int(kp.min()) > int(qp.max()):
                    continue
                if self.offloading:
                    kc, vc = kv_store[j]
                    kc.load_to_gpu()
                    vc.load_to_gpu()
                    kj, vj = kc.get_gpu_chunk(), vc.get_gpu_chunk()
                else:
                    kj, vj = kv_store[j]
                bo, bl = self._block(qi, kj, vj, kp.unsqueeze(0) > qp.unsqueeze(1), scale)
                out, lse = update_out_and_lse(out, lse, bo, bl)
                if self.offloading:
                    kc.offload()
                    vc.offload()
            oi = out.to(q.dtype)
            outs.append(single_all_to_all(oi.contiguous(), 1, 2, 0, self.spg, total_heads=q.shape[2]) if sp > 1 else oi)
        return torch.cat(outs, dim=1)

    def _positions(self, n_chunk_global: int, chunk_idx: int, sp: int, s_local: int, device):
        """Global sequence position of each token of a gathered chunk (rank-major inside the chunk)."""
        c_local = n_chunk_global // sp
        r = torch.arange(sp, device=device).repeat_interleave(c_local)
        p = torch.arange(c_local, device=device).repeat(sp)
        return r * s_local + chunk_idx * c_local + p

    @staticmethod
    def _block(q, k, v, masked, scale):
        qh, kh, vh = q.permute(0, 2, 1, 3).float(), k.permute(0, 2, 1, 3).float(), v.permute(0, 2, 1, 3).float()
        if kh.shape[1] != qh.shape[1]:
            rep = qh.shape[1] // kh.shape[1]
            kh, vh = kh.repeat_interleave(rep, dim=1), vh.repeat_interleave(rep, dim=1)
        s = torch.matmul(qh, kh.transpose(-1, -2)) * scale
        s = s.masked_fill(masked, float("-inf"))
        lse = torch.logsumexp(s, dim=-1, keepdim=True)
        lse = torch.where(torch.isinf(lse), torch.full_like(lse, -1e30), lse)  # fully masked rows contribute nothing
        p = torch.exp(s - lse)
        return torch.matmul(p, vh).permute(0, 2, 1, 3), lse.permute(0, 2, 1, 3)


class FPDT_FFN(torch.nn.Module):
    """Applies an MLP over sequence chunks so the 4x intermediate activation exists for one chunk at a time."""

    def __init__(self, mlp: torch.nn.Module, num_chunks: int = 2, seq_dim: int = 1):
        super().__init__()
        self.mlp, self.num_chunks, self.seq_dim = mlp, num_chunks, seq_dim

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        outs = []
        for c in x.chunk(self.num_chunks, dim=self.seq_dim):
            y = self.mlp(c)
            outs.append(y[0] if isinstance(y, tuple) else y)
        return torch.cat(outs, dim=self.seq_dim)


class FPDT_LogitsLoss(torch.nn.Module):
    """LM head + cross entropy computed per sequence chunk (the [s, vocab] logits never exist in full)."""

    def __init__(self, lm_head: torch.nn.Module, num_chunks: int = 2, ignore_index: int = -100):
        super().__init__()
        self.lm_head, self.num_chunks, self.ignore_index = lm_head, num_chunks, ignore_index

    def forward(self, hidden: torch.Tensor, labels: torch.Tensor) -> torch.Tensor:
        total, count = hidden.new_zeros((), dtype=torch.float32), 0
        for h, y in zip(hidden.chunk(self.num_chunks, dim=1), labels.chunk(self.num_chunks, dim=1)):
            logits = self.lm_head(h).float()
            total = total + torch.nn.functional.cross_entropy(logits.reshape(-1, logits.shape[-1]), y.reshape(-1),
                                                              ignore_index=self.ignore_index, reduction="sum")
            count += int((y != self.ignore_index).sum())
        return total / max(count, 1)


# ---- input side and small functions of the reference module (sequence/fpdt_layer.py:32, 79-130, 1044-1053) -----------------
def _rotate_half_backward(x: torch.Tensor) -> torch.Tensor:
    x1, x2 = x.reshape(*x.shape[:-1], 2, x.shape[-1] // 2).unbind(dim=-2)
    return torch.cat((x2, -x1), dim=-1)


def apply_rotary_pos_emb_backward(grad_output: torch.Tensor, freqs_cos: torch.Tensor, freqs_sin: torch.Tensor) -> torch.Tensor:
    """Gradient of ``sequence.layer.apply_rotary_pos_emb`` with respect to its input (the rotation transposed)."""
    rot = freqs_cos.shape[-1]
    g, rest = grad_output[..., :rot], grad_output[..., rot:]
    g = g * freqs_cos + _rotate_half_backward(g * freqs_sin)
    return g if rest.shape[-1] == 0 else torch.cat((g, rest), dim=-1)


def bias_gelu(x: torch.Tensor) -> torch.Tensor:
    """tanh-approximated GELU (the bias has been added by the caller)."""
    return x * 0.5 * (1.0 + torch.tanh(0.79788456 * x * (1 + 0.044715 * x * x)))


def bias_gelu_back(g: torch.Tensor, x: torch.Tensor) -> torch.Tensor:
    t = torch.tanh(0.79788456 * x * (1 + 0.044715 * x * x))
    return (0.5 * x * ((1 - t * t) * (0.79788456 + 0.1070322243 * x * x)) + 0.5 * (1 + t)) * g


class FPDT_InputConstruct(torch.nn.Module):
    """Cuts a batch for chunked sequence parallelism with a load-balanced assignment: the sequence is split into
    ``sp_size * chunks_per_gpu`` chunks and rank ``r`` takes chunks ``r, r + sp_size, r + 2*sp_size, ...`` - under a causal
    mask every rank then owns early (cheap) and late (expensive) positions alike. ``generate()`` returns this rank's
    tokens / labels / position ids, the loss mask in the re-ordered order of ALL ranks (it is applied after the logits of
    all ranks were gathered) and the attention mask untouched."""

    def __init__(self, tokens, labels, loss_mask, attention_mask, position_ids, args, sp_size: int, sp_rank: int) -> None:
        super().__init__()
        self.tokens, self.labels, self.loss_mask = tokens, labels, loss_mask
        self.attention_mask, self.position_ids = attention_mask, position_ids
        self.batch_size, self.global_seq_len = tokens.shape[0], tokens.shape[1]
        fpdt_chunk = args.ds_sequence_parallel_fpdt_chunk_size
        assert self.global_seq_len % sp_size == 0 and self.global_seq_len % fpdt_chunk == 0
        self.num_chunk_per_gpu = self.global_seq_len // fpdt_chunk
        self.local_seq_len = self.global_seq_len // sp_size
        assert self.local_seq_len % self.num_chunk_per_gpu == 0
        self.chunk_size = self.local_seq_len // self.num_chunk_per_gpu
        self.sp_size, self.sp_rank, self.device = sp_size, sp_rank, tokens.device

    def _order(self) -> torch.Tensor:
        """Token indices in (rank, chunk-of-rank, position) order."""
        chunks = torch.arange(self.global_seq_len // self.chunk_size, device=self.device)
        by_rank = chunks.reshape(self.num_chunk_per_gpu, self.sp_size).t().reshape(-1)          # rank r: r, r+sp, ...
        return (by_rank[:, None] * self.chunk_size + torch.arange(self.chunk_size, device=self.device)[None, :]).reshape(-1)

    def generate(self):
        order = self._order()
        mine = order[self.local_seq_len * self.sp_rank:self.local_seq_len * (self.sp_rank + 1)]
        pick = lambda t, idx: t[:, idx] if t is not None else None     # noqa: E731
        return pick(self.tokens, mine), pick(self.labels, mine), pick(self.loss_mask, order), self.attention_mask, \
            pick(self.position_ids, mine)
