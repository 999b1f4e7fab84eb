"""1-bit compressed all-reduce with error feedback (the communication primitive of 1-bit Adam / LAMB / 0/1 Adam).

Each rank sends only the SIGN of (gradient momentum + its residual error) plus one scale per chunk:
  1. worker compression: ``c = sign(x + e_w) * mean|x + e_w|``, ``e_w <- x + e_w - c``
  2. all-to-all of the sign bits (packed 8 per byte) + all-gather of the scales: rank r receives chunk r from everyone
  3. server averages its chunk, compresses again with its own error ``e_s``
  4. all-gather of the compressed server chunks.
One implementation over :mod:`hcache_deepspeed_b200.comm` serves NCCL (GPUs) and gloo (CPU tests); the reference has
one back end per transport (``deepspeed/runtime/comm/{nccl,mpi,hccl,compressed}.py``)."""
from __future__ import annotations

import torch

from ... import comm as dist


def pack_signs(x: torch.Tensor) -> torch.Tensor:
    """bool/sign tensor (numel % 8 == 0) -> uint8 bits."""
    b = (x >= 0).to(torch.uint8).reshape(-1, 8)
    w = torch.tensor([1, 2, 4, 8, 16, 32, 64, 128], dtype=torch.uint8, device=x.device)
    return (b * w).sum(dim=1).to(torch.uint8)


def unpack_signs(p: torch.Tensor) -> torch.Tensor:
    """uint8 bits -> float tensor of +1 / -1."""
    w = torch.tensor([1, 2, 4, 8, 16, 32, 64, 128], dtype=torch.uint8, device=p.device)
    bits = (p.reshape(-1, 1) & w) != 0
    return bits.reshape(-1).float() * 2 - 1


class CompressedBackend:

    def __init__(self, mpu=None):
        self.mpu = mpu
        self.world_group = mpu.get_data_parallel_group() if mpu is not None else None
        self.size = dist.get_world_size(group=self.world_group) if dist.is_initialized() else 1
        self.rank = dist.get_rank(group=self.world_group) if dist.is_initialized() else 0

    def padded_size(self, n: int) -> int:
        q = 8 * self.size
        return (n + q - 1) // q * q

    def compressed_allreduce(self, buffer_m: torch.Tensor, worker_error: torch.Tensor, server_error: torch.Tensor,
                             local_rank: int = 0) -> torch.Tensor:
        """In place: ``buffer_m`` <- approx mean over ranks. ``worker_error`` has ``padded_size(numel)`` elements,
        ``server_error`` has ``padded_size / world`` elements; both persist across calls."""
        shape, n = buffer_m.shape, buffer_m.numel()
        P = worker_error.numel()
        ws = self.size
        x = torch.zeros(P, dtype=torch.float32, device=buffer_m.device)
        x[:n] = buffer_m.reshape(-1).float()
        x += worker_error
        scale = x.norm() / (P ** 0.5)
        signs = torch.where(x >= 0, torch.ones_like(x), -torch.ones_like(x))
        worker_error.copy_(x - scale * signs)
        if ws == 1:
            buffer_m.copy_((scale * signs)[:n].reshape(shape).to(buffer_m.dtype))
            return buffer_m
        packed = pack_signs(signs)                       # [P/8] bytes, chunk r = bytes [r*P/8/ws, ...)
        recv = torch.empty_like(packed)
        dist.all_to_all_single(recv, packed, group=self.world_group)
        scales = [torch.zeros_like(scale) for _ in range(ws)]
        dist.all_gather(scales, scale, group=self.world_group)
        chunk = P // ws
        acc = torch.zeros(chunk, dtype=torch.float32, device=x.device)
        per = recv.numel() // ws
        for r in range(ws):
            acc += unpack_signs(recv[r * per:(r + 1) * per]) * scales[r]
        acc /= ws
        acc += server_error
        s_scale = acc.norm() / (chunk ** 0.5)
        s_signs = torch.where(acc >= 0, torch.ones_like(acc), -torch.ones_like(acc))
        server_error.copy_(acc - s_scale * s_signs)
        s_packed = pack_signs(s_signs)
        all_packed = [torch.empty_like(s_packed) for _ in range(ws)]
        dist.all_gather(all_packed, s_packed, group=self.world_group)
        s_scales = [torch.zeros_like(s_scale) for _ in range(ws)]
        dist.all_gather(s_scales, s_scale, group=self.world_group)
        out = torch.cat([unpack_signs(all_packed[r]) * s_scales[r] for r in range(ws)])
        buffer_m.copy_(out[:n].reshape(shape).to(buffer_m.dtype))
        return buffer_m


NcclBackend = MpiBackend = HcclBackend = CompressedBackend
