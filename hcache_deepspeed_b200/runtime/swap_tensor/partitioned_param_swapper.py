"""NVMe tier for ZeRO-3 parameter partitions: each parameter's slice lives in its own file; a pool of pinned buffers
stages swap-ins (async, several in flight) and swap-outs through the aio thread pool. The same machinery backs the
HCache NVMe tier (``hcache/store.py``). Reference: ``deepspeed/runtime/swap_tensor/partitioned_param_swapper.py:37-422``."""
from __future__ import annotations

import os
import shutil
from enum import Enum
from typing import Dict, List, Optional

import torch

from ...io.aio import aio_handle as AioHandle
from .utils import MIN_AIO_BYTES, AIO_ALIGNED_BYTES


class PartitionedParamStatus(Enum):
    AVAILABLE = 1
    NOT_AVAILABLE = 2
    INFLIGHT = 3


class AsyncPartitionedParameterSwapper:

    def __init__(self, ds_config=None, model_dtype=torch.float16, nvme_path: Optional[str] = None, buffer_count: int = 5,
                 buffer_size: int = int(1e8), aio_config: Optional[dict] = None):
        if ds_config is not None:
            off = ds_config.zero_config.offload_param
            nvme_path, buffer_count, buffer_size = str(off.nvme_path), off.buffer_count, off.buffer_size
            aio_config = ds_config.aio_config
        aio_config = {"block_size": 1048576, "queue_depth": 8, "intra_op_parallelism": 1, "single_submit": False,
                      "overlap_events": True, **(aio_config or {})}
        self.dtype = model_dtype
        self.swap_element_size = torch.tensor([], dtype=model_dtype).element_size()
        rank = int(os.environ.get("RANK", 0))
        self.swap_folder = os.path.join(nvme_path, "zero_stage_3", f"{str(model_dtype).split('.')[-1]}params", f"rank{rank}")
        shutil.rmtree(self.swap_folder, ignore_errors=True)
        os.makedirs(self.swap_folder, exist_ok=True)
        self.min_aio_bytes = max(MIN_AIO_BYTES, aio_config["block_size"])
        self.aligned_bytes = AIO_ALIGNED_BYTES * aio_config["intra_op_parallelism"]
        self.numel_alignment = self.aligned_bytes // self.swap_element_size
        self.elements_per_buffer = self._aligned(buffer_size)
        self.param_buffer_count = buffer_count
        buf = torch.empty(self.elements_per_buffer * buffer_count, dtype=model_dtype)
        self.buffers = buf.pin_memory() if torch.cuda.is_available() else buf
        self.available_buffer_ids = list(range(buffer_count))
        self.param_id_to_buffer_id: Dict[int, int] = {}
        self.param_id_to_swap_buffer: Dict[int, torch.Tensor] = {}
        self.param_id_to_numel: Dict[int, int] = {}
        self.id_to_path: Dict[int, str] = {}
        self.inflight_params: List = []
        self.inflight_swap_in_buffers: List = []
        self.inflight_numel = 0
        self.pending_writes = self.pending_reads = 0
        self.available_params = set()
        self.available_numel = 0
        self.aio_read_handle = AioHandle(aio_config["block_size"], aio_config["queue_depth"], aio_config["single_submit"],
                                         aio_config["overlap_events"], aio_config["intra_op_parallelism"])
        self.aio_write_handle = AioHandle(aio_config["block_size"], aio_config["queue_depth"], aio_config["single_submit"],
                                          aio_config["overlap_events"], aio_config["intra_op_parallelism"])
        self.invalid_buffer = torch.tensor(1).to(model_dtype)

    def _aligned(self, numel: int) -> int:
        a = self.numel_alignment
        return (numel + a - 1) // a * a

    def available_swap_in_buffers(self) -> int:
        return len(self.available_buffer_ids)

    def swappable_tensor(self, param=None, numel: Optional[int] = None) -> bool:
        numel = numel if numel is not None else param.ds_tensor.ds_numel
        return self.min_aio_bytes <= numel * self.swap_element_size

    def get_path(self, param, must_exist: bool = False) -> str:
        pid = param.ds_id
        if pid not in self.id_to_path:
            assert not must_exist, f"Path for param id {pid} does not exist"
            self.id_to_path[pid] = os.path.join(self.swap_folder, f"{pid}_param.tensor.swp")
        return self.id_to_path[pid]

    def _allocate(self, param) -> torch.Tensor:
        pid = param.ds_id
        numel = param.ds_tensor.ds_numel
        assert self.available_buffer_ids, f"No swap buffers to allocate for param id {pid}"
        assert self._aligned(numel) <= self.elements_per_buffer, "parameter partition larger than a swap buffer"
        bid = self.available_buffer_ids.pop()
        self.param_id_to_buffer_id[pid] = bid
        self.param_id_to_numel[pid] = numel
        buf = self.buffers.narrow(0, bid * self.elements_per_buffer, self._aligned(numel))
        self.param_id_to_swap_buffer[pid] = buf
        return buf

    def swap_out_and_release(self, params, async_op: bool = False, force_buffer_release: bool = False):
        """Write each parameter's partition to its file and drop the host copy."""
        for p in params:
            pid = p.ds_id
            if pid in self.param_id_to_swap_buffer:
                buf = self.param_id_to_swap_buffer[pid]
            else:
                buf = self._allocate(p)
                buf.narrow(0, 0, p.ds_tensor.ds_numel).copy_(p.ds_tensor.data.reshape(-1))
            self.aio_write_handle.async_pwrite(buf, self.get_path(p))
            self.pending_writes += 1
        self.aio_write_handle.wait()
        self.pending_writes = 0
        for p in params:
            self._release(p)
            p.ds_tensor.data = self.invalid_buffer.data
            p.ds_tensor.status = PartitionedParamStatus.NOT_AVAILABLE

    def _release(self, p):
        pid = p.ds_id
        if pid in self.param_id_to_buffer_id:
            self.available_buffer_ids.append(self.param_id_to_buffer_id.pop(pid))
            self.param_id_to_swap_buffer.pop(pid, None)
        if pid in self.available_params:
            self.available_params.discard(pid)
            self.available_numel -= self.param_id_to_numel.get(pid, 0)

    def remove_partition_and_release_buffers(self, params):
        for p in params:
            self._release(p)
            p.ds_tensor.data = self.invalid_buffer.data
            p.ds_tensor.status = PartitionedParamStatus.NOT_AVAILABLE

    def swap_in(self, params, async_op: bool = True, swap_in_buffers=None):
        """Start reading the partitions of ``params`` into pinned buffers; ``synchronize_reads`` completes them."""
        for p in params:
            buf = self._allocate(p)
            self.aio_read_handle.async_pread(buf, self.get_path(p, must_exist=True))
            self.pending_reads += 1
            self.inflight_params.append(p)
            self.inflight_swap_in_buffers.append(buf)
            self.inflight_numel += p.ds_tensor.ds_numel
            p.ds_tensor.status = PartitionedParamStatus.INFLIGHT
        if not async_op:
            self.synchronize_reads()

    def synchronize_reads(self):
        if self.pending_reads == 0:
            return
        self.aio_read_handle.wait()
        self.pending_reads = 0
        for p, buf in zip(self.inflight_params, self.inflight_swap_in_buffers):
            p.ds_tensor.data = buf.narrow(0, 0, p.ds_tensor.ds_numel).data
            p.ds_tensor.status = PartitionedParamStatus.AVAILABLE
            self.available_params.add(p.ds_id)
        self.available_numel += self.inflight_numel
        self.inflight_params, self.inflight_swap_in_buffers, self.inflight_numel = [], [], 0

    def synchronize_writes(self):
        if self.pending_writes:
            self.aio_write_handle.wait()
            self.pending_writes = 0

    # ---- direct-to-destination reads, buffers for partitions being built, fp32 -> lp file updates (ref: :329-422) ---------
    def swap_into_buffer(self, param, dest_buffer: torch.Tensor) -> None:
        """Read ``param``'s partition straight into ``dest_buffer`` when that is aligned (pinned where a GPU is present);
        otherwise through a pool buffer, which is given back at once. The parameter itself stays NOT_AVAILABLE."""
        assert param.ds_tensor.status == PartitionedParamStatus.NOT_AVAILABLE, f"param {param.ds_id} is already available or inflight"
        numel = param.ds_tensor.ds_numel
        direct = dest_buffer.numel() == self._aligned(numel) and dest_buffer.is_contiguous() and \
            (dest_buffer.is_pinned() or not torch.cuda.is_available())
        if direct:
            self.aio_read_handle.async_pread(dest_buffer, self.get_path(param, must_exist=True))
            self.aio_read_handle.wait()
            return
        buf = self._allocate(param)
        self.aio_read_handle.async_pread(buf, self.get_path(param, must_exist=True))
        self.aio_read_handle.wait()
        dest_buffer.reshape(-1).narrow(0, 0, numel).copy_(buf.narrow(0, 0, numel))
        self._release(param)

    def get_buffer(self, param, numel: int) -> torch.Tensor:
        """A pool buffer of ``numel`` elements assigned to ``param`` (to fill with a partition that is then swapped out)."""
        pid = param.ds_id
        assert self.available_buffer_ids, f"No swap buffers to allocate for fp16 param {pid} of numel = {numel}"
        assert self._aligned(numel) <= self.elements_per_buffer, f"More elements {numel} than buffer size {self.elements_per_buffer}"
        bid = self.available_buffer_ids.pop()
        self.param_id_to_buffer_id[pid] = bid
        self.param_id_to_numel[pid] = numel
        swap = self.buffers.narrow(0, bid * self.elements_per_buffer, self._aligned(numel))
        self.param_id_to_swap_buffer[pid] = swap
        return swap.narrow(0, 0, numel)

    def reserve_partitioned_swap_space(self, partition_num_elems) -> None:
        """One staging area large enough for the (aligned) partitions listed: ``swap_out_partitioned_params`` converts
        fp32 master slices into it and writes all files with one batch of requests."""
        total = sum(self._aligned(n) for n in partition_num_elems)
        buf = torch.zeros(total, dtype=self.dtype)
        self.partitioned_swap_buffer = buf.pin_memory() if torch.cuda.is_available() else buf

    def swap_out_partitioned_params(self, dst_fp16_params, src_fp32_params) -> None:
        """After an optimizer step: the updated fp32 slices become the low-precision partition files of their parameters."""
        assert getattr(self, "partitioned_swap_buffer", None) is not None, "partitioned swap buffers for fp16 params not initialized"
        assert len(dst_fp16_params) == len(src_fp32_params), \
            f"mismatch in number of fp16 params {len(dst_fp16_params)} and fp32 params {len(src_fp32_params)}"
        self.synchronize_writes()
        off = 0
        for p, src in zip(dst_fp16_params, src_fp32_params):
            n, span = src.numel(), self._aligned(src.numel())
            assert off + span <= self.partitioned_swap_buffer.numel(), "reserved partitioned swap space is too small"
            stage = self.partitioned_swap_buffer.narrow(0, off, span)
            stage.narrow(0, 0, n).copy_(src.detach().reshape(-1))
            self.aio_write_handle.async_pwrite(stage, self.get_path(p, must_exist=True))
            self.pending_writes += 1
            off += span
        self.synchronize_writes()
        for p in dst_fp16_params:
            p.ds_tensor.status = PartitionedParamStatus.NOT_AVAILABLE

    def reserve_available_buffers(self) -> List[torch.Tensor]:
        return [self.buffers.narrow(0, b * self.elements_per_buffer, self.elements_per_buffer) for b in self.available_buffer_ids]

    def release_reserved_buffers(self):
        pass
