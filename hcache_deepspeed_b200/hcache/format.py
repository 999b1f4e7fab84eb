"""On-disk format of one session's saved state.

The reference defines no file format: a session's hidden states are just the CPU tensor
``[num_layers, n_tokens, hidden]`` returned by ``put`` (``llama_v2/model.py:211-218``) and persistence is left
to the caller. That in-memory layout is kept (layer-major, activation dtype); this module only adds a
self-describing container so a state can live on NVMe:

    bytes 0..7     magic  b"HCACHE01"
    bytes 8..15    header length H (little-endian uint64)
    bytes 16..16+H JSON header: dtype, n_tokens, hidden, kv geometry, per-layer {kind, offset, nbytes}, tokens offset
    payload        4096-byte aligned segments (O_DIRECT friendly), one per layer

Per-layer ``kind``: ``"hidden"`` — the QKV-GEMM input ``[n_tokens, hidden]`` (restored through the K/V projection
kernel); ``"kv"`` — the layer's K/V pages ``[n_blocks, 2, kv_heads, block_size, head_dim]`` (plain KV offload:
restored by copy); ``"none"`` — nothing stored (the layer is recomputed). The planner decides the mix.
"""
from __future__ import annotations

import json
import struct
from dataclasses import dataclass, field
from typing import Dict, List, Optional

import torch

MAGIC = b"HCACHE01"
ALIGN = 4096
_DT = {"bfloat16": torch.bfloat16, "float16": torch.float16, "float32": torch.float32}


def align_up(x: int, a: int = ALIGN) -> int:
    return (x + a - 1) // a * a


@dataclass
class LayerSegment:
    kind: str          # "hidden" | "kv" | "none"
    offset: int = 0    # byte offset in the file
    nbytes: int = 0
    shape: List[int] = field(default_factory=list)


@dataclass
class StateHeader:
    dtype: str
    n_tokens: int
    hidden: int
    n_layers: int
    kv_heads: int = 0
    head_dim: int = 0
    block_size: int = 0
    layers: List[LayerSegment] = field(default_factory=list)
    tokens_offset: int = 0
    tokens_nbytes: int = 0
    extra: Dict = field(default_factory=dict)

    def to_bytes(self) -> bytes:
        d = dict(self.__dict__)
        d["layers"] = [seg.__dict__ for seg in self.layers]
        body = json.dumps(d).encode()
        return MAGIC + struct.pack("<Q", len(body)) + body

    @staticmethod
    def from_bytes(buf: bytes) -> "StateHeader":
        if buf[:8] != MAGIC:
            raise ValueError("not an HCache state file (bad magic)")
        (n, ) = struct.unpack("<Q", buf[8:16])
        d = json.loads(buf[16:16 + n].decode())
        d["layers"] = [LayerSegment(**s) for s in d["layers"]]
        return StateHeader(**d)

    @property
    def torch_dtype(self) -> torch.dtype:
        return _DT[self.dtype]

    @property
    def payload_bytes(self) -> int:
        return sum(s.nbytes for s in self.layers)


HEADER_RESERVED = 64 * 1024  # the header block at the start of the file (aligned payload begins after it)


def plan_layout(dtype: torch.dtype, n_tokens: int, hidden: int, kinds: List[str], kv_shape: Optional[List[int]] = None,
                kv_heads: int = 0, head_dim: int = 0, block_size: int = 0, tokens_nbytes: int = 0) -> StateHeader:
    """Assign aligned file offsets to every layer segment."""
    elem = torch.tensor([], dtype=dtype).element_size()
    off = HEADER_RESERVED
    segs = []
    for kind in kinds:
        if kind == "hidden":
            shape = [n_tokens, hidden]
        elif kind == "kv":
            shape = list(kv_shape)
        elif kind == "none":
            segs.append(LayerSegment("none"))
            continue
        else:
            raise ValueError(f"unknown layer kind {kind}")
        nbytes = elem
        for s in shape:
            nbytes *= s
        segs.append(LayerSegment(kind, off, nbytes, shape))
        off = align_up(off + nbytes)
    hdr = StateHeader(str(dtype).replace("torch.", ""), n_tokens, hidden, len(kinds), kv_heads, head_dim, block_size, segs,
                      off, tokens_nbytes)
    if len(hdr.to_bytes()) > HEADER_RESERVED:
        raise ValueError("state header does not fit its reserved block")
    return hdr
