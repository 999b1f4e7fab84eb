"""hds_paged_attention_splitkv_static: the split-KV attention entry point for callers that cannot allocate.

Some callers have no scratch buffer to give: an engine that handed all free HBM to its KV cache, a C host
without an allocator to borrow from, a loop that cannot afford cudaMalloc's device-wide synchronisation. For them
the library keeps the scratch in a fixed 8 MiB device array of its own. This test drives that entry point the
way such a caller would — through the C ABI with raw pointers, nothing but ctypes in between — and checks that

* two calls in a row (the second with MORE work items, so its arrival counters land where the first call left
  partial sums) both reproduce the caller-scratch entry point within the tolerance the kernel tests use;
* a problem that needs more than the built-in array is refused with HDS_ERR_TOO_LARGE and an explanation.

The file name sorts last on purpose: with `pytest -x` every other test has run before this one can fail.
"""
import ctypes
import math
import os

import pytest
import torch

pytestmark = [pytest.mark.gpu, pytest.mark.timeout(300)]

LIB = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "hcache_deepspeed_b200", "ops", "lib",
                   "libhds_b200.so")


def _load():
    P, I, L, F = ctypes.c_void_p, ctypes.c_int, ctypes.c_int64, ctypes.c_float
    lib = ctypes.CDLL(LIB)
    lib.hds_last_error.restype = ctypes.c_char_p
    lib.hds_paged_attention_splitkv_workspace_bytes.argtypes = [I, I, I, I, I, I]
    lib.hds_paged_attention_splitkv_workspace_bytes.restype = L
    lib.hds_paged_attention_splitkv_static_capacity.restype = L
    lib.hds_paged_attention_splitkv.argtypes = [P, P, P, L, P, I, P, I, I, I, I, I, I, I, F, P, L, I, P]
    lib.hds_paged_attention_splitkv.restype = I
    lib.hds_paged_attention_splitkv_static.argtypes = [P, P, P, L, P, I, P, I, I, I, I, I, I, I, F, I, P]
    lib.hds_paged_attention_splitkv_static.restype = I
    return lib


def _problem(ctx_lens, n_heads, d, bs, seed):
    g = torch.Generator().manual_seed(seed)
    nblk = [(c + 1 + bs - 1) // bs for c in ctx_lens]
    total = sum(nblk) + 3
    perm = torch.randperm(total, generator=g).tolist()
    cache = (torch.randn(total, 2, n_heads, bs, d, generator=g) * 0.5).to(torch.bfloat16).cuda()
    tables = torch.zeros(len(ctx_lens), max(nblk), dtype=torch.int32)
    items = []
    for s, (c, nb) in enumerate(zip(ctx_lens, nblk)):
        tables[s, :nb] = torch.tensor([perm.pop() for _ in range(nb)], dtype=torch.int32)
        items.append((s, s, 1, c))
    q = torch.randn(len(ctx_lens), n_heads, d, generator=g).to(torch.bfloat16).cuda()
    return q, cache, tables.cuda(), torch.tensor(items, dtype=torch.int32).cuda()


def _run_pair(lib, ctx_lens, n_splits, seed, n_heads=8, d=128, bs=64):
    q, cache, tables, items = _problem(ctx_lens, n_heads, d, bs, seed)
    n = len(ctx_lens)
    scale = 1.0 / math.sqrt(d)
    stream = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    need = lib.hds_paged_attention_splitkv_workspace_bytes(n, n_heads, n_heads, d, n_splits, 1)
    assert need <= lib.hds_paged_attention_splitkv_static_capacity()
    ws = torch.zeros(need, dtype=torch.uint8, device="cuda")
    want, got = torch.zeros_like(q), torch.zeros_like(q)
    rc = lib.hds_paged_attention_splitkv(q.data_ptr(), want.data_ptr(), cache.data_ptr(), cache.shape[0], tables.data_ptr(),
                                         tables.stride(0), items.data_ptr(), n, 1, n_heads, n_heads, d, bs, n_splits, scale,
                                         ws.data_ptr(), need, 0, stream)
    assert rc == 0, lib.hds_last_error()
    rc = lib.hds_paged_attention_splitkv_static(q.data_ptr(), got.data_ptr(), cache.data_ptr(), cache.shape[0],
                                                tables.data_ptr(), tables.stride(0), items.data_ptr(), n, 1, n_heads,
                                                n_heads, d, bs, n_splits, scale, 0, stream)
    assert rc == 0, lib.hds_last_error()
    torch.cuda.synchronize()
    torch.testing.assert_close(got.float(), want.float(), rtol=1.6e-2, atol=2e-2)


def test_static_scratch_matches_caller_scratch_twice_in_a_row():
    lib = _load()
    # call 1: few items, many splits -> partial sums spread over the front of the array
    _run_pair(lib, [4095, 777, 1500], n_splits=8, seed=1)
    # call 2: many more items -> its counters occupy bytes that call 1 filled with partial sums
    _run_pair(lib, [300 + 131 * i for i in range(40)], n_splits=4, seed=2)


def test_static_scratch_rejects_oversized_problem():
    lib = _load()
    cap = lib.hds_paged_attention_splitkv_static_capacity()
    n_items, heads, d, splits = 512, 32, 128, 16
    assert lib.hds_paged_attention_splitkv_workspace_bytes(n_items, heads, heads, d, splits, 1) > cap
    dummy = torch.zeros(16, dtype=torch.uint8, device="cuda")
    p = dummy.data_ptr()
    rc = lib.hds_paged_attention_splitkv_static(p, p, p, 1, p, 1, p, n_items, 1, heads, heads, d, 64, splits, 1.0, 0, None)
    assert rc == -4
    assert b"scratch" in lib.hds_last_error()
