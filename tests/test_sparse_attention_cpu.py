"""Block-sparse attention: packed-block products and softmax against dense masked attention."""
import math
from types import SimpleNamespace

import pytest
import torch

from hcache_deepspeed_b200.ops.sparse_attention import (BertSparseSelfAttention, FixedSparsityConfig, MatMul, Softmax,
                                                        SparseAttentionUtils, SparseSelfAttention)


def _dense_from_blocks(x, layout, block):
    B, H, nb = x.shape[0], layout.shape[0], layout.shape[1]
    out = torch.zeros(B, H, nb * block, layout.shape[2] * block)
    for k, (h, i, j) in enumerate(layout.nonzero().tolist()):
        out[:, h, i * block:(i + 1) * block, j * block:(j + 1) * block] = x[:, k]
    return out


def test_block_sparse_pipeline_matches_dense_masked_attention():
    torch.manual_seed(0)
    B, H, L, D, blk = 2, 4, 64, 16, 16
    cfg = FixedSparsityConfig(num_heads=H, block=blk, num_local_blocks=2, num_global_blocks=1, different_layout_per_head=True)
    layout = cfg.make_layout(L)
    assert 0 < int(layout.sum()) < layout.numel()
    q, k, v = (torch.randn(B, H, L, D) for _ in range(3))
    kpm = torch.zeros(B, L)
    kpm[1, -5:] = float("-inf")
    sdd, dsd = MatMul(layout, blk, "sdd", trans_b=True), MatMul(layout, blk, "dsd")
    scores = sdd(q, k)
    assert scores.shape == (B, int(layout.sum()), blk, blk)
    probs = Softmax(layout, blk)(scores, scale=1 / math.sqrt(D), key_padding_mask=kpm)
    out = dsd(probs, v)
    mask = layout.bool().repeat_interleave(blk, 1).repeat_interleave(blk, 2)
    ref_s = (q @ k.transpose(-1, -2)) / math.sqrt(D)
    ref_s = ref_s.masked_fill(~mask[None], float("-inf")) + kpm[:, None, None, :]
    ref_p = torch.softmax(ref_s, -1)
    torch.testing.assert_close(_dense_from_blocks(probs, layout, blk), ref_p, atol=1e-5, rtol=1e-4)
    torch.testing.assert_close(out, ref_p @ v, atol=1e-5, rtol=1e-4)
    # the module (mask formulation) agrees as well
    mod = SparseSelfAttention(cfg, max_seq_length=L)
    torch.testing.assert_close(mod(q, k, v, key_padding_mask=kpm), out, atol=1e-5, rtol=1e-4)
    # dds: dense @ sparse, and the transposed-sparse variants
    w = torch.randn(B, H, 24, L)
    dense_p = _dense_from_blocks(probs, layout, blk)
    torch.testing.assert_close(MatMul(layout, blk, "dds")(w, probs), w @ dense_p, atol=1e-5, rtol=1e-4)
    torch.testing.assert_close(MatMul(layout, blk, "dsd", trans_a=True)(probs, v), dense_p.transpose(-1, -2) @ v, atol=1e-5, rtol=1e-4)
    torch.testing.assert_close(MatMul(layout, blk, "dds", trans_b=True)(w, probs), w @ dense_p.transpose(-1, -2), atol=1e-5, rtol=1e-4)
    with pytest.raises(NotImplementedError):
        MatMul(layout, blk, "ddd")


def test_bert_sparse_layer_and_padding_utils():
    torch.manual_seed(0)
    conf = SimpleNamespace(hidden_size=32, num_attention_heads=4)
    layer = BertSparseSelfAttention(conf, FixedSparsityConfig(num_heads=4, block=16))
    x = torch.randn(2, 32, 32)
    out = layer(x, torch.zeros(2, 1, 1, 32))
    assert out.shape == (2, 32, 32) and torch.isfinite(out).all()
    emb = torch.nn.Embedding(10, 8)
    ids = torch.randint(1, 10, (2, 21))
    pad, ids2, am, tt, pos, ie = SparseAttentionUtils.pad_to_block_size(16, ids, torch.ones(2, 21), None, None, None, 0, emb)
    assert pad == 11 and ids2.shape == (2, 32) and am[:, -11:].sum() == 0 and tt is None and ie is None
    assert SparseAttentionUtils.unpad_sequence_output(pad, torch.zeros(2, 32, 4)).shape == (2, 21, 4)
    bert = SimpleNamespace(bert=SimpleNamespace(embeddings=SimpleNamespace(position_embeddings=torch.nn.Embedding(8, 4))),
                           config=SimpleNamespace(max_position_embeddings=8))
    SparseAttentionUtils.extend_position_embedding(bert, 20)
    w = bert.bert.embeddings.position_embeddings.weight
    assert w.shape == (20, 4) and torch.equal(w[8:16], w[:8]) and bert.config.max_position_embeddings == 20
