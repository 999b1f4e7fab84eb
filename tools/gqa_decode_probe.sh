#!/bin/bash
# One tensor-parallel rank of Llama-2-70B at TP=8 (8 query heads on 1 KV head, 64 sessions x 16K context, 8 layers):
# decode iteration with the SIMT split-KV kernel vs the head-packed tensor-core kernel.
export MODEL=llama-2-70b HEADS=8 KV_HEADS=1 INTER=3584 VOCAB=4000 B=64 CTX=16384 LAYERS=8 D=6 GRAPH=1
for packed in 0 1; do
  echo "HDS_GQA_PACKED=$packed: $(HDS_GQA_PACKED=$packed python tools/decode_profile.py 2>&1 | grep '^decode')"
done
