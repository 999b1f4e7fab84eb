#!/bin/bash
# ncu capture of the head-packed GQA decode attention kernel (one Llama-2-70B TP=8 rank: 8 query heads on 1 KV head).
mkdir -p gpurun_out
export MODEL=llama-2-70b HEADS=8 KV_HEADS=1 INTER=3584 VOCAB=4000 B=64 CTX=16384 LAYERS=2 D=2 GRAPH=0
python tools/decode_profile.py > gpurun_out/packed_plain.log 2>&1 || exit 1
# inside the NVTX "decode" range the only tcgen05 attention kernel is the head-packed one
ncu --set full --clock-control none --import-source on --nvtx --nvtx-include "decode/" -k regex:paged_attn_prefill_kernel -c 1 -s 1 \
    -o gpurun_out/ncu_packed_attn -f python tools/decode_profile.py > gpurun_out/ncu_packed_attn.log 2>&1
echo "ncu rc=$?"
ncu -i gpurun_out/ncu_packed_attn.ncu-rep --page raw --csv > gpurun_out/ncu_packed_attn_raw.csv 2>/dev/null
rm -f gpurun_out/ncu_packed_attn.ncu-rep
