#!/bin/bash
# Per-kernel device time of ONE tensor-parallel rank's decode step (Llama-2-7B at TP=8, 64 sessions), run on one GPU without
# the collectives: what the per-rank kernels cost once the weights are 8x smaller.
mkdir -p gpurun_out
export HEADS=4 KV_HEADS=4 INTER=1376 VOCAB=4000 B=64 GRAPH=0 D=3
python tools/decode_profile.py > gpurun_out/rank8_decode_plain.log 2>&1 && \
ncu --nvtx --nvtx-include "decode/" --metrics gpu__time_duration.sum --clock-control none --csv \
    --log-file gpurun_out/rank8_launches_decode.csv python tools/decode_profile.py > gpurun_out/rank8_ncu.log 2>&1
echo "rank-shape launch list rc=$?"; tail -2 gpurun_out/rank8_decode_plain.log
GRAPH=1 D=8 python tools/decode_profile.py 2>&1 | tail -1
