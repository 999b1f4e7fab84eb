#!/bin/bash
# per-kernel device time of the NVTX range $1 ("decode" or "ttft") of tools/decode_profile.py
set -o pipefail
range=$1
mkdir -p gpurun_out
python tools/decode_profile.py > gpurun_out/decode_profile_plain.log 2>&1 &&
ncu --nvtx --nvtx-include "$range/" --metrics gpu__time_duration.sum --clock-control none --csv \
    --log-file gpurun_out/launches_$range.csv python tools/decode_profile.py > gpurun_out/ncu_launches_$range.log 2>&1
echo "launch list $range rc=$?"; cat gpurun_out/decode_profile_plain.log | tail -4
