#!/bin/bash
# usage: tools/ncu_capture.sh <case> <kernel-regex>   (one GPU; plain run first, then ncu --set full)
set -o pipefail
case=$1; regex=$2
mkdir -p gpurun_out
python tools/ncu_case.py $case > gpurun_out/plain_$case.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:$regex -s 2 -c 2 -f -o gpurun_out/prof_$case python tools/ncu_case.py $case > gpurun_out/ncu_$case.log 2>&1
echo "ncu $case rc=$?"; tail -2 gpurun_out/ncu_$case.log
