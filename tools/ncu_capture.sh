#!/bin/bash
# usage: tools/ncu_capture.sh <case> <kernel-regex>   (one GPU; plain run first, then ncu --set full)
# Leaves gpurun_out/ncu_<case>_raw.csv (all metrics of the captured launch) and ncu_<case>_source.csv (per-instruction
# stall samples); the .ncu-rep itself is dropped unless KEEP_REP=1 (reports with source are ~10 MB each).
set -o pipefail
case=$1; regex=$2
mkdir -p gpurun_out
python tools/ncu_case.py $case > gpurun_out/plain_$case.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:$regex -s 2 -c 1 -f -o gpurun_out/prof_$case python tools/ncu_case.py $case > gpurun_out/ncu_$case.log 2>&1
rc=$?
if [ -f gpurun_out/prof_$case.ncu-rep ]; then
  ncu -i gpurun_out/prof_$case.ncu-rep --page raw --csv > gpurun_out/ncu_${case}_raw.csv 2>/dev/null
  ncu -i gpurun_out/prof_$case.ncu-rep --page source --csv > gpurun_out/ncu_${case}_source.csv 2>/dev/null
  [ "$KEEP_REP" = "1" ] || rm -f gpurun_out/prof_$case.ncu-rep
fi
echo "ncu $case rc=$rc"; tail -1 gpurun_out/ncu_$case.log
