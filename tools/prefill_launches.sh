#!/bin/bash
# per-kernel device time of one fresh 4160-token prefill (tools/prefill_probe.py): plain run first, then ncu (basic metric)
mkdir -p gpurun_out
PROBE_ITERS=2 python tools/prefill_probe.py > gpurun_out/prefill_probe_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none --csv -c 1200 --log-file gpurun_out/launches_prefill.csv \
    python tools/prefill_probe.py > gpurun_out/ncu_launches_prefill.log 2>&1
echo "prefill launch list rc=$?"; tail -1 gpurun_out/prefill_probe_plain.log
