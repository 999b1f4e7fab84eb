#!/bin/bash
# Multi-GPU check of the tensor-parallel path: numerics tests, then bench.py for both arms at TP = N.
# usage: tools/tp_gpu_check.sh N [steps]    (logs under gpurun_out/)
N=${1:-2}
STEPS=${2:-3}
mkdir -p gpurun_out
nvidia-smi --query-gpu=index,name --format=csv,noheader > gpurun_out/tp${N}_gpus.txt 2>&1
nvidia-smi topo -m > gpurun_out/tp${N}_topo.txt 2>&1
timeout 900 python -m pytest tests/test_gpu_tp.py -m gpu -q --timeout 600 -s -k "test_tp_matches_single_gpu" > gpurun_out/tp${N}_tests.log 2>&1
echo "tp tests rc=$? $(tail -1 gpurun_out/tp${N}_tests.log)"
grep -E "logits match|FAILED|Error|error" gpurun_out/tp${N}_tests.log | head -20
run() { timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $1 bench.py --gpus $N --steps $STEPS --warmup 3 "${@:3}" > gpurun_out/$2 2>&1; echo "$2 rc=$? $(grep -h '"metric"\|unavailable' gpurun_out/$2 | tail -1 | cut -c1-1500)"; }
run 29511 tp${N}_bench_ours.log
run 29512 tp${N}_bench_ref.log --impl reference
