#!/bin/bash
mkdir -p gpurun_out
run() { local port=$1 log=$2; shift 2; timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port $port "$@" > gpurun_out/$log 2>&1; echo "$log rc=$?"; grep '^{' gpurun_out/$log | cut -c1-1500; }
run 29561 tp8_70b_kv_bf16_packed.log bench.py --gpus 8 --model llama-2-70b --ctx 16384 --plan planner --steps 1 --warmup 1
run 29562 tp8_bench_ours_final.log bench.py --gpus 8 --steps 2 --warmup 3
