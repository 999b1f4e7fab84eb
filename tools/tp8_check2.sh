#!/bin/bash
# 8-GPU run: 7B decode probe + bench at TP=8, then the named 70B configuration (64 sessions, 16K restored context).
mkdir -p gpurun_out
run() { local port=$1 log=$2; shift 2; timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port $port "$@" > gpurun_out/$log 2>&1; echo "$log rc=$?"; grep '^{' gpurun_out/$log | cut -c1-${CUT:-1500}; }
CUT=220 run 29541 tp8_decode_probe.log tools/tp_decode_probe.py --strong
run 29542 tp8_bench_ours.log bench.py --gpus 8 --steps 2 --warmup 3
run 29543 tp8_70b_kv_bf16.log bench.py --gpus 8 --model llama-2-70b --ctx 16384 --plan planner --steps 1 --warmup 1
run 29544 tp8_70b_kv_fp8.log bench.py --gpus 8 --model llama-2-70b --ctx 16384 --plan planner --quant wf8af8 --steps 1 --warmup 1
run 29545 tp8_70b_hidden_bf16.log bench.py --gpus 8 --model llama-2-70b --ctx 16384 --plan hidden --steps 1 --warmup 1
tail -5 gpurun_out/tp8_70b_*.log | grep -v "^{" | tail -20
