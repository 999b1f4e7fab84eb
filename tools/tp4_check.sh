#!/bin/bash
# 4-GPU run: named configuration 4 (OPT-30B TP=4, hidden states on the NVMe tier, fixed hidden/KV layer split) and
# configuration 3 (Llama-2-13B TP=2).
mkdir -p gpurun_out
df -h /tmp | tail -1 > gpurun_out/tp4_disk.txt
run() { local n=$1 port=$2 log=$3; shift 3; timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $port "$@" > gpurun_out/$log 2>&1; echo "$log rc=$?"; grep '^{' gpurun_out/$log | cut -c1-1600; grep -E "Error|error" gpurun_out/$log | tail -3; }
run 4 29551 tp4_opt30b_dram.log bench.py --gpus 4 --model opt-30b --steps 2 --warmup 3
run 4 29552 tp4_opt30b_nvme_split.log bench.py --gpus 4 --model opt-30b --source nvme --force-kv-layers 12 --steps 2 --warmup 3
run 4 29553 tp4_opt30b_nvme.log bench.py --gpus 4 --model opt-30b --source nvme --steps 2 --warmup 3
run 2 29554 tp2_13b.log bench.py --gpus 2 --model llama-2-13b --steps 2 --warmup 3
run 2 29555 tp2_13b_ref.log bench.py --gpus 2 --model llama-2-13b --steps 1 --warmup 1 --impl reference
