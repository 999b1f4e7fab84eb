#!/bin/bash
# 8-GPU confirmation run: TP numerics at tp=4/8, decode probe, bench.py both arms at N=8.
mkdir -p gpurun_out
nvidia-smi topo -m > gpurun_out/tp8_topo.txt 2>&1
timeout 600 python -m pytest tests/test_gpu_tp.py -m gpu -q --timeout 500 -s -k "4-True or 8-True" > gpurun_out/tp8_tests.log 2>&1
echo "tp tests rc=$? $(tail -1 gpurun_out/tp8_tests.log)"; grep -E "logits match|^FAILED" gpurun_out/tp8_tests.log
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29541 tools/tp_decode_probe.py --strong > gpurun_out/tp8_decode_probe.log 2>&1
echo "probe rc=$?"; grep "^{" gpurun_out/tp8_decode_probe.log
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29542 bench.py --gpus 8 --steps 2 --warmup 3 > gpurun_out/tp8_bench_ours.log 2>&1
echo "ours rc=$?"; grep '^{' gpurun_out/tp8_bench_ours.log | cut -c1-1800
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29543 bench.py --gpus 8 --steps 1 --warmup 1 --impl reference > gpurun_out/tp8_bench_ref.log 2>&1
echo "ref rc=$?"; grep '^{' gpurun_out/tp8_bench_ref.log | cut -c1-900
