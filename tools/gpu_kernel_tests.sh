#!/bin/bash
# Runs each GPU kernel test group in its own process (a trapped kernel poisons the CUDA context of
# its process only). Logs land in gpurun_out/.
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,memory.total,clocks.max.sm --format=csv > gpurun_out/gpu.txt 2>&1
rc_all=0
for grp in gemm_store_splitk gemm_store gemm_bias_act gemm_gated gemm_accum norm rope_kv_append embed kv_blocks restore_kv qkv_rope splitkv_decode splitkv_multirow attention_prefill; do
  timeout 600 python -m pytest tests/test_gpu_kernels.py -m gpu -q -k "$grp" --timeout 240 -x > gpurun_out/test_$grp.log 2>&1
  rc=$?
  echo "$grp rc=$rc $(tail -1 gpurun_out/test_$grp.log)"
  [ $rc -ne 0 ] && rc_all=1
done
exit $rc_all
