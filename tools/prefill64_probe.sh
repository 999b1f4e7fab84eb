#!/bin/bash
# 64-key-tile / 2-CTAs-per-SM prefill attention: numerics tests, then kernel timing against the 128-key-tile kernel.
timeout 200 python -m pytest tests/test_gpu_kernels.py -m gpu -q -x -k "attention_prefill" --timeout 100 2>&1 | tail -3
timeout 200 python -m pytest tests/test_gpu_engine.py -m gpu -q -x --timeout 100 2>&1 | tail -2
for tile in 128 64; do
  echo "HDS_PREFILL_TILE=$tile: $(HDS_PREFILL_TILE=$tile timeout 200 python tools/prefill_probe.py 2>&1 | grep fresh_prefill | tail -1)"
  HDS_PREFILL_TILE=$tile timeout 100 python - <<'PY'
import math, os, sys, torch
sys.path.insert(0, os.getcwd())
from hcache_deepspeed_b200.ops import kernels as K
sys.path.insert(0, "tests")
from tests.test_gpu_kernels import build_paged_problem
specs=[(i*128,128) for i in range(0,33)]          # one 4224-token causal prompt as 33 work items
q,cache,tables,items=build_paged_problem([(0,4224)],32,32,128,64,torch.bfloat16,seed=1)
items=[(0,i*128,128,i*128) for i in range(33)]
it=torch.tensor(items,dtype=torch.int32,device="cuda"); out=torch.zeros_like(q)
def run(): K.paged_attention_prefill(q,out,cache,tables,it,len(items),32,32,128,64,1/math.sqrt(128))
for _ in range(3): run()
torch.cuda.synchronize(); ts=[]
for _ in range(10):
    s,e=torch.cuda.Event(enable_timing=True),torch.cuda.Event(enable_timing=True); s.record(); run(); e.record(); torch.cuda.synchronize(); ts.append(s.elapsed_time(e)*1e3)
ts.sort(); us=ts[len(ts)//2]; flops=4*128*32*sum((i+1)*128*128 for i in range(33)) - 0
print(f"  causal 4224 tokens x 32 heads x d128: {us:.1f} us, {flops/us/1e6:.0f} TFLOP/s (tile {os.environ['HDS_PREFILL_TILE']})")
PY
done
