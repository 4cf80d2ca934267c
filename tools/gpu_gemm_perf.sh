#!/bin/bash
mkdir -p gpurun_out
B="python bench.py --workload cfg3 --steps 5 --warmup 3 --no-e2e --no-cpu"
for pair in 1 0; do for split in 0 1; do
  echo "pair=$pair split=$split: $(LRS_GEMM_PAIR=$pair LRS_GEMM_SPLIT=$split timeout 300 $B 2>&1 | python -c 'import sys,json
t=sys.stdin.read()
try:
    j=json.loads([l for l in t.splitlines() if l.startswith("{")][-1]); print(j["ms_per_step"], j["roofline"]["frac"], j["gpu_launches"], j["clocks"])
except Exception as e: print("FAILED", t[-600:])')"
done; done
