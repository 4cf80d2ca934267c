#!/bin/bash
# cfg3 step time for the GEMM path variants (environment switches of gemm_tcgen05.cu)
mkdir -p gpurun_out
B="python bench.py --workload cfg3 --steps 5 --warmup 3 --no-e2e --no-cpu"
run() { echo "$1: $(env $1 timeout 300 $B 2>&1 | python -c 'import sys,json
t=sys.stdin.read()
try:
    j=json.loads([l for l in t.splitlines() if l.startswith("{")][-1]); print(j["ms_per_step"], "ms", round(j["value"]/1e6,1), "M rows/s frac", round(j["roofline"]["frac"],3), j["clocks"]["sm_mhz"], j["clocks"]["reasons"])
except Exception as e: print("FAILED", t[-600:])')"; }
run "LRS_GEMM_PAIR=1 LRS_GEMM_SPLIT=0"
run "LRS_GEMM_PAIR=1 LRS_GEMM_SPLIT=1"
run "LRS_GEMM_PAIR=0 LRS_GEMM_SPLIT=0"
run "LRS_GEMM_PAIR=0 LRS_GEMM_MC=0 LRS_GEMM_SPLIT=0"
run "LRS_GEMM_F16=0 LRS_GEMM_SPLIT=0"
run "LRS_GEMM_F16=0 LRS_GEMM_PAIR=0 LRS_GEMM_MC=0 LRS_GEMM_SPLIT=0"
