#!/bin/bash
# round 2, call B: two-pass epilogue -- parity, then cfg3 layouts interleaved with long timed regions
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_gemm_gpu.py -q --timeout=600 -x > gpurun_out/b_pytest_gemm.log 2>&1
echo "pytest exit $?" >> gpurun_out/b_pytest_gemm.log
tail -15 gpurun_out/b_pytest_gemm.log
B="python bench.py --workload cfg3 --rows 4000000 --steps 40 --warmup 5 --no-e2e --no-cpu"
for L in 0 2 0 2 1; do
  echo "layout $L: $(LRS_GEMM_LAYOUT=$L timeout 300 $B 2>&1 | tail -1 | python -c 'import sys,json
try:
    j=json.loads(sys.stdin.read()); print(j["ms_per_step"], "ms frac", round(j["roofline"]["frac"],3), j["run"]["rows_rescored_fp32"], j["clocks"])
except Exception as e: print("FAILED", e)')"
done
for S in 1 2 3; do
  echo "layout 0 skip=$S: $(LRS_GEMM_LAYOUT=0 LRS_GEMM_DEBUG_SKIP=$S timeout 300 $B 2>&1 | tail -1 | python -c 'import sys,json
try:
    j=json.loads(sys.stdin.read()); print(j["ms_per_step"], "ms frac", round(j["roofline"]["frac"],3), j["clocks"])
except Exception as e: print("FAILED", e)')"
done
