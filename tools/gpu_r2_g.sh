#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gemm_gpu.py -q --timeout=600 > gpurun_out/g_pytest_gemm.log 2>&1
echo "pytest exit $?" >> gpurun_out/g_pytest_gemm.log
tail -3 gpurun_out/g_pytest_gemm.log
timeout 600 python tools/shape_sweep.py --only 16:26,32:20,54:20,64:32,48:16,20:13,64:17,61:31,28:24 --out gpurun_out/g_sweep_short.json 2>&1 | tail -9
B="python bench.py --workload cfg3 --rows 4000000 --steps 6 --warmup 3 --no-e2e --no-cpu"
for E in "LRS_GEMM_FIXUP=1" "LRS_GEMM_FIXUP=0"; do
  echo "$E: $(env $E timeout 300 $B 2>&1 | tail -1 | python -c 'import sys,json
try:
    j=json.loads(sys.stdin.read()); print(j["ms_per_step"], "ms frac", round(j["roofline"]["frac"],3), j["run"]["rows_rescored_fp32"], j["clocks"])
except Exception as e: print("FAILED", e)')"
done
