#!/bin/bash
# round 2, call A: new dynamic-range tests, accumulation error per row length, cfg3 layouts
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_gemm_gpu.py -q --timeout=600 > gpurun_out/a_pytest_gemm.log 2>&1
echo "pytest exit $?" >> gpurun_out/a_pytest_gemm.log
tail -40 gpurun_out/a_pytest_gemm.log
timeout 600 python tools/gemm_kappa.py > gpurun_out/a_kappa.log 2>&1; tail -40 gpurun_out/a_kappa.log
B="python bench.py --workload cfg3 --rows 4000000 --steps 5 --warmup 3 --no-e2e --no-cpu"
for L in 0 2 1; do
  echo "layout $L: $(LRS_GEMM_LAYOUT=$L timeout 300 $B 2>&1 | tail -1 | python -c 'import sys,json
try:
    j=json.loads(sys.stdin.read()); print(j["ms_per_step"], "ms frac", round(j["roofline"]["frac"],3), j["run"], j["clocks"]["sm_mhz"], j["clocks"]["reasons"], j["gpu_launches"])
except Exception as e: print("FAILED", e)')"
done
timeout 900 python bench.py > gpurun_out/a_bench_default.json 2> gpurun_out/a_bench_default.err; echo "bench exit $?"; cat gpurun_out/a_bench_default.json; tail -5 gpurun_out/a_bench_default.err
