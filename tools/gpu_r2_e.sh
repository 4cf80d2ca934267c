#!/bin/bash
# round 2, call D: write-back epilogue + register-resident re-scoring; gradient GEMM; refit timing
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_refit_gpu.py -q --timeout=600 -k "grad_gemm or model_update" > gpurun_out/e_pytest_grad.log 2>&1
echo "pytest exit $?" >> gpurun_out/e_pytest_grad.log
tail -25 gpurun_out/e_pytest_grad.log
timeout 1200 python -m pytest tests/test_gemm_gpu.py tests/test_refit_gpu.py -q --timeout=600 > gpurun_out/e_pytest.log 2>&1
echo "pytest exit $?" >> gpurun_out/e_pytest.log
tail -15 gpurun_out/e_pytest.log
B="python bench.py --workload cfg3 --rows 4000000 --steps 6 --warmup 3 --no-e2e --no-cpu"
for E in "LRS_GEMM_FIXUP=1" "LRS_GEMM_FIXUP=0" "LRS_GEMM_FIXUP=1" "LRS_GEMM_FIXUP=0"; do
  sleep 3
  echo "$E: $(env $E timeout 300 $B 2>&1 | tail -1 | python -c 'import sys,json
try:
    j=json.loads(sys.stdin.read()); print(j["ms_per_step"], "ms frac", round(j["roofline"]["frac"],3), j["run"]["rows_rescored_fp32"], j["clocks"])
except Exception as e: print("FAILED", e)')"
done
timeout 600 python tools/refit_bench_general.py --out gpurun_out/e_refit_general.json 2>&1 | tail -4
CMD3="python bench.py --workload cfg3 --rows 400000 --steps 2 --warmup 3 --no-e2e --no-cpu"
timeout 300 $CMD3 > gpurun_out/e_plain3.log 2>&1 &&
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file gpurun_out/e_launches_cfg3.csv $CMD3 > gpurun_out/e_ncu_list.log 2>&1
echo "ncu list exit $?"
