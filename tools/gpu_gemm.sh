#!/bin/bash
# First contact / regression of the tcgen05 GEMM path.  Every step has its own timeout.
mkdir -p gpurun_out
timeout 180 python tools/gemm_debug.py > gpurun_out/gemm_debug.log 2>&1; echo "gemm_debug exit $?"; tail -12 gpurun_out/gemm_debug.log
timeout 600 python -m pytest tests/test_gemm_gpu.py -q -x --timeout=300 > gpurun_out/pytest_gemm.log 2>&1; echo "pytest gemm exit $?"; tail -15 gpurun_out/pytest_gemm.log
timeout 600 python bench.py --workload cfg3 --steps 5 --warmup 3 --no-e2e > gpurun_out/bench_cfg3.json 2> gpurun_out/bench_cfg3.err; echo "bench cfg3 exit $?"; cat gpurun_out/bench_cfg3.json; tail -3 gpurun_out/bench_cfg3.err
