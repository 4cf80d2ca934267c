#!/bin/bash
# round 2, call G: short-row tensor kernel -- parity and sweep
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gemm_gpu.py -q --timeout=600 -k "short_rows" > gpurun_out/g_pytest_short.log 2>&1
echo "pytest exit $?" >> gpurun_out/g_pytest_short.log
tail -30 gpurun_out/g_pytest_short.log
timeout 600 python tools/shape_sweep.py --only 16:26,32:20,54:20,64:32,48:16,20:13,64:17 --out gpurun_out/g_sweep_short.json 2>&1 | tail -8
LRS_NO_SHORT_ROWS=1 timeout 600 python tools/shape_sweep.py --only 16:26,32:20,54:20,64:32 2>&1 | tail -4
