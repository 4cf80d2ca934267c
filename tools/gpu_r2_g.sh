#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gemm_gpu.py -q --timeout=600 -k short_rows > gpurun_out/g_pytest_gemm.log 2>&1
echo "pytest exit $?" >> gpurun_out/g_pytest_gemm.log
tail -3 gpurun_out/g_pytest_gemm.log
timeout 600 python tools/shape_sweep.py --only 16:26,32:20,54:20,64:32,48:16,20:13,64:17,61:31,28:24,36:20 --out gpurun_out/g_sweep_short.json 2>&1 | tail -10
LRS_SHORT_CONV_SETS=2 timeout 600 python tools/shape_sweep.py --only 32:20,28:24 2>&1 | tail -2
LRS_SHORT_CONV_SETS=1 timeout 600 python tools/shape_sweep.py --only 48:16,36:20 2>&1 | tail -2
