mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gemm_gpu.py -q --timeout=600 -x > gpurun_out/i_pytest.log 2>&1; echo "pytest exit $?"; tail -3 gpurun_out/i_pytest.log
timeout 600 python tools/rescoring_rate.py 8192:20 16384:32 2>&1 | tail -4
timeout 600 python tools/shape_sweep.py --only 8192:20,4096:64,256:32,1024:128 2>&1 | tail -4
