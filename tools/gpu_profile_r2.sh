#!/bin/bash
# Round-2 evidence in one gpurun call (1 GPU): GPU test suite, smoke, default bench (cfg1 + cfg2 + cfg3 records) and
# reference arm, then ncu: launch list of the default bench command, and one --set full capture each of the
# kernels that changed this round (tcgen05 scoring GEMM, short-row kernel, gradient GEMM).
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,memory.total,clocks.max.sm --format=csv > gpurun_out/r2_gpu.txt 2>&1
timeout 1500 python -m pytest tests -m gpu -q -x --timeout=600 > gpurun_out/r2_pytest_gpu.log 2>&1
echo "pytest exit $?" >> gpurun_out/r2_pytest_gpu.log; tail -3 gpurun_out/r2_pytest_gpu.log
timeout 300 python __graft_entry__.py --smoke > gpurun_out/r2_smoke.log 2>&1; echo "smoke exit $?"
timeout 900 python bench.py > gpurun_out/r2_bench_default.json 2> gpurun_out/r2_bench_default.err; echo "bench exit $?"
timeout 300 python bench.py --impl reference --steps 5 --warmup 2 > gpurun_out/r2_bench_reference_arm.json 2>/dev/null
timeout 300 python bench.py --workload cfg0 --steps 20 --warmup 5 > gpurun_out/r2_bench_cfg0.json 2>/dev/null
timeout 900 python tools/shape_sweep.py --out gpurun_out/r2_shape_sweep.json > gpurun_out/r2_shape_sweep.log 2>&1; echo "sweep exit $?"
timeout 600 python tools/refit_bench_general.py --out gpurun_out/r2_refit_general.json > /dev/null 2>&1; echo "refit bench exit $?"
timeout 300 python tools/rescoring_rate.py > gpurun_out/r2_rescoring_rate.txt 2>&1
timeout 300 python tools/gemm_kappa.py > gpurun_out/r2_gemm_accumulation_error.txt 2>&1
timeout 300 python tools/gemm_kappa.py --short >> gpurun_out/r2_gemm_accumulation_error.txt 2>&1
timeout 300 python tools/latency_single.py > gpurun_out/r2_latency_single.json 2>/dev/null
timeout 300 python tools/output_patterns.py > gpurun_out/r2_output_patterns.md 2>&1
python tools/sass_markers.py > gpurun_out/r2_sass_markers.txt 2>&1
if [ "$1" != "skip-ncu" ]; then
  CMD="python bench.py --steps 2 --warmup 3 --no-e2e --no-cpu"
  timeout 600 $CMD > gpurun_out/r2_plain.log 2>&1 &&
  timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -k "regex:narrow|wide_|gemm_|x_scale|short_|fixup|w_split|w_colscale|groups_" -c 400 --csv \
      --log-file gpurun_out/r2_launches_default.csv $CMD > gpurun_out/r2_ncu_list.log 2>&1
  echo "ncu list exit $?"
  CMD3="python bench.py --workload cfg3 --rows 400000 --steps 2 --warmup 3 --no-e2e --no-cpu"
  timeout 300 $CMD3 > gpurun_out/r2_plain3.log 2>&1 &&
  timeout 600 ncu --set full --clock-control none --import-source on -k regex:gemm_softmax_kernel -s 6 -c 1 \
      -f -o gpurun_out/r2_prof_gemm $CMD3 > gpurun_out/r2_ncu_full3.log 2>&1
  echo "ncu cfg3 exit $?"
  CMDS="python tools/shape_sweep.py --gb 1 --only 16:26,64:32"
  timeout 300 $CMDS > gpurun_out/r2_plain_short.log 2>&1 &&
  timeout 600 ncu --set full --clock-control none --import-source on -k regex:short_rows_kernel -s 2 -c 1 \
      -f -o gpurun_out/r2_prof_short $CMDS > gpurun_out/r2_ncu_short.log 2>&1
  echo "ncu short exit $?"
  CMDG="python tools/refit_bench_general.py"
  timeout 300 $CMDG > gpurun_out/r2_plain_grad.log 2>&1 &&
  timeout 600 ncu --set full --clock-control none --import-source on -k regex:grad_gemm_kernel -s 1 -c 1 \
      -f -o gpurun_out/r2_prof_grad $CMDG > gpurun_out/r2_ncu_grad.log 2>&1
  echo "ncu grad exit $?"
fi
