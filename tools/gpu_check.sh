#!/bin/bash
# One gpurun call: GPU tests, smoke, bench lines, then ncu (launch list + full captures).
# Usage (from the repo root, on the GPU box): bash tools/gpu_check.sh [skip-ncu]
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,memory.total,clocks.max.sm --format=csv > gpurun_out/gpu.txt 2>&1
timeout 1500 python -m pytest tests -m gpu -q -x --timeout=600 > gpurun_out/pytest_gpu.log 2>&1
echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
tail -15 gpurun_out/pytest_gpu.log
timeout 300 python __graft_entry__.py --smoke > gpurun_out/smoke.log 2>&1; echo "smoke exit $?"; tail -3 gpurun_out/smoke.log
timeout 600 python bench.py --steps 20 --warmup 5 > gpurun_out/bench_cfg1.json 2> gpurun_out/bench_cfg1.err; echo "bench exit $?"; cat gpurun_out/bench_cfg1.json
timeout 300 python bench.py --impl reference --steps 5 --warmup 2 > gpurun_out/bench_ref.json 2>&1; cat gpurun_out/bench_ref.json
for w in cfg0 cfg2 cfg3; do
  timeout 600 python bench.py --workload $w --steps 10 --warmup 3 > gpurun_out/bench_$w.json 2> gpurun_out/bench_$w.err; cat gpurun_out/bench_$w.json
done
if [ "$1" != "skip-ncu" ]; then
  CMD="python bench.py --steps 2 --warmup 3 --no-e2e --no-cpu"
  timeout 300 $CMD > gpurun_out/plain.log 2>&1 &&
  timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv \
      --log-file gpurun_out/launches_cfg1.csv $CMD > gpurun_out/ncu_list.log 2>&1 &&
  timeout 600 ncu --set full --clock-control none --import-source on -k regex:narrow_ -s 3 -c 1 \
      -f -o gpurun_out/prof_narrow $CMD > gpurun_out/ncu_full.log 2>&1
  echo "ncu cfg1 exit $?"
  CMD2="python bench.py --workload cfg2 --rows 20000 --steps 2 --warmup 3 --no-e2e --no-cpu"
  timeout 300 $CMD2 > gpurun_out/plain2.log 2>&1 &&
  timeout 600 ncu --set full --clock-control none --import-source on -k regex:wide_kernel -s 3 -c 1 \
      -f -o gpurun_out/prof_wide $CMD2 > gpurun_out/ncu_full2.log 2>&1
  echo "ncu cfg2 exit $?"
  CMD3="python bench.py --workload cfg3 --rows 400000 --steps 2 --warmup 3 --no-e2e --no-cpu"
  timeout 300 $CMD3 > gpurun_out/plain3.log 2>&1 &&
  timeout 600 ncu --set full --clock-control none --import-source on -k regex:gemm_softmax -s 3 -c 1 \
      -f -o gpurun_out/prof_gemm $CMD3 > gpurun_out/ncu_full3.log 2>&1
  echo "ncu cfg3 exit $?"
fi
