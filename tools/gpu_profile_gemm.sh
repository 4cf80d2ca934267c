#!/bin/bash
# ncu capture of the GEMM path (cfg3, 400k rows): launch list + full set of the main kernel
mkdir -p gpurun_out
CMD3="python bench.py --workload cfg3 --rows 400000 --steps 2 --warmup 3 --no-e2e --no-cpu"
timeout 300 $CMD3 > gpurun_out/plain3.log 2>&1 &&
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:gemm_ -c 20 --csv \
    --log-file gpurun_out/launches_cfg3.csv $CMD3 > gpurun_out/ncu_list3.log 2>&1 &&
timeout 600 ncu --set full --clock-control none --import-source on -k regex:gemm_softmax -s 3 -c 1 \
    -f -o gpurun_out/prof_gemm $CMD3 > gpurun_out/ncu_full3.log 2>&1
echo "ncu cfg3 exit $?"
timeout 600 python bench.py --workload cfg3 --steps 10 --warmup 3 > gpurun_out/bench_cfg3.json 2> gpurun_out/bench_cfg3.err; cat gpurun_out/bench_cfg3.json | cut -c1-300
