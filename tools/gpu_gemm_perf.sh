#!/bin/bash
mkdir -p gpurun_out
B="python bench.py --workload cfg3 --steps 5 --warmup 3 --no-e2e --no-cpu"
for split in 1 0; do for fix in 1; do
  echo "split=$split fixup=$fix: $(LRS_GEMM_SPLIT=$split LRS_GEMM_FIXUP=$fix timeout 300 $B | python -c 'import sys,json; j=json.loads(sys.stdin.read()); print(j["ms_per_step"], j["roofline"]["frac"], j["gpu_launches"], j["clocks"])')"
done; done
CMD="python bench.py --workload cfg3 --rows 1000000 --steps 2 --warmup 3 --no-e2e --no-cpu"
timeout 300 $CMD > gpurun_out/plain_g.log 2>&1 &&
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:gemm_ -c 12 --csv --log-file gpurun_out/launches_gemm.csv $CMD > gpurun_out/ncu_g.log 2>&1
grep gemm_ gpurun_out/launches_gemm.csv | awk -F'","' '{print substr($5,1,40), $NF}' | tail -8
