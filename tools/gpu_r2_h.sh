#!/bin/bash
mkdir -p gpurun_out
CMD="python tools/shape_sweep.py --gb 1 --only 64:32,64:17,16:26"
timeout 300 $CMD > gpurun_out/h_plain.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/h_launches.csv $CMD > gpurun_out/h_ncu.log 2>&1
echo "ncu exit $?"; cat gpurun_out/h_plain.log | tail -3
