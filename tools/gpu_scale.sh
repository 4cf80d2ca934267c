#!/bin/bash
# Multi-GPU: the driver's launch line for N ranks.  Usage: bash tools/gpu_scale.sh N
N=${1:-2}
mkdir -p gpurun_out
nvidia-smi --query-gpu=index,name --format=csv > gpurun_out/gpus.txt 2>&1
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 \
    bench.py --gpus $N --steps 20 --warmup 5 > gpurun_out/scale_n$N.json 2> gpurun_out/scale_n$N.err
echo "exit $?"; cat gpurun_out/scale_n$N.json; tail -5 gpurun_out/scale_n$N.err
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 \
    bench.py --impl reference --gpus $N --steps 3 --warmup 1 > gpurun_out/scale_ref_n$N.json 2>&1; tail -2 gpurun_out/scale_ref_n$N.json
