#!/bin/bash
# two GPUs: the NCCL refit tests and the default bench line under torchrun
mkdir -p gpurun_out
nvidia-smi --query-gpu=index,name --format=csv > gpurun_out/n2_gpus.txt 2>&1
timeout 900 python -m pytest tests/test_refit_dist_gpu.py -q --timeout=600 > gpurun_out/n2_pytest_dist.log 2>&1
echo "pytest exit $?" >> gpurun_out/n2_pytest_dist.log
tail -5 gpurun_out/n2_pytest_dist.log
timeout 1200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 10 --warmup 3 > gpurun_out/n2_bench.json 2> gpurun_out/n2_bench.err
echo "bench exit $?"; tail -3 gpurun_out/n2_bench.err
python - <<'PY'
import json
j = json.loads(open("gpurun_out/n2_bench.json").read().strip().splitlines()[-1])
print("cfg1", j["n_gpus"], j["value"], j["ms_per_step"], j["e2e"]["value"], j["e2e"]["h2d_gbs"], j["e2e"]["h2d_ceiling_gbs"])
for k, v in j.get("configs", {}).items():
    print(k, v.get("value"), v.get("ms_per_step"), v.get("roofline", {}).get("frac"), v.get("e2e", {}).get("value"), v.get("error"))
print("refit", json.dumps(j.get("refit")))
PY
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --impl reference --gpus 2 --steps 2 --warmup 1 2>/dev/null | tail -1 | cut -c1-300
