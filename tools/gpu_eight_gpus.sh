#!/bin/bash
# eight GPUs of one box: the default bench line under torchrun (cfg 2 at its named size: 2M rows x 100k in one resident
# round; cfg 3: 20M rows; the refit record; the e2e figure next to its bare-copy ceiling) and /predict served by one
# engine process over all eight devices.
mkdir -p gpurun_out
nvidia-smi --query-gpu=index,name --format=csv,noheader > gpurun_out/n8_gpus.txt 2>&1; wc -l gpurun_out/n8_gpus.txt
nproc; free -g | head -2
timeout 1500 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29521 bench.py --gpus 8 --steps 10 --warmup 3 > gpurun_out/n8_bench.json 2> gpurun_out/n8_bench.err
echo "bench exit $?"; tail -2 gpurun_out/n8_bench.err
python - <<'PY'
import json
try:
    j = json.loads(open("gpurun_out/n8_bench.json").read().strip().splitlines()[-1])
    print("cfg1", j["n_gpus"], j["value"], j["ms_per_step"], "e2e", j["e2e"]["value"], j["e2e"]["h2d_gbs"], j["e2e"]["h2d_ceiling_gbs"])
    for k, v in j.get("configs", {}).items():
        print(k, v.get("value"), v.get("ms_per_step"), v.get("roofline", {}).get("frac"), "e2e", v.get("e2e", {}).get("value"), v.get("e2e", {}).get("h2d_gbs"), v.get("e2e", {}).get("h2d_ceiling_gbs"), v.get("error"))
    print("refit", json.dumps(j.get("refit")))
except Exception as e:
    print("FAILED", e)
PY
timeout 300 python tools/loadtest.py --engine --devices 0,1,2,3,4,5,6,7 --workers 6 --clients 6 --concurrency 128 --fast-client --seconds 8 --no-batcher-only --reference-restated --out gpurun_out/n8_serving_engine_8gpus.json > gpurun_out/n8_serving.log 2>&1; echo "serving exit $?"
python - <<'PY'
import json
try:
    j = json.load(open("gpurun_out/n8_serving_engine_8gpus.json"))
    for k, v in j.items():
        if isinstance(v, dict) and "req_per_s" in v:
            print(k, round(v["req_per_s"]), "req/s p50", round(v["p50_ms"], 2), "p99", round(v["p99_ms"], 2), v.get("mean_batch_under_load"))
except Exception as e:
    print("FAILED", e)
PY
