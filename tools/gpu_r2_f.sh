#!/bin/bash
# round 2, call F: whole GPU suite, serving (engine + fronts vs reference handler), sweep of 13..32-class shapes, default bench
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -x --timeout=600 > gpurun_out/f_pytest_gpu.log 2>&1
echo "pytest exit $?" >> gpurun_out/f_pytest_gpu.log
tail -6 gpurun_out/f_pytest_gpu.log
timeout 300 python tools/loadtest.py --engine --workers 6 --clients 6 --concurrency 128 --fast-client --seconds 8 --no-batcher-only --reference-restated --out gpurun_out/f_serving_engine_6fronts.json > gpurun_out/f_serving_engine.log 2>&1; echo "serving engine exit $?"
timeout 200 python tools/loadtest.py --workers 1 --clients 4 --concurrency 64 --fast-client --seconds 6 --out gpurun_out/f_serving_1worker.json > gpurun_out/f_serving_1w.log 2>&1; echo "serving 1w exit $?"
python - <<'PY'
import json
for f in ("f_serving_engine_6fronts", "f_serving_1worker"):
    try:
        j = json.load(open(f"gpurun_out/{f}.json"))
        for k, v in j.items():
            if isinstance(v, dict) and "req_per_s" in v:
                print(f, k, round(v["req_per_s"]), "req/s p50", round(v["p50_ms"], 2), "p99", round(v["p99_ms"], 2), v.get("mean_batch_under_load"), v.get("batcher", {}).get("batch_size_histogram") if isinstance(v.get("batcher"), dict) else "")
    except Exception as e:
        print(f, "FAILED", e)
PY
timeout 600 python tools/shape_sweep.py --only 1024:12,128:16,784:16,1024:16,256:13,32:20,54:20,16:26,64:32,768:256 --out gpurun_out/f_sweep.json 2>&1 | tail -14
LRS_GEMM_LAYOUT=2 timeout 300 python tools/shape_sweep.py --only 768:256,512:200 2>&1 | tail -3
timeout 900 python bench.py > gpurun_out/f_bench_default.json 2> gpurun_out/f_bench_default.err; echo "bench exit $?"
python - <<'PY'
import json
j = json.loads(open("gpurun_out/f_bench_default.json").read().strip().splitlines()[-1])
print("cfg1", j["ms_per_step"], j["roofline"]["frac"], j["e2e"]["value"], j["e2e"]["frac_of_h2d_ceiling"])
for k, v in j.get("configs", {}).items():
    print(k, v.get("ms_per_step"), v.get("roofline", {}).get("frac"), v.get("run"), v.get("e2e", {}).get("value"), v.get("clocks"))
PY
