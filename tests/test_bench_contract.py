"""bench.py's line contract, as far as it can be checked without a GPU: the reference arm runs (scikit-learn on the
host cores, a bounded sample), prints one JSON line with the contract's keys, and names the workload with exactly
the `config` object the B200 arm prints for the same arguments."""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_reference_arm_line_and_shared_config():
    sys.path.insert(0, ROOT)
    import bench
    out = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--steps", "1",
                          "--warmup", "1", "--rows", "20000"], capture_output=True, text=True, timeout=300, cwd=ROOT)
    assert out.returncode == 0, out.stderr[-2000:]
    line = json.loads(out.stdout.strip().splitlines()[-1])
    for key in ("impl", "metric", "value", "unit", "n_gpus", "steps", "warmup", "ms_per_step", "higher_is_better",
                "scaling", "vs_baseline", "dtype", "data", "config", "cpu_baseline", "e2e", "gpu_launches"):
        assert key in line, key
    assert line["impl"] == "reference" and line["metric"] == bench.METRIC and line["unit"] == bench.UNIT
    assert line["gpu_launches"] == 0 and line["e2e"]["h2d_bytes_per_step"] == 0 and line["vs_baseline"] is None
    assert line["cpu_baseline"]["kind"] == "reference" and line["cpu_baseline"]["cores"] >= 1
    assert line["run"]["rows_scored_per_step"] == 20000
    # the B200 arm builds its `config` with the same function and the same arguments
    assert line["config"] == bench.config_of("cfg1", 20000, {"proba"}, 1.0, 1)
    # the sizes the workload texts name
    assert bench.WORKLOADS["cfg2"][1] * bench.WORKLOADS["cfg2"][2] * 4 == 100_000_000_000       # 100 GB per GPU
    assert bench.WORKLOADS["cfg2"][1] * 8 == 2_000_000 and bench.WORKLOADS["cfg3"][1] * 8 == 20_000_000
    assert bench.WORKLOADS["cfg1"][4] == 136 and bench.WORKLOADS["cfg3"][5] == 2 * 1024 * 1000
