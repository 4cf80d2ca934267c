B="python bench.py --workload cfg3 --steps 5 --warmup 3 --no-e2e --no-cpu"
for sk in 0 16 2 3; do
  echo "skip=$sk: $(LRS_GEMM_MC=0 LRS_GEMM_FIXUP=0 LRS_GEMM_DEBUG_SKIP=$sk timeout 300 $B 2>&1 | python -c 'import sys,json
t=sys.stdin.read()
try:
    j=json.loads([l for l in t.splitlines() if l.startswith("{")][-1]); print(j["ms_per_step"], j["clocks"])
except Exception as e: print("FAILED", t[-600:])')"
done
