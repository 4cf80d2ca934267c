B="python bench.py --workload cfg3 --steps 5 --warmup 3 --no-e2e --no-cpu"
for sc in 3 8; do for split in 1 0; do
  echo "scale=$sc split=$split: $(LRS_GEMM_SPLIT=$split timeout 300 $B --logit-scale $sc | python -c 'import sys,json; j=json.loads(sys.stdin.read()); print(j["ms_per_step"], j["roofline"]["frac"])')"
done; done
