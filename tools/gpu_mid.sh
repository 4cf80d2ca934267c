# mid-kernel tuning: lanes per row against shape
S="64:10,256:10,784:10,3072:10,4096:10,128:16,784:16,784:6"
for l in 4 8 16 32; do
  echo "lanes=$l"; LRS_MID_LANES=$l timeout 300 python tools/shape_sweep.py --only $S 2>&1 | python -c '
import sys, json
for l in sys.stdin:
    if l.startswith("{"):
        j = json.loads(l); print("  d=%-5d C=%-3d %7.1f GB/s %6.2f TF" % (j["d"], j["C"], j["GBps"], j["TFLOPs"]))'
done
