for sp in 2 4 6 8 12 16; do
  echo "splits=$sp"
  B200_MARLIN_SPLITS=$sp timeout 120 python benchmarks/gemm_bench.py --op marlin --ms 1,16 --layers qkv,o,down 2>/dev/null | grep "^{" | python -c "
import sys,json
for l in sys.stdin:
    r=json.loads(l); print('  ', r['layer'], r['m'], round(r['us'],2))"
done
