for s in 1 2 4 8 16; do
  echo "splits=$s"
  B200_MARLIN_SPLITS=$s timeout 120 python benchmarks/gemm_bench.py --op marlin --ms 1,16 --layers o,qkv,down 2>/dev/null | python -c "
import sys,json
for l in sys.stdin:
    r=json.loads(l); print(' ',r['layer'],r['m'],round(r['us'],1))"
done
