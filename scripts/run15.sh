timeout 900 python -m pytest tests/test_quant_gpu.py -m gpu -x -q -k "24" > gpurun_out/marlin24_unroll_pytest.log 2>&1
echo "pytest rc=$?"; tail -3 gpurun_out/marlin24_unroll_pytest.log
timeout 300 python benchmarks/gemm_bench.py --op marlin24 --ms 1,16,64 > gpurun_out/gemm_marlin24_r1j.jsonl 2> gpurun_out/gemm_marlin24_r1j.err
echo "gemm rc=$?"; grep "^{" gpurun_out/gemm_marlin24_r1j.jsonl | python -c "
import sys,json
for l in sys.stdin:
    r=json.loads(l); print(r['layer'], r['m'], round(r['us'],2), round(100*r['hbm_frac'],1))"
