timeout 900 python -m pytest tests/test_quant_gpu.py tests/test_glue_gpu.py -m gpu -x -q -k "marlin or pdl" > gpurun_out/marlin_unroll_pytest.log 2>&1
echo "pytest rc=$?"; tail -3 gpurun_out/marlin_unroll_pytest.log
timeout 300 python benchmarks/gemm_bench.py --op marlin --ms 1,16,64 > gpurun_out/gemm_marlin_r1j.jsonl 2> gpurun_out/gemm_marlin_r1j.err
echo "gemm rc=$?"; grep "^{" gpurun_out/gemm_marlin_r1j.jsonl | python -c "
import sys,json
for l in sys.stdin:
    r=json.loads(l); print(r['layer'], r['m'], round(r['us'],2), round(100*r['hbm_frac'],1))"
timeout 300 python benchmarks/decode_bench.py --batches 1,8,64,256 --steps 30 > gpurun_out/decode_tp1_r1j.jsonl 2> gpurun_out/decode_tp1_r1j.err
echo "decode rc=$?"; grep "^{" gpurun_out/decode_tp1_r1j.jsonl | cut -c80-200
