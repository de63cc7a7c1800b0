N=${1:-2}
timeout 600 python -m pytest tests/test_quant_gpu.py -m gpu -x -q -k "24" > gpurun_out/marlin24_unroll_pytest.log 2>&1
echo "pytest rc=$?"; tail -2 gpurun_out/marlin24_unroll_pytest.log
timeout 300 python benchmarks/gemm_bench.py --op marlin24 --ms 1,16,64 > gpurun_out/gemm_marlin24_r1j.jsonl 2> gpurun_out/gemm_marlin24_r1j.err
echo "gemm rc=$?"; grep "^{" gpurun_out/gemm_marlin24_r1j.jsonl | python -c "
import sys,json
for l in sys.stdin:
    r=json.loads(l); print(r['layer'], r['m'], round(r['us'],2), round(100*r['hbm_frac'],1))"
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29541 benchmarks/ar_check.py > gpurun_out/ar_check_n${N}.jsonl 2> gpurun_out/ar_check_n${N}.err
echo "ar_check rc=$?"; grep -v "^\*\|OMP\|^$" gpurun_out/ar_check_n${N}.jsonl | tail -12; tail -3 gpurun_out/ar_check_n${N}.err | cut -c1-300
timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29542 benchmarks/decode_bench.py --batches 1,8,64,256 --steps 30 > gpurun_out/decode_tp${N}_r1h_fusednorm.jsonl 2> gpurun_out/decode_tp${N}_r1h_fusednorm.err
echo "decode fused-norm rc=$?"; grep "^{" gpurun_out/decode_tp${N}_r1h_fusednorm.jsonl | cut -c80-215
timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29543 benchmarks/decode_bench.py --batches 1,8,64,256 --steps 30 --no-fuse-ar-norm > gpurun_out/decode_tp${N}_r1g_oneshot.jsonl 2> gpurun_out/decode_tp${N}_r1g_oneshot.err
echo "decode one_shot rc=$?"; grep "^{" gpurun_out/decode_tp${N}_r1g_oneshot.jsonl | cut -c80-215
