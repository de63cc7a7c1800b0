N=${1:-2}
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29541 benchmarks/ar_check.py > gpurun_out/ar_check_n${N}.jsonl 2> gpurun_out/ar_check_n${N}.err
echo "ar_check rc=$?"; cat gpurun_out/ar_check_n${N}.jsonl | tail -8; tail -5 gpurun_out/ar_check_n${N}.err
timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29542 benchmarks/decode_bench.py --batches 1,8,64,256 --steps 30 > gpurun_out/decode_tp${N}_r1g_oneshot.jsonl 2> gpurun_out/decode_tp${N}_r1g_oneshot.err
echo "decode one_shot rc=$?"; grep "^{" gpurun_out/decode_tp${N}_r1g_oneshot.jsonl | cut -c1-210
timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29543 benchmarks/decode_bench.py --batches 1,8,64,256 --steps 30 --ar nccl > gpurun_out/decode_tp${N}_r1g_nccl.jsonl 2> gpurun_out/decode_tp${N}_r1g_nccl.err
echo "decode nccl rc=$?"; grep "^{" gpurun_out/decode_tp${N}_r1g_nccl.jsonl | cut -c1-210
