N=${1:-2}
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus $N --steps 50 --warmup 5 > gpurun_out/bench_n${N}.json 2> gpurun_out/bench_n${N}.err
echo "bench rc=$?"; tail -c 700 gpurun_out/bench_n${N}.json
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29513 bench.py --impl reference --gpus $N --steps 3 --warmup 1 > gpurun_out/bench_ref_n${N}.json 2> gpurun_out/bench_ref_n${N}.err
echo "ref rc=$?"
timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 benchmarks/decode_bench.py --batches 1,8,64,256 --steps 30 > gpurun_out/decode_tp${N}_r1.jsonl 2> gpurun_out/decode_tp${N}_r1.err
echo "decode rc=$?"; grep "^{" gpurun_out/decode_tp${N}_r1.jsonl | cut -c1-240
