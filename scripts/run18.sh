N=${1:-2}
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus $N --steps 50 --warmup 5 > gpurun_out/bench_n${N}.json 2> gpurun_out/bench_n${N}.err
echo "bench rc=$?"; tail -c 900 gpurun_out/bench_n${N}.json | cut -c1-900
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29513 bench.py --impl reference --gpus $N --steps 3 --warmup 1 > gpurun_out/bench_ref_n${N}.json 2> gpurun_out/bench_ref_n${N}.err
echo "ref rc=$?"; tail -c 300 gpurun_out/bench_ref_n${N}.json
timeout 300 python -m pytest tests/test_custom_ar.py -m gpu -q > gpurun_out/custom_ar_pytest.log 2>&1; echo "pytest custom_ar rc=$?"; tail -2 gpurun_out/custom_ar_pytest.log
