# GPU call 36 (2 GPUs): the multi-GPU paths on the end-of-round build: 2-GPU mix parity test, bench.py at N = 2
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -q -k "two_gpus" > gpurun_out/c36_pytest_2gpu.log 2>&1; tail -3 gpurun_out/c36_pytest_2gpu.log
timeout 1200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 3 --warmup 3 > gpurun_out/c36_bench_n2.json 2> gpurun_out/c36_bench_n2.err; echo "bench rc=$?"; tail -c 600 gpurun_out/c36_bench_n2.json | head -c 600
