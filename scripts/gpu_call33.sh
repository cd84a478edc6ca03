# GPU call 33: end-of-round lines: bench.py at N = 1 (both arms) and the launch list of the same command
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 900 python bench.py --impl reference --steps 3 --warmup 3 > gpurun_out/c33_bench_reference.json 2> gpurun_out/c33_bench_reference.err; echo "reference rc=$?"
timeout 1500 python bench.py --steps 3 --warmup 3 > gpurun_out/c33_bench_n1.json 2> gpurun_out/c33_bench_n1.err; echo "bench rc=$?"; tail -2 gpurun_out/c33_bench_n1.err
timeout 1200 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/c33_launches.csv python bench.py --steps 3 --warmup 3 --no-cpu > gpurun_out/c33_ncu_bench.log 2>&1; echo "ncu rc=$?"
