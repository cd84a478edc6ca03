# GPU call 22: round-2 end state: bench.py at N = 1 (both arms), the launch list of the same command, smoke()
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/c22_smoke.log 2>&1; tail -2 gpurun_out/c22_smoke.log
timeout 900 python bench.py --impl reference --steps 3 --warmup 3 > gpurun_out/c22_bench_reference.json 2> gpurun_out/c22_bench_reference.err; echo "reference rc=$?"
timeout 1500 python bench.py --steps 3 --warmup 3 > gpurun_out/c22_bench_n1.json 2> gpurun_out/c22_bench_n1.err; echo "bench rc=$?"; tail -2 gpurun_out/c22_bench_n1.err
timeout 1200 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/c22_launches.csv python bench.py --steps 3 --warmup 3 --no-cpu --no-workloads > gpurun_out/c22_ncu_bench.log 2>&1; echo "ncu rc=$?"
ls -la gpurun_out/c22*
