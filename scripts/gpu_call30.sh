# GPU call 30: segmented rendering test, then the full -m gpu suite and the bench at N = 1 on the end-of-round build
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -q -k "small_batch" > gpurun_out/c30_pytest_seg.log 2>&1; tail -5 gpurun_out/c30_pytest_seg.log
timeout 1500 python -m pytest tests -m gpu -q -x > gpurun_out/c30_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/c30_pytest.log; tail -4 gpurun_out/c30_pytest.log
python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -1
timeout 1500 python bench.py --steps 3 --warmup 3 > gpurun_out/c30_bench_n1.json 2> gpurun_out/c30_bench_n1.err; echo "bench rc=$?"; tail -2 gpurun_out/c30_bench_n1.err
