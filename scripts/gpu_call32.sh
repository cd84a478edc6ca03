# GPU call 32: constant-coefficient Filter on T lanes: every lane runs every turn (selects instead of a divergent branch).
# Full -m gpu suite, then W4 and the feedback-loop batch (W3b, T = 4 / 8 tile kernels)
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -x > gpurun_out/c32_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/c32_pytest.log; tail -4 gpurun_out/c32_pytest.log
timeout 300 python scripts/perf_w4.py 10000 10 441088 2>&1 | tail -1 | grep -o '"ms": [0-9.]*\|"unit_samples_per_s": [0-9.e+]*' | paste - -
timeout 300 python scripts/perf_w4.py 2500 10 441088 2>&1 | tail -1 | grep -o '"ms": [0-9.]*\|"unit_samples_per_s": [0-9.e+]*' | paste - -
python scripts/perf_probe4.py "t" w3b:1000:441000:0 w3b:10000:88064:0 2>&1 | grep -o '"workload": "[^"]*"\|"ms": [0-9.]*\|"grid": [0-9]*' | paste - - -
