# GPU call 31: Ramp's division by a shared, correctly rounded reciprocal (ops.cuh SharedDiv): full -m gpu suite (every
# Ramp case is held bit-exact), then W2 / W4
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -x > gpurun_out/c31_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/c31_pytest.log; tail -4 gpurun_out/c31_pytest.log
python scripts/perf_probe4.py "t" w2:1000:441000:0 w2:100:441000:0 w2:10000:88064:0 2>&1 | grep -o '"workload": "[^"]*"\|"ms": [0-9.]*\|"grid": [0-9]*' | paste - - -
DUSP_NO_SEGMENTS=1 python scripts/perf_probe4.py "t" w2:1000:441000:0 2>&1 | grep -o '"workload": "[^"]*"\|"ms": [0-9.]*\|"grid": [0-9]*' | paste - - -
timeout 300 python scripts/perf_w4.py 10000 10 441088 2>&1 | tail -1 | grep -o '"ms": [0-9.]*\|"unit_samples_per_s": [0-9.e+]*' | paste - -
