# GPU call 20: short rings in rounds (c_fixdelay); full -m gpu suite; W4 at 2 500 / 10 000 circuits; ticks per circuit
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -x > gpurun_out/c20_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/c20_pytest.log; tail -4 gpurun_out/c20_pytest.log
for n in 2500 10000; do timeout 600 python scripts/perf_w4.py $n 10 441088 2>&1 | tail -1 | grep -o '"workload": "[^"]*"\|"ms": [0-9.]*\|"unit_samples_per_s": [0-9.e+]*' | paste - - -; done | tee gpurun_out/c20_w4.txt
python scripts/w4_slowest.py rank 2500 > gpurun_out/c20_rank.txt 2>&1; head -8 gpurun_out/c20_rank.txt | cut -c1-220
