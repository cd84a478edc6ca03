# GPU call 9: full -m gpu suite on the interpreter that now calls the ops.cuh cores; FP64-domain rounding in the unsplit
# S-warp; bench.py at N = 1
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -x > gpurun_out/c9_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/c9_pytest.log; tail -5 gpurun_out/c9_pytest.log
W=w3:10000:88064:0
( python scripts/perf_probe4.py "p24" $W
  python scripts/perf_probe4.py "p24,DOPS_MAGIC" $W
  python scripts/perf_probe4.py "t" w2:1000:441000:0 w3b:1000:441000:0
) > gpurun_out/c9_probe.txt 2>&1
grep -o '"workload": "[^"]*"\|"ms": [0-9.]*\|Error.*' gpurun_out/c9_probe.txt | paste - -
DUSP_FUSED_DEFS=DOPS_MAGIC timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -q -k "pipelined and (w3 or filter or simple_delay or band or worm)" > gpurun_out/c9_pytest_magic.log 2>&1; tail -2 gpurun_out/c9_pytest_magic.log
timeout 1200 python bench.py --steps 3 --warmup 3 > gpurun_out/c9_bench.json 2> gpurun_out/c9_bench.err; echo "bench rc=$?"; tail -3 gpurun_out/c9_bench.err
python scripts/perf_w4.py > gpurun_out/c9_w4.txt 2>&1; tail -3 gpurun_out/c9_w4.txt
