# GPU call 13: depth-first evaluation order (W4 circuits: 28 -> 5 slots), register-capped interpreter instantiations for
# heterogeneous batches, longest-circuit-first descriptors, instruction prefetch.  Full -m gpu suite, then W4 variants.
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -x > gpurun_out/c13_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/c13_pytest.log; tail -4 gpurun_out/c13_pytest.log
run() { echo "== $*"; env "$@" python scripts/perf_w4.py 2>&1 | tail -1 | grep -o '"ms": [0-9.]*\|"unit_samples_per_s": [0-9.e+]*\|"grid": [0-9]*\|"smem_bytes": [0-9]*\|Error.*' | paste - - - -; }
( run DUSP_MULTI_DENSITY=6
  run DUSP_MULTI_DENSITY=8
  run DUSP_MULTI_DENSITY=0
  run DUSP_MULTI_DENSITY=6 DUSP_MULTI_B=8
  run DUSP_MULTI_DENSITY=8 DUSP_MULTI_B=8
  run DUSP_MULTI_DENSITY=6 DUSP_MULTI_UNSORTED=1
  run DUSP_MULTI_DENSITY=6 DUSP_KEEP_TICK_ORDER=1
  run DUSP_MULTI_DENSITY=6 DUSP_MULTI_WPC=4
) > gpurun_out/c13_w4.txt 2>&1
cat gpurun_out/c13_w4.txt
python scripts/perf_probe4.py "t" w2:1000:441000:0 w3b:1000:441000:0 2>&1 | grep -o '"workload": "[^"]*"\|"ms": [0-9.]*' | paste - -
