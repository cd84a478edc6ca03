# GPU call 21: pairs of independent constant-coefficient Filters as one pass in the multi-program interpreter
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -q -k "w4 or multi or control_ops" > gpurun_out/c21_pytest.log 2>&1; tail -3 gpurun_out/c21_pytest.log
run() { echo "== $*"; n=$1; shift; env "$@" timeout 300 python scripts/perf_w4.py $n 10 441088 2>&1 | tail -1 | grep -o '"ms": [0-9.]*\|"unit_samples_per_s": [0-9.e+]*\|"smem_bytes": [0-9]*\|Error.*' | paste - - -; }
( run 10000 A=1
  run 2500 A=1
  run 10000 DUSP_MULTI_NO_PAIRS=1
  run 10000 DUSP_MULTI_NO_PAIRS=1 DUSP_NO_FILTER_GROUPS=1
  run 2500 DUSP_MULTI_NO_PAIRS=1 DUSP_NO_FILTER_GROUPS=1
  run 10000 DUSP_NO_FILTER_GROUPS=1
) > gpurun_out/c21_w4.txt 2>&1
cat gpurun_out/c21_w4.txt
