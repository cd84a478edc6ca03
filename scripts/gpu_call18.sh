# GPU call 18: runs of independent constant-coefficient Filters as one pass (a recurrence per lane) in the multi-program
# interpreter: parity on the heterogeneous-batch tests, then W4 with 4 / 2 / 0 scratch rows
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -q -k "w4 or multi or control_ops" > gpurun_out/c18_pytest.log 2>&1; tail -4 gpurun_out/c18_pytest.log
run() { echo "== $*"; env "$@" timeout 300 python scripts/perf_w4.py 10000 10 441088 2>&1 | tail -1 | grep -o '"ms": [0-9.]*\|"unit_samples_per_s": [0-9.e+]*\|"grid": [0-9]*\|"smem_bytes": [0-9]*\|Error.*' | paste - - - -; }
( run A=1
  run DUSP_MULTI_GROUP_ROWS=2
  run DUSP_MULTI_GROUP_ROWS=1
  run DUSP_MULTI_GROUP_ROWS=0
  run DUSP_MULTI_GROUP_ROWS=0 DUSP_NO_FILTER_GROUPS=1
) > gpurun_out/c18_w4.txt 2>&1
cat gpurun_out/c18_w4.txt
timeout 300 python scripts/perf_w4.py 2500 10 441088 2>&1 | tail -1 | grep -o '"ms": [0-9.]*'
