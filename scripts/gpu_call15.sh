# GPU call 15: cooperative recurrences in the multi-program interpreter (constant-coefficient Filters of the 8 circuits of
# a CTA served one circuit per lane): parity on the heterogeneous-batch tests, then W4
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
DUSP_MULTI_COOP=1 timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -q -k "w4 or multi_program or control_ops" > gpurun_out/c15_pytest_coop.log 2>&1; tail -5 gpurun_out/c15_pytest_coop.log
run() { echo "== $*"; env "$@" timeout 300 python scripts/perf_w4.py 10000 10 441088 2>&1 | tail -1 | grep -o '"ms": [0-9.]*\|"unit_samples_per_s": [0-9.e+]*\|"grid": [0-9]*\|"smem_bytes": [0-9]*\|Error.*' | paste - - - -; }
( run DUSP_MULTI_COOP=1
  run DUSP_MULTI_COOP=0
) > gpurun_out/c15_w4.txt 2>&1
cat gpurun_out/c15_w4.txt
