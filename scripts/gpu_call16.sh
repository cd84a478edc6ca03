# GPU call 16: a circuit's words in shared memory for the multi-program launch: parity on the heterogeneous-batch tests, W4
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_gpu_events.py -m gpu -q -x -k "w4 or multi or control_ops or heterogeneous or event" > gpurun_out/c16_pytest.log 2>&1; tail -4 gpurun_out/c16_pytest.log
run() { echo "== $*"; env "$@" timeout 300 python scripts/perf_w4.py 10000 10 441088 2>&1 | tail -1 | grep -o '"ms": [0-9.]*\|"unit_samples_per_s": [0-9.e+]*\|"grid": [0-9]*\|"smem_bytes": [0-9]*\|Error.*' | paste - - - -; }
( run A=1
  run DUSP_MULTI_GLOBAL_WORDS=1
  run DUSP_MULTI_DENSITY=0
  run DUSP_MULTI_WPC=2
) > gpurun_out/c16_w4.txt 2>&1
cat gpurun_out/c16_w4.txt
