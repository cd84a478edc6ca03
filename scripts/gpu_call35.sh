# GPU call 35: W4 against the interpreter's register cap again (64 / 72 / 80 / uncapped) now that the launch's floor counts
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
run() { echo "== $*"; n=$1; shift; env "$@" timeout 300 python scripts/perf_w4.py $n 10 441088 2>&1 | tail -1 | grep -o '"ms": [0-9.]*\|"unit_samples_per_s": [0-9.e+]*' | paste - -; }
( run 10000 DUSP_MULTI_DENSITY=8
  run 10000 DUSP_MULTI_DENSITY=7
  run 10000 DUSP_MULTI_DENSITY=6
  run 10000 DUSP_MULTI_DENSITY=0
  run 2500 DUSP_MULTI_DENSITY=8
  run 2500 DUSP_MULTI_DENSITY=6
  run 2500 DUSP_MULTI_DENSITY=0
) 2>&1 | tee gpurun_out/c35_w4.txt
