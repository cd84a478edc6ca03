# GPU call 23: quarter-step form of the pipelined kernel (DUSP_PIPED_FFP=2): parity on the pipelined subset, launch time
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
DUSP_PIPED_FFP=2 timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -q -k "pipelined" > gpurun_out/c23_pytest_ffq.log 2>&1; tail -4 gpurun_out/c23_pytest_ffq.log
W=w3:10000:88064:0
( python scripts/perf_probe4.py "p24" $W
  DUSP_PIPED_FFP=2 python scripts/perf_probe4.py "p24" $W w3:10000:88064:1
  DUSP_PIPED_FFP=2 DUSP_PIPED_MINB=2 python scripts/perf_probe4.py "p24" $W
  DUSP_PIPED_FFP=2 python scripts/perf_probe4.py "p28" $W
  DUSP_PIPED_FFP=2 python scripts/perf_probe4.py "p20" $W
) > gpurun_out/c23_probe.txt 2>&1
grep -o '"workload": "[^"]*"\|"ms": [0-9.]*\|"block": [0-9]*\|"grid": [0-9]*\|Error.*' gpurun_out/c23_probe.txt | paste - - - -
