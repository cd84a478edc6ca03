# GPU call 24: quarter-step form with stage 2's loads issued a quarter ahead: parity on the pipelined subset, launch time
cd "$(dirname "$0")/.."
mkdir -p gpurun_out/c24_dump
DUSP_PIPED_FFP=2 timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -q -k "pipelined" > gpurun_out/c24_pytest_ffq.log 2>&1; tail -4 gpurun_out/c24_pytest_ffq.log
W=w3:10000:88064:0
( python scripts/perf_probe4.py "p24" $W
  DUSP_PIPED_FFP=2 python scripts/perf_probe4.py "p24" $W w3:10000:88064:1
) > gpurun_out/c24_probe.txt 2>&1
grep -o '"workload": "[^"]*"\|"ms": [0-9.]*\|"block": [0-9]*\|"grid": [0-9]*\|Error.*' gpurun_out/c24_probe.txt | paste - - - -
DUSP_PIPED_FFP=2 DUSP_FUSED_DUMP=gpurun_out/c24_dump ncu --set full --import-source on --clock-control none -k regex:dusp_piped -s 1 -c 1 -f -o gpurun_out/c24_ffq python scripts/perf_probe4.py "p24" $W > gpurun_out/c24_ncu.log 2>&1
ls gpurun_out/c24_dump gpurun_out/c24_ffq.ncu-rep
