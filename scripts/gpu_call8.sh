# GPU call 8: S/F split with shared-memory counters instead of named barriers; register caps; ring depth; parity
cd "$(dirname "$0")/.."
mkdir -p gpurun_out/c8_dump
W=w3:10000:88064:0
( python scripts/perf_probe4.py "p24" $W
  DUSP_PIPED_SPLIT=1 python scripts/perf_probe4.py "p24" $W
  DUSP_PIPED_SPLIT=1 DUSP_PIPED_MAXNREG=112 python scripts/perf_probe4.py "p24" $W
  DUSP_PIPED_SPLIT=1 DUSP_PIPED_MAXNREG=112 DUSP_PIPED_RING=8 python scripts/perf_probe4.py "p24" $W
  DUSP_PIPED_SPLIT=1 DUSP_PIPED_MAXNREG=112 DUSP_PIPED_MAGIC=0 python scripts/perf_probe4.py "p24" $W
  DUSP_PIPED_SPLIT=1 DUSP_PIPED_MAXNREG=112 python scripts/perf_probe4.py "p24,DOPS_NO_SHIFT" $W
  DUSP_PIPED_MAXNREG=136 python scripts/perf_probe4.py "p24" $W
) > gpurun_out/c8_probe.txt 2>&1
grep -o '"workload": "[^"]*"\|"ms": [0-9.]*\|Error.*' gpurun_out/c8_probe.txt | paste - -
DUSP_PIPED_SPLIT=1 DUSP_PIPED_MAXNREG=112 timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -q -x -k "pipelined or full_size or sixty" > gpurun_out/c8_pytest_split.log 2>&1; echo "split pytest rc=$?" >> gpurun_out/c8_pytest_split.log; tail -3 gpurun_out/c8_pytest_split.log
DUSP_PIPED_SPLIT=1 DUSP_PIPED_MAXNREG=112 DUSP_FUSED_DUMP=gpurun_out/c8_dump ncu --set full --import-source on --clock-control none -k regex:dusp_piped -s 1 -c 1 -f -o gpurun_out/c8_split python scripts/perf_probe4.py "p24" $W > gpurun_out/c8_ncu.log 2>&1
ls gpurun_out/c8_dump gpurun_out/c8_split.ncu-rep
