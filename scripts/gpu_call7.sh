# GPU call 7: the no-alloc mix push (alone in a fresh process, then after the whole suite); S/F split with a register cap
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
python -m pytest tests/test_zz_no_alloc.py -m gpu -q > gpurun_out/c7_noalloc_alone.log 2>&1; echo "alone rc=$?" >> gpurun_out/c7_noalloc_alone.log; tail -3 gpurun_out/c7_noalloc_alone.log
W=w3:10000:88064:0
( DUSP_PIPED_SPLIT=1 DUSP_PIPED_MAXNREG=112 python scripts/perf_probe4.py "p24" $W
  DUSP_PIPED_SPLIT=1 DUSP_PIPED_MAXNREG=104 python scripts/perf_probe4.py "p24" $W
  DUSP_PIPED_SPLIT=1 DUSP_PIPED_MAXNREG=112 DUSP_PIPED_MAGIC=0 python scripts/perf_probe4.py "p24" $W
  DUSP_PIPED_MAXNREG=136 python scripts/perf_probe4.py "p24" $W
) > gpurun_out/c7_probe.txt 2>&1
grep -o '"workload": "[^"]*"\|"ms": [0-9.]*\|Error.*' gpurun_out/c7_probe.txt | paste - -
timeout 1500 python -m pytest tests -m gpu -q -x > gpurun_out/c7_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/c7_pytest.log; tail -4 gpurun_out/c7_pytest.log
