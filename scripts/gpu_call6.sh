# GPU call 6: full -m gpu suite with the a11 / Noise / PickChannel cases and the two-phase prefetched ring reads;
# W3 timing; ncu --set full capture of dusp_piped (W3, 10 000 voices x 88 064 samples) + launch list of bench.py
cd "$(dirname "$0")/.."
mkdir -p gpurun_out/c6_dump
timeout 1500 python -m pytest tests -m gpu -q > gpurun_out/c6_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/c6_pytest.log
tail -6 gpurun_out/c6_pytest.log
W=w3:10000:88064:0
( python scripts/perf_probe4.py "p24" $W w3:10000:88064:1
  python scripts/perf_probe4.py "p24,DOPS_NO_SHIFT" $W
  DUSP_PIPED_SPLIT=1 python scripts/perf_probe4.py "p24" $W
) > gpurun_out/c6_probe.txt 2>&1
grep -o '"workload": "[^"]*"\|"ms": [0-9.]*\|"mode": [0-9]\|Error.*' gpurun_out/c6_probe.txt | paste - - -
DUSP_FUSED_DUMP=gpurun_out/c6_dump ncu --set full --import-source on --clock-control none -k regex:dusp_piped -s 1 -c 1 -f -o gpurun_out/c6_piped python scripts/perf_probe4.py "p24" $W > gpurun_out/c6_ncu.log 2>&1
ls -la gpurun_out/*.ncu-rep gpurun_out/c6_dump
