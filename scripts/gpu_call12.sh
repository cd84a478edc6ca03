# GPU call 12: ffp form with more P-warps; ncu of the ffp form (who waits now) and of the interpreter on W4
cd "$(dirname "$0")/.."
mkdir -p gpurun_out/c12_dump
W=w3:10000:88064:0
( DUSP_PIPED_FFP=1 DUSP_PIPED_PWARPS=6 DUSP_PIPED_MINB=3 python scripts/perf_probe4.py "p24" $W
  DUSP_PIPED_FFP=1 DUSP_PIPED_PWARPS=8 DUSP_PIPED_MINB=3 python scripts/perf_probe4.py "p24" $W
  DUSP_PIPED_FFP=1 DUSP_PIPED_PWARPS=6 python scripts/perf_probe4.py "p24" $W
  DUSP_PIPED_FFP=1 DUSP_PIPED_MINB=3 python scripts/perf_probe4.py "p24" $W
  DUSP_PIPED_PWARPS=6 python scripts/perf_probe4.py "p24" $W
) > gpurun_out/c12_probe.txt 2>&1
grep -o '"workload": "[^"]*"\|"ms": [0-9.]*\|"block": [0-9]*\|"grid": [0-9]*\|Error.*' gpurun_out/c12_probe.txt | paste - - - -
DUSP_PIPED_FFP=1 DUSP_FUSED_DUMP=gpurun_out/c12_dump ncu --set full --import-source on --clock-control none -k regex:dusp_piped -s 1 -c 1 -f -o gpurun_out/c12_ffp python scripts/perf_probe4.py "p24" $W > gpurun_out/c12_ncu_ffp.log 2>&1
python scripts/perf_w4.py 2000 2 > gpurun_out/c12_w4_2000.txt 2>&1; tail -1 gpurun_out/c12_w4_2000.txt | cut -c1-300
ncu --set full --import-source on --clock-control none -k regex:dusp_render_tile -s 2 -c 1 -f -o gpurun_out/c12_w4 python scripts/perf_w4.py 2000 2 > gpurun_out/c12_ncu_w4.log 2>&1
ls -la gpurun_out/*.ncu-rep gpurun_out/c12_dump
