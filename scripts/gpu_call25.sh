# GPU call 25: W2 (1 000 voices, one voice per warp): tile loop of the specialised kernel unrolled 2x / 4x
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
W="w2:1000:441000:0 w2:10000:88064:0"
( python scripts/perf_probe4.py "t" $W
  DUSP_FUSED_UNROLL=2 python scripts/perf_probe4.py "t" $W
  DUSP_FUSED_UNROLL=4 python scripts/perf_probe4.py "t" $W
) > gpurun_out/c25_probe.txt 2>&1
grep -o '"workload": "[^"]*"\|"ms": [0-9.]*\|"block": [0-9]*\|"grid": [0-9]*\|Error.*' gpurun_out/c25_probe.txt | paste - - - -
