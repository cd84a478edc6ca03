# GPU call 11: "ffp" form of the pipelined kernel (feed-forward half in the P-warps' stage 0, f64 brackets in the exchange
# rows, S-warp = recurrence only): parity on the pipelined subset, then the launch time against the default
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
DUSP_PIPED_FFP=1 timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -q -k "pipelined" > gpurun_out/c11_pytest_ffp.log 2>&1; tail -4 gpurun_out/c11_pytest_ffp.log
W=w3:10000:88064:0
( python scripts/perf_probe4.py "p24" $W
  DUSP_PIPED_FFP=1 python scripts/perf_probe4.py "p24" $W
  DUSP_PIPED_FFP=1 DUSP_PIPED_MAGIC=1 python scripts/perf_probe4.py "p24" $W
  DUSP_PIPED_FFP=1 python scripts/perf_probe4.py "p16" $W
  DUSP_PIPED_FFP=1 python scripts/perf_probe4.py "p32" $W
  python scripts/perf_probe4.py "t" w2:1000:441000:0
) > gpurun_out/c11_probe.txt 2>&1
grep -o '"workload": "[^"]*"\|"ms": [0-9.]*\|"block": [0-9]*\|"grid": [0-9]*\|Error.*' gpurun_out/c11_probe.txt | paste - - - -
python scripts/perf_w4.py > gpurun_out/c11_w4.txt 2>&1; tail -2 gpurun_out/c11_w4.txt | cut -c1-400
