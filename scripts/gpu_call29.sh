# GPU call 29: segmented rendering (a voice's timeline on several warps) for small batches of jumpable circuits:
# parity of the specialised mode over every case, then W2 (1 000 voices) and the bench's sub-records
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_gpu_parity.py -m gpu -q -x -k "specialised" > gpurun_out/c29_pytest.log 2>&1; echo "rc=$?" >> gpurun_out/c29_pytest.log; tail -12 gpurun_out/c29_pytest.log
( python scripts/perf_probe4.py "t" w2:1000:441000:0 w2:100:441000:0 w2:3000:441000:0
  DUSP_NO_SEGMENTS=1 python scripts/perf_probe4.py "t" w2:1000:441000:0 w2:100:441000:0
  DUSP_SEGMENTS=8 python scripts/perf_probe4.py "t" w2:1000:441000:0
  DUSP_SEGMENTS=2 python scripts/perf_probe4.py "t" w2:1000:441000:0
) > gpurun_out/c29_probe.txt 2>&1
grep -o '"workload": "[^"]*"\|"ms": [0-9.]*\|"grid": [0-9]*\|Error.*' gpurun_out/c29_probe.txt | paste - - -
