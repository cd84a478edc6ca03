# GPU call 4: S/F split of the pipelined kernel (two-step rotation, named-barrier ring), parity + timing; bench.py dry run
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -q -x -k "pipelined or full_size or misaligned or sixty" > gpurun_out/c4_pytest_piped.log 2>&1; echo "piped pytest rc=$?" >> gpurun_out/c4_pytest_piped.log
tail -4 gpurun_out/c4_pytest_piped.log
W=w3:10000:88064:0
( python scripts/perf_probe4.py "p24" $W w3:10000:88064:1
  DUSP_PIPED_MAGIC=0 python scripts/perf_probe4.py "p24" $W
  DUSP_PIPED_SPLIT=0 python scripts/perf_probe4.py "p24" $W
  python scripts/perf_probe4.py "p32;p16" $W
  python scripts/perf_probe4.py "p24,DOPS_NO_SHIFT" $W
) > gpurun_out/c4_probe.txt 2>&1
grep -o '"workload": "[^"]*"\|"ms": [0-9.]*\|"mode": [0-9]\|Error.*' gpurun_out/c4_probe.txt | paste - - -
timeout 1200 python bench.py --steps 2 --warmup 3 > gpurun_out/c4_bench.json 2> gpurun_out/c4_bench.err; echo "bench rc=$?"; tail -3 gpurun_out/c4_bench.err
python - <<'PY'
import json
try:
    d = json.loads(open('gpurun_out/c4_bench.json').read().strip().splitlines()[-1])
    print({k: d[k] for k in ('value', 'ms_per_step')}, d['roofline']['frac'], d['roofline']['avg_launch_ms'], d['e2e'], [ (w['workload'][:10], w['value']) for w in d['workloads']], d['mixdown']['value'], d['mixdown']['e2e_mix'])
except Exception as e:
    print('bench parse failed', e)
PY
