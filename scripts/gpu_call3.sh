# GPU call 3: 32-bit tick generator + per-mode kernels: parity suite, W3 / W2 probes, magic-rounding recurrence in situ
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
W=w3:10000:88064:0
( python scripts/perf_probe4.py "p24" $W w3:10000:88064:1
  python scripts/perf_probe4.py "p24,DOPS_MAGIC" $W
  python scripts/perf_probe4.py "p24,DOPS_SHIFT" $W
  python scripts/perf_probe4.py "p24,DOPS_MAGIC,DOPS_SHIFT" $W
  python scripts/perf_probe4.py "t" w2:1000:441000:0 w2:10000:88064:0 w3b:1000:441000:0
) > gpurun_out/c3_probe.txt 2>&1
grep -o '"workload": "[^"]*"\|"ms": [0-9.]*\|"mode": [0-9]\|Error.*' gpurun_out/c3_probe.txt | paste - - -
DUSP_FUSED_DEFS=DOPS_MAGIC timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -q -k "w3 or filter or simple_delay or band or worm or ap_web or misaligned" > gpurun_out/c3_pytest_magic.log 2>&1; echo "magic pytest rc=$?" >> gpurun_out/c3_pytest_magic.log
tail -3 gpurun_out/c3_pytest_magic.log
timeout 1500 python -m pytest tests -m gpu -q > gpurun_out/c3_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/c3_pytest.log
tail -4 gpurun_out/c3_pytest.log
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
