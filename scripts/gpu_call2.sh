# GPU call 2 of round 2: warp placement / XU throughput microbenchmarks, chain variants on clean data, S-warp placement
# and integer-widening experiments in situ, the full -m gpu suite through the dusp_run protocol
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
./scripts/ubench/place > gpurun_out/c2_place.txt 2>&1
./scripts/ubench/chain > gpurun_out/c2_chain.txt 2>&1
W=w3:10000:88064:0
( python scripts/perf_probe4.py "p24,DOPS_NO_SHIFT" $W
  python scripts/perf_probe4.py "p24" $W
  DUSP_PIPED_SLOT=1 python scripts/perf_probe4.py "p24" $W
  DUSP_PIPED_SLOT=2 python scripts/perf_probe4.py "p24" $W
  python scripts/perf_probe4.py "p24,DOPS_WIDEN_CHAIN" $W
  python scripts/perf_probe4.py "p24,DOPS_WIDEN_CHAIN,DOPS_WIDEN_FF" $W
  DUSP_PIPED_SLOT=1 python scripts/perf_probe4.py "p24,DOPS_WIDEN_CHAIN,DOPS_WIDEN_FF" $W
  DUSP_PIPED_SLOT=2 python scripts/perf_probe4.py "p24,DOPS_WIDEN_CHAIN,DOPS_WIDEN_FF" $W
) > gpurun_out/c2_probe.txt 2>&1
DUSP_FUSED_DEFS=DOPS_WIDEN_CHAIN,DOPS_WIDEN_FF timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -q -k "w3 or filter or simple_delay or band or worm or ap_web" > gpurun_out/c2_pytest_widen.log 2>&1; echo "widen pytest rc=$?" >> gpurun_out/c2_pytest_widen.log
tail -3 gpurun_out/c2_pytest_widen.log
timeout 1500 python -m pytest tests -m gpu -q > gpurun_out/c2_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/c2_pytest.log
tail -4 gpurun_out/c2_pytest.log; cat gpurun_out/c2_place.txt gpurun_out/c2_chain.txt; grep -o '"workload": "[^"]*"\|"ms": [0-9.]*' gpurun_out/c2_probe.txt | paste - -
