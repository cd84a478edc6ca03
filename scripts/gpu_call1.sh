# GPU call 1 of round 2: smoke, the -m gpu suite with the new cases, chain-latency microbenchmarks, baseline W3 probe
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.max.sm --format=csv > gpurun_out/c1_gpu.txt
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/c1_smoke.log 2>&1; echo "smoke rc=$?" >> gpurun_out/c1_smoke.log
./scripts/ubench/lat > gpurun_out/c1_lat.txt 2>&1
./scripts/ubench/chain > gpurun_out/c1_chain.txt 2>&1
python scripts/perf_probe4.py "p24" w3:10000:88064:0 > gpurun_out/c1_probe.txt 2>&1
python scripts/exp_tighten.py > gpurun_out/c1_tighten.txt 2>&1
timeout 1200 python -m pytest tests -m gpu -q -x > gpurun_out/c1_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/c1_pytest.log
tail -3 gpurun_out/c1_pytest.log; cat gpurun_out/c1_smoke.log | tail -2; cat gpurun_out/c1_chain.txt gpurun_out/c1_lat.txt; tail -2 gpurun_out/c1_probe.txt
