# GPU call 17: is W4 bound by throughput or by its longest circuits?  5 000 / 10 000 / 20 000 circuits, same 10 s
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
for n in 2500 5000 10000 20000; do timeout 600 python scripts/perf_w4.py $n 10 2>&1 | tail -1 | grep -o '"workload": "[^"]*"\|"ms": [0-9.]*\|"unit_samples_per_s": [0-9.e+]*' | paste - - -; done | tee gpurun_out/c17_w4_scaling.txt
