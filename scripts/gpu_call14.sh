# GPU call 14: W4 in one launch per pass; ncu of the interpreter on W4 after the evaluation-order / density changes
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
python scripts/perf_w4.py 10000 10 441088 2>&1 | tail -1 | cut -c1-420
python scripts/perf_w4.py 10000 10 2>&1 | tail -1 | cut -c1-420
python scripts/perf_w4.py 2000 2 > gpurun_out/c14_w4_2000.txt 2>&1; tail -1 gpurun_out/c14_w4_2000.txt | cut -c1-300
ncu --set full --import-source on --clock-control none -k regex:dusp_render_tile -s 2 -c 1 -f -o gpurun_out/c14_w4 python scripts/perf_w4.py 2000 2 > gpurun_out/c14_ncu_w4.log 2>&1
ls -la gpurun_out/c14*
