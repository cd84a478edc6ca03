# GPU call 39: ncu of the W2 kernels (segmented render + scan), 1 000 voices x 10 s
cd "$(dirname "$0")/.."
mkdir -p gpurun_out/c39_dump
DUSP_FUSED_DUMP=gpurun_out/c39_dump ncu --set full --import-source on --clock-control none -k regex:dusp_fused -s 2 -c 2 -f -o gpurun_out/c39_w2 python scripts/perf_probe4.py "t" w2:1000:441000:0 > gpurun_out/c39_ncu.log 2>&1
ls gpurun_out/c39_dump gpurun_out/c39_w2.ncu-rep
