# GPU call 27: end-of-round capture of the default pipelined kernel (same build as the bench): ncu --set full, roles
cd "$(dirname "$0")/.."
mkdir -p gpurun_out/c27_dump
W=w3:10000:88064:0
python scripts/perf_probe4.py "p24" $W 2>&1 | grep -o '"ms": [0-9.]*'
DUSP_FUSED_DUMP=gpurun_out/c27_dump ncu --set full --import-source on --clock-control none -k regex:dusp_piped -s 1 -c 1 -f -o gpurun_out/c27_piped python scripts/perf_probe4.py "p24" $W > gpurun_out/c27_ncu.log 2>&1
ls gpurun_out/c27_dump gpurun_out/c27_piped.ncu-rep
