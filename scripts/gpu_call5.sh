# GPU call 5: ncu --set full captures of dusp_piped (W3, 10 000 voices x 88 064 samples), without and with the S/F split
cd "$(dirname "$0")/.."
mkdir -p gpurun_out/c5_a gpurun_out/c5_b
W=w3:10000:88064:0
DUSP_FUSED_DUMP=gpurun_out/c5_a ncu --set full --import-source on --clock-control none -k regex:dusp_piped -s 1 -c 1 -f -o gpurun_out/c5_nosplit python scripts/perf_probe4.py "p24" $W > gpurun_out/c5_ncu_a.log 2>&1
DUSP_PIPED_SPLIT=1 DUSP_FUSED_DUMP=gpurun_out/c5_b ncu --set full --import-source on --clock-control none -k regex:dusp_piped -s 1 -c 1 -f -o gpurun_out/c5_split python scripts/perf_probe4.py "p24" $W > gpurun_out/c5_ncu_b.log 2>&1
ls -la gpurun_out/*.ncu-rep gpurun_out/c5_a gpurun_out/c5_b
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_golden.py -m gpu -q -k "shape_then or attack_then or full_size" > gpurun_out/c5_pytest.log 2>&1; tail -3 gpurun_out/c5_pytest.log
