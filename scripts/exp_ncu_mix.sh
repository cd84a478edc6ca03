# ncu captures of the mixdown path's kernels (DUSP_OUT_MIX epilogues with vector reductions; the one-sided push).
cd "$(dirname "$0")/.."
DUSP_PIPED=0 ncu --set full --clock-control none -k regex:dusp_fused -s 1 -c 1 -f -o gpurun_out/v12_w2mix python scripts/perf_probe4.py "t" w2:6250:88064:1 > gpurun_out/v12_ncu_w2mix.log 2>&1
ncu --set full --clock-control none -k regex:dusp_piped -s 1 -c 1 -f -o gpurun_out/v12_w3mix python scripts/perf_probe4.py "p24" w3:6250:88064:1 > gpurun_out/v12_ncu_w3mix.log 2>&1
ncu --set full --clock-control none -k regex:mix_push -c 1 -f -o gpurun_out/v12_mixpush python -m pytest tests/test_gpu_parity.py -q -k mix_push > gpurun_out/v12_ncu_mixpush.log 2>&1
ls -la gpurun_out/*.ncu-rep
