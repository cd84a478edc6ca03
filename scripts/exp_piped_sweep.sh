# Development aid: sweep of the pipelined kernel's switches on W3 (10 000 voices x 88 064 samples, per-voice output).
# Results of round 1 are tabulated in DESIGN.md section 5.  usage (GPU box): bash scripts/exp_piped_sweep.sh
cd "$(dirname "$0")/.."
W=w3:10000:88064:0
python scripts/perf_probe4.py "p24;p32;p16" $W
DUSP_PIPED_STEP=128 python scripts/perf_probe4.py "p24" $W
for n in 2 3 6; do DUSP_PIPED_PWARPS=$n python scripts/perf_probe4.py "p24" $W; done
for v in rot 0 1 2 3; do DUSP_PIPED_SROT=$v python scripts/perf_probe4.py "p24" $W; done
