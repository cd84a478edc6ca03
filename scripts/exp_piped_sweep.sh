cd /root/repo
W=w3:10000:88064:0
python scripts/perf_probe4.py "p24;p24,DOPS_FILT_FENCE;p32,DOPS_FILT_FENCE;p16,DOPS_FILT_FENCE" $W
DUSP_PIPED_PWARPS=3 python scripts/perf_probe4.py "p24,DOPS_FILT_FENCE" $W
DUSP_PIPED_PWARPS=6 python scripts/perf_probe4.py "p24,DOPS_FILT_FENCE" $W
DUSP_PIPED_PWARPS=2 python scripts/perf_probe4.py "p24,DOPS_FILT_FENCE" $W
