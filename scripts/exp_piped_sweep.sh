cd /root/repo
W=w3:10000:88064:0
python scripts/perf_probe4.py "p24;p24,DOPS_SPEC_FILTER" $W
DUSP_PIPED_STEP=128 python scripts/perf_probe4.py "p24;p24,DOPS_SPEC_FILTER;p32" $W
DUSP_PIPED_STEP=128 DUSP_PIPED_PWARPS=6 python scripts/perf_probe4.py "p24" $W
DUSP_PIPED_STEP=128 DUSP_PIPED_PWARPS=8 python scripts/perf_probe4.py "p24;p32;p16" $W
DUSP_PIPED_PWARPS=6 python scripts/perf_probe4.py "p24" $W
