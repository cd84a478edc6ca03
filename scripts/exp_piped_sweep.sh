cd /root/repo
W=w3:10000:88064:0
python scripts/perf_probe4.py "p24" $W
for v in rot 0 1 2 3; do DUSP_PIPED_SROT=$v python scripts/perf_probe4.py "p24" $W; done
