cd /root/repo
W=w3:10000:88064:0
python scripts/perf_probe4.py "p24;p24,DOPS_FILT_INTW" $W
DUSP_FUSED_DEFS=DOPS_FILT_INTW python -m pytest tests -m gpu -x -q -k "w3 or pipelined or full_size" 2>&1 | tail -3
