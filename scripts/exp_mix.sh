cd "$(dirname "$0")/.."
python -m pytest tests -m gpu -x -q -k "mix or Mix or w5 or W5" 2>&1 | tail -2
# mix-mode probes of the two W5 voice kinds (6250 voices each, 2 s block)
DUSP_PIPED=0 python scripts/perf_probe4.py "t" w2:6250:88064:1 w2:6250:88064:0
python scripts/perf_probe4.py "p24" w3:6250:88064:1 w3:6250:88064:0
python bench.py --no-e2e --no-cpu --steps 2 > gpurun_out/v11c.json 2>gpurun_out/v11c.err
DUSP_NO_CONCURRENT_GROUPS=1 python bench.py --no-e2e --no-cpu --steps 2 > gpurun_out/v11d.json 2>gpurun_out/v11d.err
python -c "
import json
for f in ['v11c','v11d']:
    d=json.loads(open('gpurun_out/%s.json'%f).read().strip().splitlines()[-1]); print(f, d['value'], d['roofline']['frac'], d['mixdown']['value'], d['mixdown']['ms_per_step'])
"
