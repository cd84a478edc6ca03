# Round-end 8-GPU bench line (W3 weak scaling + e2e + W5 mixdown with the one-sided transport).
cd "$(dirname "$0")/.."
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29761 bench.py --gpus 8 --steps 3 --warmup 3 > gpurun_out/v14_n8.json 2> gpurun_out/v14_n8.err
tail -2 gpurun_out/v14_n8.err
python -c "
import json
for f in ['v14_n8']:
    d=json.loads(open('gpurun_out/%s.json'%f).read().strip().splitlines()[-1]); print(f, d['value'], d['roofline']['frac'], d['e2e']['value'], d['mixdown']['value'], d['mixdown']['ms_per_step'], d['mixdown']['collective'][:30], d['clocks'])
"
