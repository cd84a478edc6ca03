# Development aid: W5 mixdown at 8 GPUs with both transports of the sum (one-sided NVLink adds / ncclReduce).
cd "$(dirname "$0")/.."
for C in p2p nccl; do
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29751 bench.py --gpus 8 --steps 2 --warmup 3 --no-e2e --no-cpu --mix-collective $C > gpurun_out/v11_n8_$C.json 2> gpurun_out/v11_n8_$C.err
tail -2 gpurun_out/v11_n8_$C.err
done
python -c "
import json
for f in ['p2p','nccl']:
    d=json.loads(open('gpurun_out/v11_n8_%s.json'%f).read().strip().splitlines()[-1]); print(f, d['value'], d['mixdown']['value'], d['mixdown']['ms_per_step'], d['mixdown']['collective'][:40], d['clocks'])
"
