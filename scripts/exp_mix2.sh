cd "$(dirname "$0")/.."
python -m pytest tests -m gpu -x -q -k "mix" 2>&1 | tail -3
for C in p2p nccl; do
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29741 bench.py --gpus 2 --steps 2 --warmup 3 --no-e2e --no-cpu --mix-collective $C > gpurun_out/v11_n2_$C.json 2> gpurun_out/v11_n2_$C.err
tail -3 gpurun_out/v11_n2_$C.err
done
python -c "
import json
for f in ['p2p','nccl']:
    d=json.loads(open('gpurun_out/v11_n2_%s.json'%f).read().strip().splitlines()[-1]); print(f, d['value'], d['mixdown']['value'], d['mixdown']['ms_per_step'], d['mixdown']['collective'][:40])
"
