# GPU call 10 (8 GPUs): 2-GPU mixdown check kept on record; bench.py at N = 2 and N = 8 (mix_check in the line);
# device->host copy ceiling with 1 / 2 / 4 / 8 ranks copying at once
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
$TR --nproc-per-node 2 --master-port 29741 tests/mgpu_mix_check.py > gpurun_out/c10_mixcheck_2gpu.log 2>&1; echo "mixcheck rc=$?"; grep "MIX-CHECK" gpurun_out/c10_mixcheck_2gpu.log
$TR --nproc-per-node 2 --master-port 29742 bench.py --gpus 2 --steps 3 --warmup 3 > gpurun_out/c10_bench_n2.json 2> gpurun_out/c10_bench_n2.err; echo "bench2 rc=$?"
$TR --nproc-per-node 8 --master-port 29743 bench.py --gpus 8 --steps 3 --warmup 3 > gpurun_out/c10_bench_n8.json 2> gpurun_out/c10_bench_n8.err; echo "bench8 rc=$?"
python scripts/perf_d2h.py > gpurun_out/c10_d2h_n1.txt 2>&1
for n in 2 4 8; do $TR --nproc-per-node $n --master-port 2975$n scripts/perf_d2h.py > gpurun_out/c10_d2h_n$n.txt 2>&1; done
grep -h summary gpurun_out/c10_d2h_n*.txt | cut -c1-600; tail -1 gpurun_out/c10_d2h_n1.txt | cut -c1-600
python - <<'PY'
import json
for n in (2, 8):
    try:
        d = json.loads([l for l in open('gpurun_out/c10_bench_n%d.json' % n) if l.startswith('{')][-1])
        print(n, d['value'], d['roofline']['frac'], d['e2e']['value'], d['mixdown']['value'], d['mixdown']['mix_check'], d['mixdown']['e2e_mix']['value'])
    except Exception as e:
        print(n, 'parse failed', e)
PY
nvidia-smi topo -m > gpurun_out/c10_topo.txt 2>&1; numactl -H > gpurun_out/c10_numa.txt 2>&1; nproc >> gpurun_out/c10_numa.txt; free -g >> gpurun_out/c10_numa.txt
