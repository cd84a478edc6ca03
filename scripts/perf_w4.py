"""W4 probe (BASELINE configs[3]): N random circuits in one session, multi-program interpreter launch.
usage: perf_w4.py [n_circuits=10000] [seconds=10] [block_samples=88064]"""
import sys, os, time, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from dusp_b200 import capi, workloads as W
from dusp_b200.render import circuit_blob

n = int(sys.argv[1]) if len(sys.argv) > 1 else 10000
secs = float(sys.argv[2]) if len(sys.argv) > 2 else 10.0
t0 = time.perf_counter()
progs = [capi.Program(circuit_blob(W.w4_circuit(i))[0]) for i in range(n)]
t1 = time.perf_counter()
s = capi.Session(progs, [1] * n, 0)
t2 = time.perf_counter()
units = s.unit_count
blk = int(sys.argv[3]) if len(sys.argv) > 3 else 88064
total = int(secs * 44100)
buf = torch.empty(n * blk, dtype=torch.float32, device="cuda")
st = torch.cuda.current_stream()
times = []
for rep in range(3):
    s.reset()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for t in range(0, total, blk):
        s.render_device(min(blk, total - t), buf.data_ptr(), capi.OUT_VOICES, st.cuda_stream)
    e1.record()
    torch.cuda.synchronize()
    times.append(e0.elapsed_time(e1))
ms = min(times[1:])
print(json.dumps({"workload": "W4: %d random circuits x %g s mono" % (n, secs), "build_s": t1 - t0, "session_s": t2 - t1,
                  "units_total": units, "units_mean": units / n, "ms": ms, "unit_samples_per_s": units * total / (ms / 1e3),
                  "realtime_circuits": n * total / (ms / 1e3) / 44100, "launch": s.launch_info()}))
s.close()
