"""e2e probe: dusp_session_render into pinned host memory (per-voice output), W3 10k voices x S seconds."""
import sys, os, time, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from dusp_b200 import capi, workloads as W
from dusp_b200.render import circuit_blob
nv, secs = 10000, float(sys.argv[1]) if len(sys.argv) > 1 else 4.0
blob, _ = circuit_blob(W.w3_circuit())
prog = capi.Program(blob)
sess = capi.Session([prog], [nv], 0)
ov = W.w3_overrides(blob, W.w3_params(0, nv))
sess.set_params(0, list(ov), np.stack([ov[k] for k in ov], axis=1))
n = int(secs * 44100)
hbuf = torch.empty(nv * 2 * n, dtype=torch.float32).pin_memory()
for rep in range(3):
    sess.reset()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    sess.render_into_host(hbuf.data_ptr(), n)
    dt = time.perf_counter() - t0
    print(json.dumps({"staging_mb": os.environ.get("DUSP_STAGING_MB", "256"), "s": dt, "GBps": nv * 2 * n * 4 / dt / 1e9,
                      "unit_samples_per_s": nv * n * 8 / dt}), flush=True)
