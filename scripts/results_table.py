"""Fills BASELINE.md section 4: every BASELINE config on one B200 (device-timed, CUDA events, output per voice into a
recycled device block buffer) beside the oracle port on all host cores (bounded samples), plus a parity spot check.
usage: python scripts/results_table.py > gpurun_out/results_table.json"""
import json, os, sys, time
from concurrent.futures import ThreadPoolExecutor

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import torch
from dusp_b200 import capi, workloads as W
from dusp_b200.render import circuit_blob
from oracle import oracle  # CPU baseline + parity spot check (checker only)

SR = 44100
CORES = os.cpu_count() or 1
oracle.build()


def gpu_time(sess, total, n_rows, mode=capi.OUT_VOICES, blk=88064, reps=3):
    blk = min(blk, (total + 255) // 256 * 256)
    rows = n_rows if mode == capi.OUT_VOICES else 2
    buf = torch.empty(rows * blk, dtype=torch.float32, device="cuda")
    st = torch.cuda.current_stream()
    best = 1e30
    for _ in range(reps):
        sess.reset()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for t in range(0, total, blk):
            sess.render_device(min(blk, total - t), buf.data_ptr(), mode, st.cuda_stream)
        e1.record()
        torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1))
    return best / 1e3


def cpu_rate(blobs, secs, budget=6.0):
    """unit-samples/s of the oracle over `blobs` (one circuit each), all cores, bounded to ~budget seconds."""
    t0 = time.perf_counter()
    oracle.render(blobs[0], secs)
    one = time.perf_counter() - t0
    n = max(1, min(len(blobs), int(budget / max(one, 1e-4)) * CORES))
    t0 = time.perf_counter()
    with ThreadPoolExecutor(CORES) as ex:
        list(ex.map(lambda b: oracle.render(b, secs)[0][:1], blobs[:n]))
    return n * int(secs * SR) / (time.perf_counter() - t0), n


def spot(sess_row, blob, secs):
    want = oracle.render(blob, secs)
    n = len(want[0])
    err = max(float(np.max(np.abs(sess_row[c][:n].astype(np.float64) - want[c]))) for c in range(len(want)))
    return err


rows = []


def record(name, units_per_voice, voices, secs, gpu_s, cpu_vs, cpu_n, err, bpvs, note=""):
    total = int(secs * SR)
    vs = voices * total / gpu_s
    rows.append({"config": name, "units_per_voice": units_per_voice, "voices": voices, "seconds": secs,
                 "gpu_unit_samples_per_s": vs * units_per_voice, "gpu_realtime_voices": vs / SR,
                 "cpu_unit_samples_per_s": cpu_vs * units_per_voice, "cpu_cores": CORES, "cpu_sample_voices": cpu_n,
                 "hbm_frac": vs * bpvs / 6541.5e9 if bpvs else None, "bytes_per_voice_sample": bpvs,
                 "max_abs_err_vs_oracle": err, "note": note})
    print(json.dumps(rows[-1]), flush=True)


def head(sess, n):
    sess.reset()
    return sess.render_host(n)


# ---- W1
blob, _ = circuit_blob(W.w1_circuit())
p = capi.Program(blob); s = capi.Session([p], [1], 0)
g = gpu_time(s, 441000, 1, blk=441088)
t0 = time.perf_counter(); oracle.render(blob, 10.0); c1 = 441000 / (time.perf_counter() - t0)
record("W1 [Osc 440]", 1, 1, 10.0, g, c1, 1, spot(head(s, 22050)[0], blob, 0.5), 4.0, "CPU: one core (single circuit)")
s.close()

# ---- W2 literal / live, 1000 voices
blob, _ = circuit_blob(W.w2_literal_circuit())
p = capi.Program(blob); s = capi.Session([p], [1000], 0)
g = gpu_time(s, 441000, 1000)
c, n = cpu_rate([blob] * 64, 10.0)
record("W2-lit FM ring-mod (inert Ramp => 1 Hz)", p.num_units, 1000, 10.0, g, c, n, spot(head(s, 22050)[0], blob, 0.5), 4.0)
s.close()
blob, _ = circuit_blob(W.w2_live_circuit())
p = capi.Program(blob); s = capi.Session([p], [1000], 0)
prm = W.w2_params(0, 1000); ov = W.w2_overrides(blob, prm)
s.set_params(0, list(ov), np.stack([ov[k] for k in ov], axis=1))
g = gpu_time(s, 441000, 1000)
vb = [circuit_blob(W.w2_live_circuit(float(prm["f0"][v])))[0] for v in range(64)]
c, n = cpu_rate(vb, 10.0)
record("W2-live FM ring-mod", p.num_units, 1000, 10.0, g, c, n, spot(head(s, 22050)[0], vb[0], 0.5), 4.0)
s.close()

# ---- W3 10k x 60 s stereo
blob, _ = circuit_blob(W.w3_circuit())
p = capi.Program(blob); s = capi.Session([p], [10000], 0)
prm = W.w3_params(0, 10000); ov = W.w3_overrides(blob, prm)
s.set_params(0, list(ov), np.stack([ov[k] for k in ov], axis=1))
g = gpu_time(s, 60 * SR, 20000, reps=2)
vb = [circuit_blob(W.w3_voice_circuit(prm, v))[0] for v in range(256)]
c, n = cpu_rate(vb, 10.0)
s.close()
s = capi.Session([p], [96], 0)
prm96 = W.w3_params(0, 96); ov = W.w3_overrides(blob, prm96)
s.set_params(0, list(ov), np.stack([ov[k] for k in ov], axis=1))
record("W3 Noise->LP->feedback delay->Pan", p.num_units, 10000, 60.0, g, c, n, spot(head(s, 22050)[0], vb[0], 0.5), 24.0)
s.close()

# ---- W3b cyclic graph, 1000 voices
blob, _ = circuit_blob(W.w3b_circuit())
p = capi.Program(blob); s = capi.Session([p], [1000], 0)
g = gpu_time(s, 441000, 2000)
c, n = cpu_rate([blob] * 64, 10.0)
record("W3b SimpleDelay (cyclic graph)", p.num_units, 1000, 10.0, g, c, n, spot(head(s, 22050)[0], blob, 0.5),
       p.bytes_per_voice_sample(capi.OUT_VOICES))
s.close()

# ---- W4 10k random circuits
nc = 10000
blobs = [circuit_blob(W.w4_circuit(i))[0] for i in range(nc)]
progs = [capi.Program(b) for b in blobs]
s = capi.Session(progs, [1] * nc, 0)
g = gpu_time(s, 441000, nc, reps=2)
upv = s.unit_count / nc
c, n = cpu_rate(blobs[:512], 10.0)
cu = sum(pp.num_units for pp in progs[:n]) / n
s.close()
s = capi.Session(progs[:8], [1] * 8, 0)
h = head(s, 22050)
err = max(spot(h[i][:1], blobs[i], 0.5) for i in range(8))
s.close()
total = 441000
rows.append({"config": "W4 random circuits depth 4-8", "units_per_voice": upv, "voices": nc, "seconds": 10.0,
             "gpu_unit_samples_per_s": nc * upv * total / g, "gpu_realtime_voices": nc * total / g / SR,
             "cpu_unit_samples_per_s": c * cu, "cpu_cores": CORES, "cpu_sample_voices": n, "hbm_frac": None,
             "bytes_per_voice_sample": None, "max_abs_err_vs_oracle": err,
             "note": "one multi-program interpreter launch per block (a warp per circuit)"})
print(json.dumps(rows[-1]), flush=True)
