"""Quick device-timed throughput probe of the render kernel on a few workloads
(development aid; the judged numbers come from bench.py)."""
import sys, os, time, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import dusp_b200 as d
from dusp_b200 import capi, components as C, patches as P
from dusp_b200.render import circuit_blob

sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import circuits


def probe(name, blob, n_voices, n_samples, mode, reps=3):
    p = capi.Program(blob)
    s = capi.Session([p], [n_voices], 0)
    floats = (n_voices * p.num_channels if mode == capi.OUT_VOICES else 2) * n_samples
    out = torch.empty(floats, dtype=torch.float32, device="cuda")
    times = []
    for r in range(reps + 1):
        s.reset()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        s.render_device(n_samples, out.data_ptr(), mode, torch.cuda.current_stream().cuda_stream)
        e1.record()
        torch.cuda.synchronize()
        times.append(e0.elapsed_time(e1))
    t = min(times[1:]) / 1e3
    vs = n_voices * n_samples / t
    bpvs = p.bytes_per_voice_sample(mode)
    print(json.dumps({"workload": name, "voices": n_voices, "samples": n_samples, "mode": mode, "ms": t * 1e3,
                      "voice_samples_per_s": vs, "unit_samples_per_s": vs * p.num_units, "GBps_alg": vs * bpvs / 1e9,
                      "frac_hbm": vs * bpvs / 6547.2e9, "launch": s.launch_info(), "slots": p.num_slots,
                      "instr": p.num_instructions}), flush=True)
    s.close()


if __name__ == "__main__":
    b_w2, _ = circuit_blob(circuits.w2_live())
    b_w3, _ = circuit_blob(circuits.w3_voice())
    b_osc, _ = circuit_blob(circuits.osc440())
    probe("osc440", b_osc, 1, 441000, capi.OUT_VOICES)
    probe("w2_live", b_w2, 1000, 441000, capi.OUT_VOICES)
    probe("w2_live", b_w2, 10000, 44100 * 2, capi.OUT_VOICES)
    probe("w2_live", b_w2, 100000, 44100, capi.OUT_VOICES)
    probe("w3", b_w3, 10000, 44100 * 2, capi.OUT_VOICES)
    probe("w3", b_w3, 10000, 44100 * 2, capi.OUT_MIX)
    probe("w3", b_w3, 100000, 44100 // 2, capi.OUT_MIX)
