#!/usr/bin/env python
"""bench.py -- headline benchmark of the render path (contract: task prompt section 4).

Workload (BASELINE.json configs[2], the "10k-voice filter batch"): 10 000
voices x 60 s of  seeded Noise -> LP Filter -> CircleBuffer feedback delay ->
Pan (stereo), 8 reference units per voice, per-voice parameters from
dusp_b200.workloads (seed 1002).  One step = one pass over the whole batch:
the 60 s are rendered in time blocks; in DUSP_OUT_VOICES mode every voice's
stereo channel data is written to HBM (block buffers are recycled, the full
211 GB of per-voice audio does not fit in 180 GB).

  value    unit-samples/s, kernel launches only, parameters/state/rings resident in HBM
  e2e      same metric through the C ABI with HOST (pinned) output buffers: the
           per-voice channel data crosses PCIe inside the timed region
  roofline algorithmic bytes (24 B per voice-sample) / measured launch time
  cpu_baseline / --impl reference: the oracle (C restatement of the reference's
           JavaScript path) on all host cores, on a bounded sample of the same voices
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

SR = 44100
METRIC = "unit_samples_per_sec"


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--voices", type=int, default=10000, help="voices per GPU")
    ap.add_argument("--seconds", type=float, default=60.0)
    ap.add_argument("--block-seconds", type=float, default=2.0)
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--e2e-seconds", type=float, default=2.5, help="audio rendered per e2e step (the same at every N)")
    ap.add_argument("--no-workloads", action="store_true", help="skip the W1 / W2 / W4 sub-records (N = 1 only)")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--no-mix", action="store_true", help="skip the W5 stereo-mixdown + NCCL reduce leg")
    ap.add_argument("--mix-voices", type=int, default=12500, help="W5 voices per GPU (100 000 / 8)")
    ap.add_argument("--w4-circuits", type=int, default=10000)
    ap.add_argument("--mix-collective", default="auto", choices=["auto", "p2p", "nccl"],
                    help="transport of the mixdown sum: one-sided adds into rank 0's accumulator over NVLink peer memory "
                         "(default for N > 1) or ncclReduce per block")
    return ap.parse_args()


# ------------------------------------------------------------------ clocks
class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.proc, self.lines = index, None, []

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=lambda: self.lines.extend(self.proc.stdout), daemon=True)
            self.t.start()
        except OSError:
            self.proc = None

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        self.t.join(timeout=2)
        sm, mx, reasons = [], None, set()
        for l in self.lines:
            f = [x.strip() for x in l.split(",")]
            if len(f) < 6:
                continue
            try:
                sm.append(float(f[0])); mx = float(f[1])
            except ValueError:
                continue
            for name, val in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[2:6]):
                if val.lower().startswith("active"):
                    reasons.add(name)
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": mx, "reasons": sorted(reasons),
                "samples": len(sm)}


# ------------------------------------------------------------- CPU baseline
def cpu_baseline(n_units_per_voice, params_fn, budget_s=12.0, cores=None):
    """The oracle port on all host cores, bounded sample of the SAME voices at the
    SAME duration (voices are independent, so throughput scales linearly)."""
    from concurrent.futures import ThreadPoolExecutor
    from dusp_b200 import workloads as W
    from dusp_b200.render import circuit_blob
    from oracle import oracle  # CPU baseline leg: the one place bench.py executes oracle/
    oracle.lib()
    cores = cores or os.cpu_count() or 1
    secs = 10.0
    probe = W.w3_params(0, 1)
    blob, _ = circuit_blob(W.w3_voice_circuit(probe, 0))
    t0 = time.perf_counter()
    oracle.render(blob, secs)  # ctypes releases the GIL: threads scale across cores
    one = time.perf_counter() - t0
    per_core = max(1, int(budget_s / max(one, 1e-3)))
    n = per_core * cores
    params = params_fn(0, n)
    blobs = [circuit_blob(W.w3_voice_circuit(params, v))[0] for v in range(n)]
    t0 = time.perf_counter()
    with ThreadPoolExecutor(cores) as ex:
        list(ex.map(lambda b: oracle.render(b, secs)[0][:1], blobs))
    dt = time.perf_counter() - t0
    vs = n * int(secs * SR) / dt
    return {"value": vs * n_units_per_voice, "unit": "unit-samples/s", "cores": cores, "kind": "port",
            "sample": "%d voices x %g s of the W3 patch (same generator, voices 0..%d), C restatement of the "
                      "reference's JS render path (Node unavailable), %d threads, %.1f s wall" % (n, secs, n - 1, cores, dt),
            "voice_samples_per_s": vs, "realtime_voices": vs / SR}


def run_reference(args, rank, world):
    if rank != 0:
        return
    from dusp_b200 import workloads as W
    steps = []
    cb = None
    for i in range(args.warmup + args.steps):
        t0 = time.perf_counter()
        cb = cpu_baseline(W.W3_UNITS, W.w3_params, budget_s=2.0 if i < args.warmup else 6.0)
        steps.append(time.perf_counter() - t0)
    line = {
        "impl": "reference", "metric": METRIC, "value": cb["value"], "unit": "unit-samples/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * sum(steps[args.warmup:]) / max(args.steps, 1),
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": "W3 (BASELINE configs[2]): Noise->LP->CircleBuffer feedback delay->Pan, stereo; "
                               "bounded sample of the same voices on the host cores", "units_per_voice": W.W3_UNITS},
        "cpu_baseline": cb,
        "e2e": {"value": cb["value"], "unit": "unit-samples/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "realtime_voices": cb["realtime_voices"],
    }
    print(json.dumps(line), flush=True)


# ------------------------------------------------- W5: mixdown + the one collective
def measure_mixdown(args, rank, world, local_rank, barrier):
    """BASELINE configs[4] shape at N GPUs (weak scaling): every GPU renders 12 500 voices of W5 (even global
    index: W2-live FM voice, odd: W3 filter + feedback-delay voice), 10 s, stereo mixdown only; each 2 s
    block's [2][n] mix is sum-reduced onto rank 0 (NCCL over NVLink) while the next block renders."""
    import numpy as np
    import torch
    import torch.distributed as dist
    from dusp_b200 import capi, mixdown as MD, workloads as W
    from dusp_b200.render import circuit_blob

    nv = args.mix_voices
    voice0 = rank * nv
    i2, i3 = W.w5_split(voice0, nv)
    b2, _ = circuit_blob(W.w2_live_circuit())
    b3, _ = circuit_blob(W.w3_circuit())
    p2, p3 = capi.Program(b2), capi.Program(b3)
    sess = capi.Session([p2, p3], [len(i2), len(i3)], local_rank)
    o2 = W.w2_overrides(b2, W.w2_params(0, 0, i2))
    o3 = W.w3_overrides(b3, W.w3_params(0, 0, i3))
    sess.set_params(0, list(o2), np.stack([o2[k] for k in o2], axis=1))
    sess.set_params(1, list(o3), np.stack([o3[k] for k in o3], axis=1))
    secs = min(args.seconds, 10.0)
    n_total = int(secs * SR)
    blk = max(256, int(args.block_seconds * SR) // 256 * 256)
    r = MD.MixdownRenderer(sess, blk, torch.device("cuda", local_rank), collective=args.mix_collective)
    peak = [0.0]

    def on_block(t0, n, t):  # rank 0 consumes the reduced mix (device-side peak; the audio itself stays on the GPU)
        if rank == 0:
            peak[0] = max(peak[0], float(t.abs().max()))

    def step():
        sess.reset()
        r.render(n_total, on_block)

    stream = torch.cuda.current_stream()
    for _ in range(2):
        step()
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    k = max(1, min(args.steps, 3))
    e0.record(stream)
    for _ in range(k):
        step()
    e1.record(stream)
    barrier()
    ms = torch.tensor([e0.elapsed_time(e1)], device="cuda")
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    ms_step = float(ms.item()) / k
    r.close()

    # ---- self-check of the sum (every N > 1 run): block 0 of every rank's LOCAL mix is all-gathered and added in
    # f64 on rank 0; both transports must deliver that sum (1e-5 of the mix's full scale).
    mix_check = None
    if world > 1:
        n0 = min(blk, n_total)
        local = torch.zeros(2 * n0, dtype=torch.float32, device="cuda")
        sess.reset()
        sess.render_device(n0, local.data_ptr(), capi.OUT_MIX, stream.cuda_stream)
        torch.cuda.synchronize()
        gathered = [torch.empty_like(local) for _ in range(world)] if rank == 0 else None
        dist.gather(local, gathered, dst=0)
        want = torch.stack(gathered).double().sum(dim=0) if rank == 0 else None
        mix_check = {}
        for transport in ("p2p", "nccl"):
            got = {}

            def keep(t0, n, t):
                if t0 == 0:
                    got["b0"] = t.reshape(-1).clone()

            rr = MD.MixdownRenderer(sess, blk, torch.device("cuda", local_rank), collective=transport)
            sess.reset()
            rr.render(n0, keep)
            torch.cuda.synchronize()
            used = rr.collective
            rr.close()
            barrier()
            if rank == 0:
                full = max(1.0, float(want.abs().max()))
                err = float((got["b0"].double() - want).abs().max())
                mix_check[transport] = {"max_abs_err": err, "full_scale": full, "ok": err <= 1e-5 * full, "transport_used": used}

    # ---- e2e of the mix path: the whole W5 render through dusp_session_render with a HOST [2][n] buffer (this is the
    # output mode whose result fits through PCIe: 2 * n floats whatever the voice count)
    hmix = torch.empty(2 * n_total, dtype=torch.float32).pin_memory()

    def step_host():
        sess.reset()
        sess.render_into_host(hmix.data_ptr(), n_total, capi.OUT_MIX)

    step_host()
    barrier()
    t0 = time.perf_counter()
    for _ in range(k):
        step_host()
    barrier()
    dt = torch.tensor([time.perf_counter() - t0], device="cuda")
    if world > 1:
        dist.all_reduce(dt, op=dist.ReduceOp.MAX)
    units = (len(i2) * p2.num_units + len(i3) * p3.num_units) * world
    e2e_mix = {"value": units * n_total * k / float(dt.item()), "unit": "unit-samples/s", "d2h_bytes_per_step": 2 * n_total * 4,
               "note": "every rank renders its shard's stereo mix into its own pinned host buffer (no cross-rank sum: "
                       "the host-side sum of N [2][n] arrays is not part of the timed region)"}
    out = {"workload": "W5 (BASELINE configs[4] shape): %d voices/GPU (W2-live on even, W3 on odd global indices) x %g s, "
                       "stereo mixdown only" % (nv, secs),
           "value": units * n_total / (ms_step / 1e3), "unit": "unit-samples/s", "voices_total": nv * world,
           "realtime_voices": nv * world * n_total / (ms_step / 1e3) / SR, "ms_per_step": ms_step,
           "collective": ("none (one GPU)" if world == 1 else
                          "one-sided: every rank adds its [2][%d] f32 block into rank 0's accumulator over NVLink peer memory "
                          "(system-scope float reductions + arrival counters, csrc/mixpush.cu); no rendezvous kernel" % blk
                          if r.collective == "p2p" else
                          "ncclReduce(sum) of the [2][%d] f32 mix per block onto rank 0, async (overlaps the next block)" % blk),
           "mix_peak": peak[0], "mix_check": mix_check, "e2e_mix": e2e_mix}
    sess.close()
    return out


# ------------------------------------------- the other BASELINE configs (N = 1): W1, W2, W4
def measure_workloads(args):
    """BASELINE configs[0], [1] and [3] as short device-timed legs beside the headline (one GPU): per-voice output
    into a recycled device block buffer, state resident, CUDA events, best of 3 passes after one warm-up pass."""
    import numpy as np
    import torch
    from dusp_b200 import capi, workloads as W
    from dusp_b200.render import circuit_blob

    st = torch.cuda.current_stream()

    def timed(sess, total, rows, blk=88064):
        blk = min(blk, (total + 255) // 256 * 256)
        buf = torch.empty(rows * blk, dtype=torch.float32, device="cuda")
        best = 1e30
        for rep in range(4):
            sess.reset()
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(st)
            for t in range(0, total, blk):
                sess.render_device(min(blk, total - t), buf.data_ptr(), capi.OUT_VOICES, st.cuda_stream)
            e1.record(st)
            torch.cuda.synchronize()
            if rep:
                best = min(best, e0.elapsed_time(e1))
        return best / 1e3

    out = []
    total = 10 * SR
    # W1: `[Osc 440]`, 10 s mono, one circuit (a latency test on a GPU)
    blob, _ = circuit_blob(W.w1_circuit())
    p = capi.Program(blob)
    s = capi.Session([p], [1], 0)
    g = timed(s, total, 1, blk=441088)
    out.append({"workload": "W1 (BASELINE configs[0]): [Osc 440], 10 s mono, one circuit", "voices": 1, "units_per_voice": 1,
                "value": total / g, "unit": "unit-samples/s", "ms": g * 1e3, "kernel": s.launch_info()})
    s.close()
    # W2: FM ring-mod patch, 1 000 voices x 10 s, every voice its own live Ramp
    blob, _ = circuit_blob(W.w2_live_circuit())
    p = capi.Program(blob)
    s = capi.Session([p], [1000], 0)
    ov = W.w2_overrides(blob, W.w2_params(0, 1000))
    s.set_params(0, list(ov), np.stack([ov[k] for k in ov], axis=1))
    g = timed(s, total, 1000)
    out.append({"workload": "W2 (BASELINE configs[1]): [Multiply A:[Osc #o f:<live Ramp>] B:#o], 1 000 voices x 10 s",
                "voices": 1000, "units_per_voice": p.num_units, "value": 1000 * p.num_units * total / g, "unit": "unit-samples/s",
                "ms": g * 1e3, "bytes_per_voice_sample": p.bytes_per_voice_sample(capi.OUT_VOICES), "kernel": s.launch_info()})
    s.close()
    # W4: 10 000 random Dusp circuits (depth 4-8), 10 s mono, one multi-program launch
    nc = args.w4_circuits
    t0 = time.perf_counter()
    progs = [capi.Program(circuit_blob(W.w4_circuit(i))[0]) for i in range(nc)]
    s = capi.Session(progs, [1] * nc, 0)
    build_s = time.perf_counter() - t0
    g = timed(s, total, nc, blk=441088)  # 17.6 GB of channel data: one launch per pass, so the longest circuit is the tail once
    out.append({"workload": "W4 (BASELINE configs[3]): %d random Dusp circuits (depth 4-8), 10 s mono" % nc, "voices": nc,
                "units_per_voice": s.unit_count / nc, "value": s.unit_count * total / g, "unit": "unit-samples/s", "ms": g * 1e3,
                "host_build_s": build_s, "kernel": s.launch_info()})
    s.close()
    return out


# ------------------------------------------------------------------ B200 arm
def run_b200(args, rank, world, local_rank):
    import numpy as np
    import torch
    import torch.distributed as dist
    from dusp_b200 import capi, workloads as W
    from dusp_b200.render import circuit_blob

    torch.cuda.set_device(local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    hbm_peak = float(peaks.get("hbm_gbs", 6650.0))
    peak_src = "measured (MEASURED_PEAKS.json hbm_gbs)" if "hbm_gbs" in peaks else "fallback 6650 GB/s"

    nv = args.voices
    voice0 = rank * nv  # shard by global voice index: no data-path collective
    params = W.w3_params(voice0, nv)
    blob, _ = circuit_blob(W.w3_circuit())
    prog = capi.Program(blob)
    sess = capi.Session([prog], [nv], local_rank)
    ov = W.w3_overrides(blob, params)
    sess.set_params(0, list(ov), np.stack([ov[k] for k in ov], axis=1))

    n_total = int(args.seconds * SR)
    blk = max(256, int(args.block_seconds * SR) // 256 * 256)
    blocks = [(t, min(blk, n_total - t)) for t in range(0, n_total, blk)]
    nch = prog.num_channels
    dbuf = [torch.empty(nv * nch * blk, dtype=torch.float32, device="cuda") for _ in range(2)]
    stream = torch.cuda.current_stream()

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def step_device(events=None):
        sess.reset()
        for i, (t, n) in enumerate(blocks):
            if events is not None:
                e0 = torch.cuda.Event(enable_timing=True); e0.record(stream)
            sess.render_device(n, dbuf[i & 1].data_ptr(), capi.OUT_VOICES, stream.cuda_stream)
            if events is not None:
                e1 = torch.cuda.Event(enable_timing=True); e1.record(stream)
                events.append((e0, e1))

    def timed(fn, steps):
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        for _ in range(steps):
            fn()
        e1.record(stream)
        barrier()
        ms = torch.tensor([e0.elapsed_time(e1)], device="cuda")
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return float(ms.item())

    for _ in range(args.warmup):
        step_device()
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    launches0 = sess.launch_count
    launch_events = []
    total_ms = timed(lambda: step_device(launch_events), args.steps)
    launches = sess.launch_count - launches0
    clocks = sampler.stop() if rank == 0 else None
    launch_ms = [a.elapsed_time(b) for a, b in launch_events]

    voice_samples_step = nv * n_total * world
    units = prog.num_units
    ms_per_step = total_ms / args.steps
    value = voice_samples_step * units / (ms_per_step / 1e3)

    # roofline of the dominant (only) kernel: algorithmic bytes per launch / average launch time
    bpvs = prog.bytes_per_voice_sample(capi.OUT_VOICES)
    full = [m for m, (t, n) in zip(launch_ms, blocks * args.steps) if n == blk] or launch_ms
    avg_launch_ms = sum(full) / len(full)
    achieved = bpvs * nv * blk / (avg_launch_ms / 1e3) / 1e9
    # dram__bytes_read.sum + dram__bytes_write.sum of one such launch, from the committed `ncu --set full` capture
    # (profiles/ncu_traffic.json names the build it was taken from; ncu cannot run inside the timed bench)
    traffic, traffic_build = None, None
    try:
        tj = json.load(open(os.path.join(ROOT, "profiles", "ncu_traffic.json")))
        if tj["voices"] == nv and tj["samples"] == blk:
            traffic = tj["dram_bytes_read"] + tj["dram_bytes_write"]
            traffic_build = tj.get("build")
    except Exception:
        pass
    roofline = {"bound": "hbm", "achieved": achieved, "peak": hbm_peak, "unit": "GB/s", "frac": achieved / hbm_peak,
                "traffic": traffic, "traffic_capture": traffic_build, "algorithmic_bytes_per_launch": bpvs * nv * blk, "peak_source": peak_src, "kernel": "dusp_piped" if sess.launch_info().get("block") == 160 else "dusp_fused",
                "bytes_per_voice_sample": bpvs, "avg_launch_ms": avg_launch_ms,
                "kernel_share_of_step": sum(launch_ms) / total_ms}

    # e2e: the C-ABI call a host shim makes, HOST output buffers (pinned), D2H inside the timed region
    e2e = None
    if not args.no_e2e:
        # the SAME workload at every N (so the four scaling figures compare): 10k voices x 2.5 s stereo per rank
        # = 8.8 GB of pinned host memory per rank, 70 GB per box at N = 8
        e2e_secs = min(args.seconds, args.e2e_seconds)
        n_e2e = int(e2e_secs * SR)
        hbuf = torch.empty(nv * nch * n_e2e, dtype=torch.float32).pin_memory()
        dwords = np.stack([ov[k] for k in ov], axis=1)

        def step_e2e():
            sess.reset()                          # a step is a whole render from clock 0
            sess.set_params(0, list(ov), dwords)  # per-step inputs: the voices' parameters (H2D on first launch)
            sess.render_into_host(hbuf.data_ptr(), n_e2e, capi.OUT_VOICES)

        step_e2e()
        barrier()
        t0 = time.perf_counter()
        k = max(1, min(args.steps, 2))
        for _ in range(k):
            step_e2e()
        barrier()
        dt = torch.tensor([time.perf_counter() - t0], device="cuda")
        if world > 1:
            dist.all_reduce(dt, op=dist.ReduceOp.MAX)
        e2e_val = nv * n_e2e * world * units * k / float(dt.item())
        e2e = {"value": e2e_val, "unit": "unit-samples/s", "h2d_bytes_per_step": int(dwords.size * 8),
               "d2h_bytes_per_step": int(nv * nch * n_e2e * 4), "seconds_rendered": e2e_secs,
               "GBps_d2h_per_rank": nv * nch * n_e2e * 4 * k / float(dt.item()) / 1e9,
               "note": "per-voice stereo channel data copied to pinned host memory (PCIe-bound); the same per-rank "
                       "workload at every N"}
        del hbuf

    mixdown = None
    if not args.no_mix:
        mixdown = measure_mixdown(args, rank, world, local_rank, barrier)

    workloads = None
    if world == 1 and not args.no_workloads:
        workloads = measure_workloads(args)

    cb = None
    if rank == 0 and world == 1 and not args.no_cpu:
        cb = cpu_baseline(units, W.w3_params)
    if rank == 0:
        line = {
            "metric": METRIC, "value": value, "unit": "unit-samples/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": "W3 (BASELINE configs[2]): %d voices/GPU x %g s stereo, seeded Noise->LP Filter->"
                                   "CircleBuffer feedback delay->Pan, 8 units/voice, per-voice output written to HBM"
                                   % (nv, args.seconds),
                       "voices_per_gpu": nv, "seconds": args.seconds, "units_per_voice": units,
                       "time_block_samples": blk, "sharding": "voices by global index, no data-path collective",
                       "l2": "working set (rings %.2f GB + block outputs %.1f GB) larger than L2"
                             % (nv * 44100 * 4 / 1e9, 2 * nv * nch * blk * 4 / 1e9)},
            "realtime_voices": value / units / SR, "voice_samples_per_sec": value / units,
            "roofline": roofline, "cpu_baseline": cb, "e2e": e2e, "mixdown": mixdown, "workloads": workloads,
            "gpu_launches": int(launches),
            "launch": sess.launch_info(), "clocks": clocks,
        }
        print(json.dumps(line), flush=True)
    sess.close()
    if world > 1:
        dist.destroy_process_group()


def main():
    args = parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        run_reference(args, rank, world)
    else:
        run_b200(args, rank, world, local_rank)


if __name__ == "__main__":
    main()
