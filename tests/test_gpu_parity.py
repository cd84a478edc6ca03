"""Parity tests proper: the CUDA path, called through the C ABI, against the
oracle on the same inputs.  Bar: bit-exact wherever the device evaluates only
IEEE +,-,*,/,sqrt and conversions; max-abs <= 1e-5 of full scale (north_star)
where a libdevice transcendental (tan / pow) is involved."""
import os

import numpy as np
import pytest

import dusp_b200 as d
from dusp_b200 import capi, components as C, patches as P
from dusp_b200.render import circuit_blob, n_samples_for

import circuits

pytestmark = pytest.mark.gpu
SR = 44100
TOL = 1e-5


def gpu_render(blob, seconds, n_voices=1):
    from dusp_b200.render import SessionRun
    run = SessionRun([blob], [n_voices], 0)  # events -> pokes between segments, graph edits -> adopted sessions
    try:
        return run.render(n_samples_for(seconds))
    finally:
        run.close()


@pytest.fixture(params=["interpreter", "specialised", "pipelined"])
def kernel_mode(request, monkeypatch):
    """The device paths: the op-program interpreter (render_sm100.cu), the straight-line tile kernel
    specialised for the circuit shape (fused.cc + ops.cuh, NVRTC) and the pipelined kernel (piped.cc:
    lanes along time for the parallel units, a lane-per-voice warp for the recurrences; circuits with
    no 3-stage split fall back to the tile kernel).  DUSP_FUSED / DUSP_PIPED force one."""
    monkeypatch.setenv("DUSP_FUSED", "0" if request.param == "interpreter" else "1")
    monkeypatch.setenv("DUSP_PIPED", "1" if request.param == "pipelined" else "0")
    return request.param


def compare(got, want, exact, what=""):
    assert got.shape[0] == len(want), what
    for c, w in enumerate(want):
        g = got[c]
        assert g.shape == w.shape
        assert not np.isnan(g).any()
        if exact:
            if not np.array_equal(g, w):
                bad = np.flatnonzero(g != w)
                raise AssertionError("%s ch%d: %d/%d samples differ, first at %d: %r vs %r" %
                                     (what, c, bad.size, g.size, bad[0], g[bad[0]], w[bad[0]]))
        else:
            full = max(1.0, float(np.max(np.abs(w))) if w.size else 1.0)
            err = float(np.max(np.abs(g.astype(np.float64) - w.astype(np.float64)))) if w.size else 0.0
            assert err <= TOL * full, "%s ch%d: max-abs error %g (full scale %g)" % (what, c, err, full)


@pytest.mark.parametrize("name", sorted(circuits.CASES))
def test_case_matches_oracle(lib, oracle, name, kernel_mode):
    fac, seconds, exact = circuits.CASES[name]
    blob, _ = circuit_blob(fac())
    compare(gpu_render(blob, seconds)[0], oracle.render(blob, seconds), exact, name)


def test_osc440_ten_seconds_bit_exact(lib, oracle):
    """BASELINE config 0: [Osc 440] rendered 10 s mono."""
    blob, _ = circuit_blob(circuits.osc440())
    got = gpu_render(blob, 10)[0]
    assert got.shape == (1, 441000)
    compare(got, oracle.render(blob, 10), True, "osc440 10s")


def test_render_channel_data_api(lib, oracle):
    cd = d.renderChannelData(d.unDusp("(O200 + O250) @ 0"), 1)
    assert len(cd) == 2 and cd[0].shape == (SR,) and cd.sampleRate == SR and cd[0].dtype == np.float32
    blob, _ = circuit_blob(d.unDusp("(O200 + O250) @ 0"))
    compare(np.stack(cd), oracle.render(blob, 1), False, "sum_pan api")
    with pytest.raises(d.DuspError, match="expects an outlet"):
        d.renderChannelData(None, 1)


@pytest.mark.parametrize("seconds", [0, 1 / SR, 255 / SR, 256 / SR, 257 / SR, 0.0123])
def test_ragged_durations(lib, oracle, seconds):
    blob, _ = circuit_blob(circuits.w2_live())
    got = gpu_render(blob, seconds)[0]
    want = oracle.render(blob, seconds)
    if not want:
        assert got.shape[-1] == 0
    else:
        compare(got, want, True, "ragged %g" % seconds)


def test_block_wise_render_continues_state(lib, oracle, kernel_mode):
    blob, _ = circuit_blob(circuits.w3_voice())
    p = capi.Program(blob)
    s = capi.Session([p], [1], 0)
    a = s.render_host(1024)
    b = s.render_host(2048)
    s.reset()
    whole = s.render_host(3072)
    s.close()
    assert np.array_equal(np.concatenate([a, b], axis=-1), whole)


def unit_index(blob, kind, nth=0):
    kinds = [l.split()[1] for l in blob.splitlines() if l.startswith("U ")]
    return [i for i, k in enumerate(kinds) if k == kind][nth]


@pytest.mark.parametrize("tile", ["8,1", "16,1", "8,4", "8,8", "8,32"])
@pytest.mark.parametrize("name", ["w2_live", "w3_voice", "simple_delay", "fixed_delays", "envelope_fm", "filter_modulated",
                                  "noise_sample_hold", "delay_mixer", "ahd_timer", "fm_osc"])
def test_tile_shapes(lib, oracle, name, tile, kernel_mode, monkeypatch):
    """Every V x T tile shape must give the reference's result: time-lanes only change how the work
    is spread over a warp (exact scans, serial turns), never the arithmetic."""
    monkeypatch.setenv("DUSP_TILE", tile)
    fac, seconds, exact = circuits.CASES[name]
    blob, _ = circuit_blob(fac())
    compare(gpu_render(blob, min(seconds, 0.5))[0], oracle.render(blob, min(seconds, 0.5)), exact, name + "@" + tile)


def test_w2_live_batch_with_per_voice_parameters(lib, oracle, kernel_mode):
    """BASELINE config 1 shape (FM ring-mod), 96 voices with their own ramp, bit-exact."""
    rng = np.random.default_rng(1001)
    f0 = rng.uniform(100, 800, 96)
    blob, _ = circuit_blob(circuits.w2_live(300.0))
    ramp = unit_index(blob, "Ramp")
    p = capi.Program(blob)
    s = capi.Session([p], [96], 0)
    s.set_params(0, ["u%d.y0" % ramp, "u%d.y1" % ramp], np.stack([f0, f0 / 2], axis=1))
    got = s.render_host(SR // 2)
    s.close()
    for v in (0, 1, 31, 32, 63, 95):
        vb, _ = circuit_blob(circuits.w2_live(float(f0[v])))
        compare(got[v], oracle.render(vb, 0.5), True, "w2 voice %d" % v)


def w3_overrides(blob, f, dly, g, pan, seeds):
    filt, pan_u = unit_index(blob, "Filter"), unit_index(blob, "Pan")
    reader, mult, noise = unit_index(blob, "CircleBufferReader"), unit_index(blob, "Multiply"), unit_index(blob, "Noise")
    names = ["u%d.f[0]" % filt, "u%d.offset[0]" % reader, "u%d.b[0]" % mult, "u%d.pan[0]" % pan_u,
             "u%d.seed_lo" % noise, "u%d.seed_hi" % noise, "u%d.y" % noise]
    y0 = np.array([2 * d.config.rand_tick(int(sd), 0) - 1 for sd in seeds])
    vals = np.stack([f, dly, g, pan, (seeds & 0xFFFFFFFF).astype(np.float64), (seeds >> 32).astype(np.float64), y0], axis=1)
    return names, vals


def test_w3_batch_with_per_voice_parameters(lib, oracle, kernel_mode):
    """BASELINE config 2 shape: Noise -> LP -> CircleBuffer feedback delay -> Pan, 40 voices."""
    rng = np.random.default_rng(1002)
    n = 40
    f, dly = rng.uniform(200, 4000, n), rng.uniform(0.05, 0.5, n)
    g, pan = rng.uniform(0.2, 0.7, n), rng.uniform(-1, 1, n)
    seeds = rng.integers(0, 2**63, n, dtype=np.uint64)
    blob, _ = circuit_blob(circuits.w3_voice())
    p = capi.Program(blob)
    s = capi.Session([p], [n], 0)
    names, vals = w3_overrides(blob, f, dly, g, pan, seeds)
    s.set_params(0, names, vals)
    got = s.render_host(SR)
    mix = None
    s.reset()
    mix = s.render_host(SR, capi.OUT_MIX)
    s.close()
    ref_mix = np.zeros((2, SR))
    for v in range(n):
        filt = C.Filter(C.Noise(seed=int(seeds[v])), float(f[v]), "LP")
        mt = P.MultiTapDelay(1, 1.0, filt)
        tap = mt.addFeedback(float(dly[v]), float(g[v]), 0)
        vb, _ = circuit_blob(C.Pan(C.Sum(filt, tap), float(pan[v])))
        want = oracle.render(vb, 1.0)
        if v % 7 == 0:
            compare(got[v], want, True, "w3 voice %d" % v)
        ref_mix += np.stack(want).astype(np.float64)
    assert np.max(np.abs(mix - ref_mix)) <= TOL * max(1.0, np.max(np.abs(ref_mix)))
    assert np.max(np.abs(mix - got.astype(np.float64).sum(axis=0))) <= 1e-5 * max(1.0, np.max(np.abs(ref_mix)))


def test_small_batch_renders_a_voice_on_several_warps(lib, oracle, monkeypatch):
    """Segmented rendering (ops.cuh "segments of one voice's timeline"): 70 FM voices with their own Ramp, rendered in
    three launches -- every launch cuts a voice's timeline into segments on different warps (a first kernel leaves each
    segment's wrapped sum of the oscillator's increments).  Bit for bit the render in one piece and the oracle; a voice
    whose increments are too small for exact partial sums (f = 1e-6) is rendered whole, in the same launches."""
    blob, _ = circuit_blob(circuits.w2_live())
    ramp = unit_index(blob, "Ramp")
    n = 70
    f0 = np.linspace(110.0, 790.0, n)
    f0[5] = 1e-6   # below 2^-13: this voice's sums are not exact -> one piece
    vals = np.stack([f0, f0 / 2], axis=1)
    lens = [256 * 40 + 3, 256 * 7, 256 * 61 + 200]

    def render():
        s = capi.Session([capi.Program(blob)], [n], 0)
        s.set_params(0, ["u%d.y0" % ramp, "u%d.y1" % ramp], vals)
        l0 = s.launch_count
        out = [s.render_host(k) for k in lens]
        launches = s.launch_count - l0
        info = s.launch_info()
        s.close()
        return out, launches, info
    seg, seg_launches, info = render()
    monkeypatch.setenv("DUSP_NO_SEGMENTS", "1")
    whole, whole_launches, _ = render()
    monkeypatch.delenv("DUSP_NO_SEGMENTS")
    assert info["specialised"] and seg_launches == 2 * whole_launches == 6
    for a, b in zip(seg, whole):
        assert np.array_equal(a, b)
    for v in (0, 5, 33, 69):  # the first launch of a few voices against the oracle
        osc = C.Osc(C.Ramp(88200, float(f0[v]), float(f0[v]) / 2).trigger())
        vb, _ = circuit_blob(C.Multiply(osc, osc))
        want = oracle.render(vb, 0.25)
        m = min(lens[0], len(want[0]))
        compare(seg[0][v][:, :m], [w[:m] for w in want], True, "voice %d" % v)


def test_w3_three_voices_at_the_benchmarked_sixty_seconds(lib, oracle, kernel_mode):
    """BASELINE config 2 renders 60 s: three voices of bench.py's generator at that length, bit for bit (the
    44 100-sample CircleBuffer wraps 60 times, the noise stream runs 2.6 M draws, the session renders it in the
    bench's 2 s blocks)."""
    from dusp_b200 import workloads as W
    idx = [0, 4999, 9999]
    blob, _ = circuit_blob(W.w3_circuit())
    params = W.w3_params(0, 0, idx)
    ov = W.w3_overrides(blob, params)
    p = capi.Program(blob)
    s = capi.Session([p], [len(idx)], 0)
    s.set_params(0, list(ov), np.stack([ov[k] for k in ov], axis=1))
    n, blk = 60 * SR, 88064
    got = np.concatenate([s.render_host(min(blk, n - t)) for t in range(0, n, blk)], axis=-1)
    s.close()
    for j in range(len(idx)):
        vb, _ = circuit_blob(W.w3_voice_circuit(params, j))
        compare(got[j], oracle.render(vb, 60.0), True, "w3 voice %d at 60 s" % idx[j])


def test_render_into_a_misaligned_device_buffer(lib, oracle, kernel_mode):
    """include/dusp_b200.h: any 4-byte aligned out_device is accepted (the 16-byte vector stores of the copy-out
    fall back to scalar stores).  A torch slice offset by one float, n_samples % 4 == 0."""
    import torch
    blob, _ = circuit_blob(circuits.w3_voice())
    p = capi.Program(blob)
    n, nv = 4096, 70
    want = oracle.render(blob, n / SR)
    for mode, rows in ((capi.OUT_VOICES, nv * 2), (capi.OUT_MIX, 2)):
        s = capi.Session([p], [nv], 0)
        buf = torch.full((rows * n + 8,), 7.0, dtype=torch.float32, device="cuda")
        s.render_device(n, buf.data_ptr() + 4, mode)
        torch.cuda.synchronize()
        s.close()
        out = buf.cpu().numpy()
        assert out[0] == 7.0 and np.all(out[rows * n + 1:] == 7.0)  # nothing outside the buffer was touched
        got = out[1:rows * n + 1].reshape(-1, 2, n)
        if mode == capi.OUT_VOICES:
            for v in (0, 33, 69):
                compare(got[v], want, True, "misaligned voice %d" % v)
        else:
            ref = np.stack(want).astype(np.float64) * nv
            assert float(np.max(np.abs(got[0] - ref))) <= 1e-5 * max(1.0, float(np.max(np.abs(ref))))


def test_set_params_after_a_render_needs_a_reset(lib):
    """dusp_session_set_params re-initialises ONE group at clock 0: refused once the session has rendered (the
    other groups would keep their advanced state), fine again after dusp_session_reset."""
    blob, _ = circuit_blob(circuits.w2_live())
    p = capi.Program(blob)
    s = capi.Session([p, p], [2, 2], 0)
    ramp = unit_index(blob, "Ramp")
    s.set_params(0, ["u%d.y0" % ramp], np.array([[300.0], [400.0]]))
    a = s.render_host(512)
    with pytest.raises(d.DuspError, match="dusp_session_reset"):
        s.set_params(1, ["u%d.y0" % ramp], np.array([[300.0], [400.0]]))
    s.reset()
    s.set_params(1, ["u%d.y0" % ramp], np.array([[300.0], [400.0]]))
    b = s.render_host(512)
    s.close()
    assert np.array_equal(a[:2], b[:2]) and np.array_equal(b[:2], b[2:])


def test_state_block_checkpoint_and_resume(lib, kernel_mode):
    """dusp_session_state_export / _import (include/dusp_b200.h): a render continued in a NEW session from an exported
    block is bit for bit the render that never stopped -- words (phases, biquad memory, ring heads, noise draw
    counters), rings and stale chunks all travel.  Two program groups (feedback delay voices and FM voices); a block
    from a session of another shape is refused."""
    blob3, _ = circuit_blob(circuits.w3_voice())
    blob2, _ = circuit_blob(circuits.w2_live())
    progs = [capi.Program(blob3), capi.Program(blob2)]
    n = 3 * 256 * 7 + 100  # not chunk-aligned at the end: the clock runs in whole chunks (Circuit.js:38)
    whole = capi.Session(progs, [3, 2], 0)
    a1 = whole.render_host(256 * 9)
    a2 = whole.render_host(n)
    first = capi.Session(progs, [3, 2], 0)
    b1 = first.render_host(256 * 9)
    block = first.export_state()
    clock = first.clock
    first.close()
    second = capi.Session(progs, [3, 2], 0)
    second.import_state(block)
    assert second.clock == clock == whole.clock - (n + 255) // 256 * 256
    b2 = second.render_host(n)
    assert np.array_equal(a1, b1) and np.array_equal(a2, b2)
    fresh = capi.Session(progs, [3, 2], 0)   # a session that never rendered exports its initial state
    third = capi.Session(progs, [3, 2], 0)
    third.import_state(fresh.export_state())
    assert np.array_equal(third.render_host(256 * 9), a1)
    other = capi.Session(progs, [2, 2], 0)
    with pytest.raises(d.DuspError, match="another shape"):
        other.import_state(block)
    with pytest.raises(d.DuspError, match="not a state block"):
        other.import_state(np.zeros(other.export_state().size, dtype=np.uint8))
    for x in (whole, second, fresh, third, other):
        x.close()


def test_multi_program_session(lib, oracle, kernel_mode):
    cases = ["w2_live", "w3_voice", "fixed_delays", "stereo_concat"]
    blobs = [circuit_blob(circuits.CASES[c][0]())[0] for c in cases]
    progs = [capi.Program(b) for b in blobs]
    s = capi.Session(progs, [3, 2, 1, 2], 0)
    got = s.render_host(4096)
    s.close()
    assert got.shape == (8, 2, 4096)
    row = 0
    for c, b, nv in zip(cases, blobs, [3, 2, 1, 2]):
        want = oracle.render(b, 4096 / SR)
        for _ in range(nv):
            compare(got[row][:len(want)], want, circuits.CASES[c][2], c)
            assert np.all(got[row][len(want):] == 0)
            row += 1


def test_concurrent_group_launches_give_the_same_render(lib, oracle, kernel_mode, monkeypatch):
    """DUSP_CONCURRENT_GROUPS=1 (opt-in, capi.cc enqueue_block): the program groups of a session launch on side
    streams between a fork and a join event instead of in sequence.  Same output, block after block."""
    cases = ["w2_live", "w3_voice", "fixed_delays", "stereo_concat"]
    blobs = [circuit_blob(circuits.CASES[c][0]())[0] for c in cases]

    def render():
        progs = [capi.Program(b) for b in blobs]
        s = capi.Session(progs, [3, 2, 1, 2], 0)
        a = s.render_host(3072)
        b = s.render_host(1536)  # a second block (whole chunks) continues the state through the join
        s.close()
        return np.concatenate([a, b], axis=2)

    seq = render()
    monkeypatch.setenv("DUSP_CONCURRENT_GROUPS", "1")
    par = render()
    assert np.array_equal(seq, par)
    want = oracle.render(blobs[1], 4608 / SR)
    compare(par[3][:len(want)], want, circuits.CASES["w3_voice"][2], "w3_voice on a side stream")


def test_unsupported_voice_parameters_are_rejected(lib):
    blob, _ = circuit_blob(circuits.w3_voice())
    p = capi.Program(blob)
    s = capi.Session([p], [2], 0)
    reader = unit_index(blob, "CircleBufferReader")
    with pytest.raises(capi.UnsupportedCircuit, match="closer than one"):
        s.set_params(0, ["u%d.offset[0]" % reader], np.array([[0.2], [0.001]]))
    s.close()


def test_full_size_w2_properties(lib):
    """BASELINE config 1 at full size (1000 voices x 10 s): size-independent properties.
    Voices with equal parameters are identical; integer-frequency voices equal the
    closed-form table read-out of Osc.js:39-45 squared."""
    import math
    n, secs = 1000, 10
    blob, _ = circuit_blob(circuits.w2_live(400.0))
    ramp = unit_index(blob, "Ramp")
    p = capi.Program(blob)
    s = capi.Session([p], [n], 0)
    rng = np.random.default_rng(5)
    y0 = rng.uniform(100, 800, n)
    y0[::50] = 440.0
    y1 = y0 / 2
    y1[::50] = 440.0  # flat ramp: constant integer frequency
    s.set_params(0, ["u%d.y0" % ramp, "u%d.y1" % ramp], np.stack([y0, y1], axis=1))
    got = s.render_host(secs * SR)
    s.close()
    tab = np.array([math.sin(2 * math.pi * t / 44101) for t in range(44101)], dtype=np.float32)
    k = np.arange(secs * SR, dtype=np.int64)
    o = tab[(440 * (k + 1)) % SR]
    assert np.array_equal(got[0, 0], o * o)
    for v in range(50, n, 50):
        assert np.array_equal(got[v, 0], got[0, 0])
    assert not np.isnan(got).any() and float(np.abs(got).max()) <= 1.0
    assert len({got[v, 0].tobytes() for v in range(1, 50)}) == 49  # distinct voices differ


def test_full_size_w3_properties(lib, oracle):
    """BASELINE config 2 at its full voice count (10 000 voices of the Noise -> LP -> feedback delay -> Pan patch,
    bench.py's generator, 1 s): a spread of voices equals the oracle's stand-alone render of that voice's circuit
    bit for bit; rendering in two blocks equals rendering at once for EVERY voice; the stereo mixdown of the same
    batch equals the sum of the per-voice renders; no NaN anywhere."""
    import torch
    from dusp_b200 import workloads as W
    n, ns = 10000, SR
    blob, _ = circuit_blob(W.w3_circuit())
    params = W.w3_params(0, n)
    ov = W.w3_overrides(blob, params)
    p = capi.Program(blob)
    s = capi.Session([p], [n], 0)
    s.set_params(0, list(ov), np.stack([ov[k] for k in ov], axis=1))
    whole = torch.empty(n * 2 * ns, dtype=torch.float32, device="cuda")
    s.render_device(ns, whole.data_ptr())
    torch.cuda.synchronize()
    assert s.launch_info()["specialised"] and s.launch_info()["block"] == 160  # the pipelined kernel, as benchmarked
    whole = whole.view(n, 2, ns)
    for v in (0, 1, 23, 24, 4999, 7777, 9983, 9999):  # first / last lanes and CTAs of the 24-voice tiling
        vb, _ = circuit_blob(W.w3_voice_circuit(params, v))
        compare(whole[v].cpu().numpy(), oracle.render(vb, 1.0), True, "full-size w3 voice %d" % v)
    assert not bool(torch.isnan(whole).any())
    # two blocks (second one not chunk-aligned in length) == one render, for every voice
    s.reset()
    n1 = 20480
    a = torch.empty(n * 2 * n1, dtype=torch.float32, device="cuda")
    b = torch.empty(n * 2 * (ns - n1), dtype=torch.float32, device="cuda")
    s.render_device(n1, a.data_ptr())
    s.render_device(ns - n1, b.data_ptr())
    torch.cuda.synchronize()
    assert torch.equal(a.view(n, 2, n1), whole[:, :, :n1])
    assert torch.equal(b.view(n, 2, ns - n1), whole[:, :, n1:])
    del a, b
    # mixdown of the batch == sum over voices (float summation order differs: tolerance of the mix's full scale)
    s.reset()
    mix = torch.empty(2 * ns, dtype=torch.float32, device="cuda")
    s.render_device(ns, mix.data_ptr(), capi.OUT_MIX)
    torch.cuda.synchronize()
    want = whole.double().sum(dim=0)
    scale = max(1.0, float(want.abs().max()))
    assert float((mix.view(2, ns).double() - want).abs().max()) <= 1e-5 * scale
    s.close()


@pytest.mark.parametrize("n", [1001, 4999, 256 * 9 + 3])
def test_mixdown_ragged_lengths_all_kernel_paths(lib, kernel_mode, n):
    """DUSP_OUT_MIX at lengths that are no multiple of 4: channel 1 of the [2][n] mix then starts unaligned and the
    last run of every lane is partial, so the epilogue's 16-byte vector reductions must fall back to scalar adds
    exactly there (ops.cuh / render_sm100.cu mix_add).  Mix == sum over the per-voice renders, mono (W2: copied to
    both channels) and stereo (W3) programs in one session."""
    from dusp_b200 import workloads as W
    i2, i3 = W.w5_split(0, 14)
    b2, _ = circuit_blob(W.w2_live_circuit())
    b3, _ = circuit_blob(W.w3_circuit())
    p2, p3 = capi.Program(b2), capi.Program(b3)
    s = capi.Session([p2, p3], [len(i2), len(i3)], 0)
    o2, o3 = W.w2_overrides(b2, W.w2_params(0, 0, i2)), W.w3_overrides(b3, W.w3_params(0, 0, i3))
    s.set_params(0, list(o2), np.stack([o2[k] for k in o2], axis=1))
    s.set_params(1, list(o3), np.stack([o3[k] for k in o3], axis=1))
    voices = s.render_host(n).astype(np.float64)          # [14][2][n]; a mono program leaves channel 1 zero
    s.reset()
    mix = s.render_host(n, capi.OUT_MIX).astype(np.float64)
    s.close()
    want = voices.sum(axis=0)
    want[1] += voices[:len(i2), 0].sum(axis=0)             # mono voices go to both mix channels
    scale = max(1.0, float(np.max(np.abs(want))))
    assert mix.shape == (2, n)
    assert float(np.max(np.abs(mix - want))) <= 1e-5 * scale
    assert float(np.max(np.abs(want))) > 0.1


def test_mixdown_renderer_blocks_match_whole_render(lib):
    """dusp_b200.mixdown.MixdownRenderer (the multi-GPU mix path; here one GPU, no process group): block-wise
    OUT_MIX renders into rotating device buffers equal one whole-length mix render."""
    import torch
    from dusp_b200 import mixdown, workloads as W
    i2, i3 = W.w5_split(10, 12)
    b2, _ = circuit_blob(W.w2_live_circuit())
    b3, _ = circuit_blob(W.w3_circuit())
    p2, p3 = capi.Program(b2), capi.Program(b3)
    s = capi.Session([p2, p3], [len(i2), len(i3)], 0)
    o2, o3 = W.w2_overrides(b2, W.w2_params(0, 0, i2)), W.w3_overrides(b3, W.w3_params(0, 0, i3))
    s.set_params(0, list(o2), np.stack([o2[k] for k in o2], axis=1))
    s.set_params(1, list(o3), np.stack([o3[k] for k in o3], axis=1))
    n = 5000
    whole = s.render_host(n, capi.OUT_MIX)
    s.reset()
    got = np.zeros((2, n), dtype=np.float32)

    def on_block(t0, k, t):
        got[:, t0:t0 + k] = t.cpu().numpy()

    mixdown.MixdownRenderer(s, 2048, torch.device("cuda", 0)).render(n, on_block)
    torch.cuda.synchronize()
    s.close()
    assert np.max(np.abs(got - whole)) <= 1e-5 * max(1.0, float(np.max(np.abs(whole))))
    assert float(np.max(np.abs(whole))) > 0.1


def test_mix_push_adds_blocks_and_counts_arrivals(lib):
    """csrc/mixpush.cu on one GPU: three pushes of different blocks into the accumulator (the transport of the
    multi-GPU mixdown sum, here without a peer) add up exactly and raise the block's arrival counter to 3;
    dusp_mix_wait then lets the stream through."""
    import torch
    n = 2 * 3000 + 2  # not a multiple of the grid stride
    acc = capi.MixAccumulator(0, 2 * n, 2)
    ticket = torch.zeros(1, dtype=torch.int32, device="cuda")
    rng = np.random.default_rng(5)
    parts = [rng.integers(-8, 9, size=n).astype(np.float32) / 4 for _ in range(3)]  # dyadic: float sums are exact
    st = torch.cuda.current_stream().cuda_stream
    acc.clear(st)
    for part in parts:
        t = torch.from_numpy(part).cuda()
        acc.push(t.data_ptr(), n, n, 1, ticket.data_ptr(), st)  # into the second block
        torch.cuda.synchronize()
    acc.wait(1, 3, st)
    torch.cuda.synchronize()
    from dusp_b200.mixdown import _DevArray
    got = torch.as_tensor(_DevArray(acc.acc, 2 * n), device="cuda").cpu().numpy()
    ctr = np.frombuffer(torch.as_tensor(_DevArray(acc.counters, 2), device="cuda").cpu().numpy().tobytes(), dtype=np.uint32)
    acc.close()
    assert np.array_equal(got[:n], np.zeros(n, dtype=np.float32))
    assert np.array_equal(got[n:], parts[0] + parts[1] + parts[2])
    assert ctr.tolist() == [0, 3] and int(ticket.item()) == 0


def test_mixdown_two_gpus_p2p_matches_nccl_and_single_process(lib):
    """The W5 mixdown on two GPUs (torchrun, one process per GPU): the one-sided NVLink sum and the ncclReduce
    transport give the same [2][n] mix as one process rendering all voices (1e-5 of the mix's full scale)."""
    import subprocess, sys, torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    script = os.path.join(os.path.dirname(os.path.abspath(__file__)), "mgpu_mix_check.py")
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2",
                        "--master-addr", "127.0.0.1", "--master-port", "29731", script],
                       capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    assert "MIX-CHECK OK" in r.stdout


def test_w4_random_circuits_in_one_multi_program_launch(lib, oracle):
    """BASELINE configs[3] shape at test size: 24 different random circuits (depth 4-8, Osc / Ramp / Shape / Noise /
    Multiply / Sum / Filter / AllPass / FixedDelay) in ONE session: every circuit is its own program, all rendered by one
    multi-program launch (a warp per circuit, vm.h DuspWarpDesc).  Every op here is in the bit-exact class."""
    from dusp_b200 import workloads as W
    n, secs = 24, 0.3
    blobs = [circuit_blob(W.w4_circuit(i))[0] for i in range(n)]
    progs = [capi.Program(b) for b in blobs]
    s = capi.Session(progs, [1] * n, 0)
    l0 = s.launch_count
    got = s.render_host(n_samples_for(secs))
    launches = s.launch_count - l0
    s.close()
    assert launches == 1, "heterogeneous batch must render in one launch per block"
    for i, b in enumerate(blobs):
        compare(got[i][:1], oracle.render(b, secs), True, "w4 circuit %d" % i)


def test_batch_with_scheduled_events(lib, oracle, kernel_mode):
    """trigger / state-assigning events after the first chunk, for a whole batch: the session renders
    chunk-aligned segments and pokes the events' assignments into every voice between them (dusp_session_poke)."""
    fac, seconds, exact = circuits.CASES["scheduled_retriggers"]
    b = d.Batch(fac(), 70)
    try:
        got = b.render(seconds)
    finally:
        b.close()
    want = oracle.render(b.blob, seconds)
    for v in (0, 33, 69):
        compare(got[v], want, exact, "events voice %d" % v)


def test_poke_rejects_specialised_parameters(lib):
    blob, _ = circuit_blob(circuits.w3_voice())
    p = capi.Program(blob)
    s = capi.Session([p], [2], 0)
    filt = unit_index(blob, "Filter")
    with pytest.raises(capi.UnsupportedCircuit, match="specialised"):
        s.poke(0, -1, ["u%d.f[0]" % filt], [500.0])
    with pytest.raises(capi.UnsupportedCircuit, match="does not hold"):
        s.poke(0, 0, ["u%d.nonexistent" % filt], [1.0])
    s.close()


def test_retriggerer_batch_with_per_voice_rates(lib, oracle, kernel_mode):
    """Retriggerer on the device for a batch (whole-chunk tiles whatever kernel the batch would
    otherwise get), every voice with its own retrigger rate and envelope length, bit-exact."""
    def build(rate, dur):
        env = d.components.Shape("decay", dur).trigger()
        d.components.Retriggerer(env, rate)
        return d.components.Multiply(d.components.Osc(330), env)
    rng = np.random.default_rng(5)
    rates, durs = rng.uniform(1, 30, 80), rng.uniform(0.02, 0.3, 80)
    blob, _ = circuit_blob(build(4.0, 0.1))
    rt, shape = unit_index(blob, "Retriggerer"), unit_index(blob, "Shape")
    p = capi.Program(blob)
    s = capi.Session([p], [80], 0)
    assert s.launch_info()["sub_block"] * s.launch_info()["time_lanes"] == 256 and not s.launch_info()["specialised"]
    s.set_params(0, ["u%d.rate[0]" % rt, "u%d.duration[0]" % shape], np.stack([rates, durs], axis=1))
    got = s.render_host(SR)
    s.close()
    for v in (0, 1, 40, 79):
        vb, _ = circuit_blob(build(float(rates[v]), float(durs[v])))
        compare(got[v], oracle.render(vb, 1.0), True, "retrigger voice %d" % v)


def test_retriggered_circuits_next_to_others_in_one_session(lib, oracle):
    """Circuits with per-chunk control ops stay out of the multi-program launch (its tiles are half chunks)."""
    names = ["bang_retrigger", "osc440", "sporadic_retrigger", "w2_live", "retrigger_modulated"]
    blobs = [circuit_blob(circuits.CASES[n][0]())[0] for n in names]
    progs = [capi.Program(b) for b in blobs]
    s = capi.Session(progs, [1] * len(progs), 0)
    got = s.render_host(SR // 2)
    s.close()
    for i, n in enumerate(names):
        want = oracle.render(blobs[i], 0.5)
        compare(got[i][:len(want)], want, True, n)


def test_graph_edit_for_a_batch_with_per_voice_parameters(lib, oracle, kernel_mode):
    """`then` for a whole batch.  Every voice has its own filter cutoff and noise stream; after the
    edit the adopted session keeps the per-voice constants, the filter state and the delay line of every voice."""
    rng = np.random.default_rng(8)
    fs = rng.uniform(300, 3000, 72)
    seeds = rng.integers(1, 2 ** 62, 72)

    def circuit(f, seed):
        noise = d.components.Noise(seed=int(seed))
        noise.scheduleFinish(0.2)
        out = d.components.Repeater(noise)
        osc = d.components.Multiply(d.components.Osc(150), 0.7)
        noise.onFinish = lambda: setattr(out, "IN", osc)
        ap = d.components.AllPass(0.013, 0.6)
        ap.IN = d.components.Filter(out, f, "LP")
        return ap

    b = d.Batch(circuit(1000.0, 1), 72)
    blob0 = b.blob
    filt, noise = unit_index(blob0, "Filter"), unit_index(blob0, "Noise")
    y0 = np.array([2 * d.config.rand_tick(int(sd), 0) - 1 for sd in seeds])
    b.run.set_params(0, ["u%d.f[0]" % filt, "u%d.seed_lo" % noise, "u%d.seed_hi" % noise, "u%d.y" % noise],
                     np.stack([fs, (seeds & 0xFFFFFFFF).astype(np.float64), (seeds >> 32).astype(np.float64), y0], axis=1))
    try:
        got = b.render(0.5)
    finally:
        b.close()
    for v in (0, 35, 71):
        vb, _ = circuit_blob(circuit(float(fs[v]), seeds[v]))
        compare(got[v], oracle.render(vb, 0.5), True, "graph edit voice %d" % v)


def test_render_stream_crosses_a_graph_edit(lib, oracle):
    """The streaming sink renders block-wise: the edit falls inside a later block."""
    from dusp_b200.render import circuit_blob as cb
    blob, _ = cb(circuits.for_then())
    want = oracle.render(blob, 1.0, raw=True)[0]
    rs = d.RenderStream(circuits.for_then(), 1, block_chunks=32)
    got = np.concatenate([rs.read() for _ in range(172)])
    rs.stop()
    assert np.array_equal(got[:44032], want[:44032])


def test_spectral_batch_with_per_voice_parameters(lib, oracle):
    """70 voices of Osc -> STFT -> HardLowPass -> ISTFT -> ReChunk, each with its own oscillator
    frequency and cut-off (per-voice words); frames are per-voice rings, one warp per voice."""
    rng = np.random.default_rng(1006)
    n = 70
    f = rng.uniform(100, 4000, n)
    cut = rng.uniform(200, 6000, n)
    blob, _ = circuit_blob(circuits.spectral_lowpass())
    osc0 = unit_index(blob, "Osc", 0)
    lp = unit_index(blob, "HardLowPass")
    p = capi.Program(blob)
    s = capi.Session([p], [n], 0)
    s.set_params(0, ["u%d.f[0]" % osc0, "u%d.f[0]" % lp], np.stack([f, cut], axis=1))
    a = s.render_host(19 * 256)   # launches are chunk-aligned, not aligned to the 1024-sample hop
    b = s.render_host(SR // 2 + 6 - 19 * 256)
    s.close()
    got = np.concatenate([a, b], axis=-1)
    for v in (0, 1, 33, 69):
        def mid(x, c=float(cut[v])):
            l = C.HardLowPass(c); l.IN = x
            return l
        u = circuits._stft(C.Sum(C.Osc(float(f[v])), C.Osc(3000.7)), 4096, 1024, mid)
        r = C.ReChunk(1024, 256); r.IN = u
        vb, _ = circuit_blob(r)
        compare(got[v], oracle.render(vb, 0.5 + 6 / SR), True, "spectral voice %d" % v)
