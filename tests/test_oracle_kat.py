"""Pins the CPU oracle against closed-form known answers derived from the
reference source -- the reference itself ships no golden vectors
and cannot run in this image, so these KATs are what the oracle is pinned to."""
import math

import numpy as np
import pytest

import dusp_b200 as d
from dusp_b200 import components as C, patches as P
from dusp_b200.config import rand_tick
from dusp_b200.graphblob import strip_order
from dusp_b200.render import circuit_blob

import circuits

SR = 44100


def sine_table():
    return np.array([math.sin(2 * math.pi * t / 44101) for t in range(44101)], dtype=np.float32)  # waveTables.js:6-8


def render(oracle, outlet, seconds):
    blob, _ = circuit_blob(outlet)
    return oracle.render(blob, seconds)


def test_tables_match_reference_formulas(oracle):
    sin = oracle.wave_table("sin")
    assert np.array_equal(sin, sine_table())
    assert sin[44100] == np.float32(math.sin(2 * math.pi * 44100 / 44101)) and abs(sin[44100] + 1.4247e-4) < 1e-7
    saw = oracle.wave_table("saw")
    assert saw[0] == -1 and saw[44100] == 0 and saw[22050] == np.float32(-1 + 22050 * 2 / 44101)
    sq = oracle.wave_table("square")
    assert sq[22049] == 1 and sq[22050] == -1 and sq[44100] == -1
    tri = oracle.wave_table("triangle")
    assert tri[11025] == 1 and tri[22050] == 0 and tri[33075] == -1 and tri[44100] == 0
    assert np.array_equal(oracle.wave_table("8bit"), (np.floor(sin.astype(np.float64) * 128 + 0.5) / 128).astype(np.float32))
    assert oracle.shape_table("decay")[0] == 1 and oracle.shape_table("decay")[44100] == 0
    assert oracle.shape_table("decaySquared")[22050] == np.float32(0.25)
    assert oracle.shape_table("semiSine")[22050] == 1


def test_osc440_closed_form(oracle):
    """Osc.js:39-45: phase advances BEFORE the read; integer f => exact table read-out."""
    out = render(oracle, circuits.osc440(), 10)
    assert len(out) == 1 and out[0].shape == (441000,)
    k = np.arange(441000, dtype=np.int64)
    expect = sine_table()[(440 * (k + 1)) % SR]
    assert np.array_equal(out[0], expect)
    assert np.allclose(out[0][:4], [0.0626469, 0.1250477, 0.18695726, 0.24813233], atol=1e-8)


def test_readme_fm_is_a_1hz_sine(oracle):
    """from/to are dead properties, duration=2 SAMPLES, never triggered -> f == 1."""
    a = render(oracle, circuits.readme_fm_literal(), 1)[0]
    b = render(oracle, C.Osc(1), 1)[0]
    assert np.array_equal(a, b)
    k = np.arange(SR, dtype=np.int64)
    assert np.array_equal(a, sine_table()[(k + 1) % SR])


def test_w2_literal_is_square_of_1hz(oracle):
    a = render(oracle, circuits.w2_literal(), 0.25)[0]
    o = render(oracle, C.Osc(1), 0.25)[0]
    assert np.array_equal(a, o * o)  # one f32 multiply == JS double multiply + one rounding


def test_sum_pan_shape_and_value(oracle):
    """test/test_renderChannelData.js:4-8: 2 channels x 44100; Pan.js:18-24 centre gain."""
    out = render(oracle, circuits.sum_pan(), 1)
    assert len(out) == 2 and out[0].shape == (SR,)
    tab = sine_table()
    k = np.arange(SR, dtype=np.int64)
    s = (tab[(200 * (k + 1)) % SR].astype(np.float64) + tab[(250 * (k + 1)) % SR].astype(np.float64)).astype(np.float32)
    comp = math.pow(10, 1.5 / 20)
    expect = (s.astype(np.float64) * 1 / 2 * comp).astype(np.float32)
    assert np.array_equal(out[0], expect) and np.array_equal(out[1], expect)


def test_live_ramp_closed_form(oracle):
    r = C.Ramp(1000, 200, 100).trigger()
    out = render(oracle, r, 0.05)[0]
    t = np.minimum(np.arange(1, len(out) + 1), 1000).astype(np.float64)
    assert np.array_equal(out, (200 + (t / 1000) * (100 - 200)).astype(np.float32))
    assert out[-1] == 100


def test_untriggered_ramp_is_constant(oracle):
    out = render(oracle, C.Ramp(2), 0.01)[0]
    assert np.all(out == 1)


def test_noise_uses_injected_stream(oracle):
    n = C.Noise(seed=77)
    out = render(oracle, n, 0.01)[0]
    expect = np.array([2 * rand_tick(77, k + 1) - 1 for k in range(len(out))]).astype(np.float32)
    assert np.array_equal(out, expect)
    assert oracle.rand(77, 5) == rand_tick(77, 5)


def test_noise_sample_and_hold_period(oracle):
    out = render(oracle, C.Noise(11025, seed=5), 0.01)[0]  # new draw every 4th sample
    assert out[0] == out[1] == out[2] == np.float32(2 * rand_tick(5, 0) - 1)
    assert out[3] == np.float32(2 * rand_tick(5, 1) - 1) and out[7] == np.float32(2 * rand_tick(5, 2) - 1)


def test_filter_unset_f_renders_zeros(oracle):
    out = render(oracle, circuits.filter_unset_f(), 0.1)[0]
    assert np.all(out == 0)


def test_fixed_delay_is_pure_delay(oracle):
    fd = C.FixedDelay(100 / SR)
    fd.IN = C.Osc(441)
    out = render(oracle, fd, 0.05)[0]
    src = render(oracle, C.Osc(441), 0.05)[0]
    assert np.all(out[:100] == 0) and np.array_equal(out[100:], src[:-100])


def test_shape_decay_shorthand(oracle):
    out = render(oracle, d.unDusp("D1"), 1.5)[0]
    assert out[0] == np.float32(1 - 1 / SR) and out[SR - 1] == 0 and np.all(out[SR:] == 0)


def test_render_length_semantics(oracle):
    """renderChannelData.js:24-29,44: floor(duration*SR) samples; partial last chunk dropped."""
    assert render(oracle, C.Osc(100), 0.01)[0].shape == (441,)
    assert render(oracle, C.Osc(100), 256 / SR)[0].shape == (256,)
    assert render(oracle, C.Osc(100), 0) == []


def test_simple_delay_tick_order(oracle):
    """the order derived from Unit.js:171-209 + Circuit.js:67-107 for SimpleDelay."""
    sd = P.SimpleDelay(C.Osc(440), 0.25, 0.5, 0.4)
    blob, circuit = circuit_blob(sd)
    kinds = [type(u).__name__ for u in circuit.units]
    assert kinds == ["Osc", "SecondsToSamples", "Repeater", "Delay", "CrossFader", "Multiply", "Sum"]
    assert oracle.schedule(strip_order(blob)) == list(range(7))


def test_simple_delay_feedback_is_one_chunk_stale(oracle):
    """With delay 0 samples the wet path is the dry input (Delay.js:27-38 reads then writes),
    and the feedback edge adds the PREVIOUS chunk's sum: chunk 0 has no feedback at all."""
    sd = P.SimpleDelay(C.Osc(441), 1e-9, 0.5, 1.0)  # DELAY inlet || default would replace 0
    out = render(oracle, sd, 0.02)[0]
    assert np.all(out[:1] == 0)


@pytest.mark.parametrize("name", sorted(circuits.CASES))
def test_host_scheduler_matches_oracle_scheduler(oracle, name):
    fac, _, _ = circuits.CASES[name]
    from dusp_b200.render import split_epochs
    blob, circuit = circuit_blob(fac())
    first = split_epochs(blob)[0][1]  # (later epochs grow by Circuit.add, one unit at a time: host mirror only)
    assert oracle.schedule(strip_order(first)) == list(range(sum(l.startswith("U ") for l in first.splitlines())))


def test_trigger_event_after_the_first_chunk_closed_form(oracle):
    """`x at 0.5` = stop() + scheduleTrigger(0.5) (constructOperation.js:63-69): the event's time is sample 22050,
    so it runs before the tick of the chunk that starts at 22016 (Circuit.js:23,57-65: events with t < clock + 256).
    Until then the stopped Ramp holds y0; from sample 22016 on it counts t = 1, 2, ... (Ramp.js:25-40)."""
    import dusp_b200 as d
    from dusp_b200.render import circuit_blob
    blob, circ = circuit_blob(d.unDusp("([Ramp 44100 400 200] at 0.5)"))
    assert circ.device_events == [(86, 0, "playing", 1.0), (86, 0, "t", 0.0)]
    out = oracle.render(blob, 1.0)[0]
    n = np.arange(44100)
    t = np.clip(n - 22016 + 1, 0, 44100).astype(np.float64)
    want = np.where(n < 22016, 400.0, 400.0 + (t / 44100.0) * (200.0 - 400.0)).astype(np.float32)
    assert np.array_equal(out, want)


def test_retriggerer_closed_form(oracle):
    """`x ! r` = trigger() + new Retriggerer(x, r) (constructOperation.js:71-76).  With r = 44100/512 (exact in
    f32, every partial sum exact) the accumulator reaches the sample rate on the 512th sample, i.e. inside the
    SECOND chunk of every pair; the Retriggerer is chained before its target (Retriggerer.js:31), so the Ramp's
    counter restarts at the first sample of chunks 1, 3, 5, ... and counts on through chunks 2, 4, ... (Ramp.js:25-40)."""
    import dusp_b200 as d
    from dusp_b200.render import circuit_blob
    blob, _ = circuit_blob(d.unDusp("[Ramp 1000 3 1] ! 86.1328125"))
    assert [l for l in blob.splitlines() if l.startswith("TG")] == ["TG 0 1 playing 1", "TG 0 1 t 0"]
    out = oracle.render(blob, 0.1)[0]
    n = np.arange(len(out))
    start = np.where(n < 256, 0, ((n // 256 - 1) // 2 * 2 + 1) * 256)  # sample at which the running count began
    t = np.minimum(n - start + 1, 1000).astype(np.float64)
    want = (3.0 + (t / 1000.0) * (1.0 - 3.0)).astype(np.float32)
    assert np.array_equal(out, want)


def test_sporadic_retriggerer_uses_injected_stream(oracle):
    """SporadicRetrigger.js:12-16: one draw per tick, trigger iff draw < rate[0] * 256 / 44100."""
    from dusp_b200 import components as C, config
    from dusp_b200.render import circuit_blob
    ramp = C.Ramp(131072, 1, 131073).trigger()  # out = 1 + t exactly
    sr = C.SporadicRetriggerer(ramp, 40, seed=77)
    blob, _ = circuit_blob(ramp)
    out = oracle.render(blob, 0.5)[0]
    t, want = 0, []
    for k in range((len(out) + 255) // 256):
        if config.rand_tick(77, k) < float(np.float32(40)) * 256 / 44100:
            t = 0
        want += [1 + t + 1 + i for i in range(256)]
        t += 256
    assert np.array_equal(out, np.array(want[:len(out)], dtype=np.float32))
    assert len(set(np.flatnonzero(np.diff(out) < 0))) > 3  # it did fire a few times


def test_for_then_closed_form(oracle):
    """`(O440 for 0.5) then O220` (constructOperation.js:36-60): the finish event's time is sample 22050, so it runs
    before the tick of chunk 86 (Circuit.js:23,57-65) and hands the Repeater the second oscillator, which joins
    the circuit there with phase 0 (Inlet.js:40-74): from sample 22016 on the output is a 220 Hz table read-out
    that starts from its first sample."""
    import dusp_b200 as d
    from dusp_b200.render import circuit_blob, split_epochs
    blob, _ = circuit_blob(d.unDusp("(O440 for 0.5) then O220"))
    assert [k for k, _ in split_epochs(blob)] == [0, 86]
    out = oracle.render(blob, 1.0)[0]
    tab = oracle.wave_table("sin")
    n = np.arange(44100)
    want = np.where(n < 86 * 256, tab[(440 * (n + 1)) % 44100], tab[(220 * (n - 86 * 256 + 1)) % 44100])
    assert np.array_equal(out, want.astype(np.float32))


def test_then_number_and_downstream_state(oracle):
    """`[Osc f:((100 for 0.25) then 300)]`: the Repeater's inlet becomes the constant 300 at chunk 43; the Osc
    downstream keeps its phase (integer frequencies: every phase is exact)."""
    import dusp_b200 as d
    from dusp_b200.render import circuit_blob
    blob, _ = circuit_blob(d.unDusp("[Osc f:((100 for 0.25) then 300)]"))
    out = oracle.render(blob, 0.5)[0]
    tab = oracle.wave_table("sin")
    n = np.arange(22050)
    n0 = 43 * 256
    phase = np.where(n < n0, 100 * (n + 1), 100 * n0 + 300 * (n - n0 + 1)) % 44100
    assert np.array_equal(out, tab[phase].astype(np.float32))


@pytest.mark.parametrize("n", [2, 4, 8, 16, 64, 256, 1024, 4096])
def test_fft_restatement_matches_the_dft(oracle, n):
    """Third-party arithmetic: spectral/FFT.js:2,16 and IFFT.js:2,12 call the un-vendored package `fft.js`
    (package.json "^4.0.3").  The oracle restates its radix-4 realTransform + completeSpectrum, transform and
    inverseTransform; this pins the restatement against the definition of the DFT (numpy) at the sizes the
    reference uses (config.js:10-14: window 4096) and the small ones its radix-2 / radix-4 base cases hit."""
    rng = np.random.default_rng(n)
    x = rng.standard_normal(n)
    ref = np.fft.fft(x)
    got = oracle.fft(x, "real")
    assert np.max(np.abs((got[0::2] + 1j * got[1::2]) - ref)) <= 1e-12 * max(1.0, float(np.max(np.abs(ref))))
    z = rng.standard_normal(2 * n)
    zc = z[0::2] + 1j * z[1::2]
    ref = np.fft.fft(zc)
    got = oracle.fft(z, "complex")
    assert np.max(np.abs((got[0::2] + 1j * got[1::2]) - ref)) <= 1e-12 * max(1.0, float(np.max(np.abs(ref))))
    got = oracle.fft(z, "inverse")  # fft.js divides by N, like numpy's ifft
    assert np.max(np.abs((got[0::2] + 1j * got[1::2]) - np.fft.ifft(zc))) <= 1e-12


def test_fft_rejects_sizes_fft_js_rejects(oracle):
    from oracle.oracle import OracleError
    for n in (1, 3, 12, 100):
        with pytest.raises(OracleError):
            oracle.fft(np.zeros(n), "real")


def test_shape_finishing_by_itself_switches_inside_that_tick(oracle):
    """`D0.3 then O440`: the envelope's running sum t += 1/duration passes 44100 at sample 13230 (the same sequential
    f64 additions as Shape/index.js:31-32); finish() -> onFinish hands the Repeater the oscillator in the middle of
    Circuit.tick; Circuit.add sorts the oscillator between the Shape and the Repeater, and the loop goes on with it
    (Circuit.js:34-37): the whole chunk that holds the finish sample (51) already is the oscillator's first chunk."""
    import math
    import dusp_b200 as d
    from dusp_b200.render import circuit_blob
    blob, _ = circuit_blob(d.unDusp("D0.3 then O440"))
    out = oracle.render(blob, 0.5)[0]
    inc, t, n = np.float64(1.0) / np.float64(np.float32(0.3)), np.float64(0), 0
    while True:
        t += inc
        if t > 44100:
            break
        n += 1
    assert n == 13230 and "EPOCH %d FIN 0" % (n // 256) in blob
    k = n // 256
    tab = np.array([math.sin(2 * math.pi * i / 44101) for i in range(44101)], dtype=np.float32)
    j = np.arange(len(out) - k * 256, dtype=np.int64)
    assert np.array_equal(out[k * 256:], tab[(440 * (j + 1)) % 44100])
    dec = np.array([1 - i / 44100 for i in range(44101)], dtype=np.float32)  # shapeTables.js "decay"
    tt = np.cumsum(np.full(k * 256, inc))
    fr = tt % 1
    env = (dec[np.ceil(tt).astype(int)].astype(np.float64) * fr + dec[np.floor(tt).astype(int)].astype(np.float64) * (1 - fr))
    assert np.array_equal(out[:k * 256], env.astype(np.float32))


def test_shared_reciprocal_division_is_ieee_division(tmp_path):
    """ops.cuh SharedDiv (the Ramp's t / duration as q + (t - duration q) / duration with one correctly rounded
    reciprocal per call) must equal the hardware's division bit for bit: tests/kat_div.c, 2.6e7 random and hard cases
    here (4.2e8 when run by hand)."""
    import os, subprocess
    here = os.path.dirname(os.path.abspath(__file__))
    exe = str(tmp_path / "kat_div")
    subprocess.check_call(["gcc", "-O2", "-ffp-contract=off", "-o", exe, os.path.join(here, "kat_div.c"), "-lm"])
    r = subprocess.run([exe, "10"], capture_output=True, text=True, timeout=120)
    assert r.returncode == 0 and " 0 mismatches" in r.stdout, r.stdout[-500:]


# ---- the reference's own per-sample statements, executed ----------------------------------------------------------
def _js_statement_to_python(stmt, aliases, helpers):
    """The innermost statement of a stateless unit's _tick (verbatim JavaScript, read from the reference's source at test time: tests/ref_text.py)
    as the source of a Python function of the unit's inlets.  JavaScript's arithmetic operators on Numbers ARE IEEE
    double operations, as Python's on floats; only the surface differs."""
    import re
    s = stmt
    s = re.sub(r"//[^\n]*", "", s)
    s = re.sub(r"[{}]", "", s)
    for local, port in aliases.items():
        s = re.sub(r"\b%s\[t\]" % re.escape(local), "P_" + port, s)
    s = re.sub(r"this\.(\w+)\[(\d)\]\[t\]", r"P_\1\2", s)          # this.out[0][t]
    s = re.sub(r"this\.(\w+)\[c\]\[t\]", r"P_\1", s)
    s = re.sub(r"this\.(\w+)\[t\]", r"P_\1", s)
    s = re.sub(r"\b(out\w*|fOut)\[t\]", r"P_out", s)                 # local names of the output channel
    s = s.replace("config.sampleRate", "44100").replace("Math.abs", "abs").replace("Math.pow", "jspow")
    s = re.sub(r"this\.(\w+)", r"F_\1", s)                           # plain fields: this.sf, this.compensationDB
    s = re.sub(r"\bvar\s+", "", s)
    lines = []
    text = " ".join(l.strip() for l in s.splitlines())
    m = re.match(r"if\((.*?)\)\s*(P_\w+\s*=.*?)\s*else\s*(P_\w+\s*=.*)$", text)   # if(..) a = x else a = y
    if m:
        lines = ["if %s:" % m.group(1), "    " + m.group(2), "else:", "    " + m.group(3)]
    else:
        for part in re.split(r"\s+(?=(?:P_\w+|\w+)\s*=[^=])", text):   # one assignment per line
            part = part.strip()
            if not part:
                continue
            t = re.match(r"(\w+)\s*=\s*(.*?)\s*\?\s*(.*?)\s*:\s*(.*)$", part)   # a = cond ? x : y
            if t:
                part = "%s = (%s) if (%s) else (%s)" % (t.group(1), t.group(3), t.group(2), t.group(4))
            lines.append(part)
    src = "def f(I, F):\n"
    names = sorted(set(re.findall(r"\bP_(\w+)", "\n".join(lines))))
    assigned = set(re.findall(r"(?:^|\s)P_(\w+)\s*=[^=]", "\n".join(lines), re.M))
    for n in names:
        if n not in assigned:
            src += "    P_%s = I[%r]\n" % (n, n)
    for n in sorted(set(re.findall(r"\bF_(\w+)", "\n".join(lines)))):
        src += "    F_%s = F[%r]\n" % (n, n)
    src += "".join("    " + l + "\n" for l in lines)
    outs = [n for n in names if n in assigned]
    src += "    return [%s]\n" % ", ".join("P_" + n for n in outs)
    for h, d in helpers.items():
        src = "def %s(%s):\n    return %s\n" % (h, d["params"], d["expr"].replace("Math.pow", "jspow").replace("Math.abs", "abs")) + src
    return src


def _jspow(a, b):
    import math
    try:
        return math.pow(a, b)
    except (ValueError, OverflowError):  # Math.pow: NaN for a negative base with a fractional exponent, Infinity on overflow
        if a < 0 and b != int(b):
            return float("nan")
        return float("inf") if abs(a) > 1 else 0.0
    except ZeroDivisionError:
        return float("inf")


def test_oracle_matches_the_references_own_tick_statements(oracle):
    """The strongest pin available without a JavaScript engine: the per-sample statements of the stateless units'
    `_tick`s, taken VERBATIM from the reference's source at test time (tests/ref_text.py runs
    tests/golden/make_ref_structure.py on /root/reference; nothing of it is committed; skipped where the reference is absent), are translated token for token into Python (IEEE double arithmetic on both sides),
    run on known inputs -- seeded Noise units, whose samples the test computes itself from the injected generator --
    and compared with the oracle's render of the same unit: bit for bit, or within an ulp of f32 where Math.pow runs."""
    import json, math, os
    import numpy as np
    import dusp_b200 as d
    from dusp_b200 import components as C, config
    from dusp_b200.render import circuit_blob
    import ref_text
    ref = ref_text.reference_text()["tick_expr"]
    n = 512
    checked = 0

    def noise_samples(seed):  # Noise.js:16-28 at the default F = sampleRate: a new draw every sample; draw 0 was the constructor's
        return [float(np.float32(2 * config.rand_tick(seed, k + 1) - 1)) for k in range(n)]
    for name, r in sorted(ref.items()):
        src = _js_statement_to_python(r["stmt"], r["aliases"], r["helpers"])
        env = {"jspow": _jspow, "abs": abs}
        exec(src, env)
        f = env["f"]
        u = getattr(C, name)(*{"FixedMultiply": (0.37,), "Clip": (0.5,), "Pow": (1, 2)}.get(name, ()))
        inputs, seed = {}, 1000
        for inlet in u.inletsOrdered:
            if inlet.noData:
                continue
            seed += 1
            setattr(u, inlet.name.upper(), C.Noise(seed=seed))
            inputs[inlet.name] = noise_samples(seed)
        fields = {k: float(v) for k, v in vars(u).items() if isinstance(v, (int, float)) and not k.startswith("_")}
        blob, _ = circuit_blob(u)
        got = oracle.render(blob, n / 44100)
        want = [f({k: v[t] for k, v in inputs.items()}, fields) for t in range(n)]
        for ch in range(len(want[0])):
            w = np.array([x[ch] for x in want], dtype=np.float64)
            w32 = w.astype(np.float32)
            w32 = np.where(np.isnan(w32) | (w32 == 0), np.float32(0), w32)  # renderChannelData.js:44: x || 0
            g = got[ch][:n]
            if "pow" in src:
                ok = np.isclose(g, w32, rtol=2e-7, atol=0) | (np.isinf(w32) & np.isinf(g)) | (g == w32)
                assert ok.all(), (name, ch, int(np.flatnonzero(~ok)[0]))
            else:
                assert np.array_equal(g, w32), (name, ch, int(np.flatnonzero(g != w32)[0]), src)
        checked += 1
    assert checked == len(ref) >= 19


def test_oracle_matches_the_references_stateful_tick_bodies(oracle):
    """The same for the hot-path units that carry state: the WHOLE `_tick` bodies
    of Osc, Ramp, Shape, Filter (with its LP coefficient function), Noise, FixedDelay, CombFilter, AllPass and Timer,
    verbatim from the reference's source (tests/ref_text.py, read at test time), translated line by line (tests/js2py.py) and
    run chunk after chunk on known inputs; the oracle's render of the same circuit must equal the result bit for bit --
    phase wrap, the biquad's f32-rounded feedback, `x || 0` on undefined history, the delay lines' index arithmetic and
    the sample-and-hold of Noise included.  The sine table is built by the reference's own loop (waveTables.js:3-9)."""
    import json, os, sys
    import numpy as np
    sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
    import js2py
    from js2py import F32, Bag, JSArray
    from dusp_b200 import components as C, config
    from dusp_b200.render import circuit_blob
    import ref_text
    bodies = ref_text.reference_text()["tick_body"]
    SR, CH = 44100, 256

    def fn(name, args="S", tail=""):
        env = dict(js2py.ENV)
        exec(js2py.translate(bodies[name] + tail, "f", args), env)
        return env["f"], env

    def noise(seed, n):  # a Noise unit at the default F: draw k + 1 at sample k, stored as f32
        return np.array([2 * config.rand_tick(seed, k + 1) - 1 for k in range(n)], dtype=np.float64).astype(np.float32)

    def js_out(x):  # renderChannelData.js:44
        x = np.asarray(x, dtype=np.float32)
        return np.where(np.isnan(x) | (x == 0), np.float32(0), x)

    def check(name, unit, run_chunk, n_chunks, channel=0):
        blob, _ = circuit_blob(unit)
        got = oracle.render(blob, n_chunks * CH / SR)[channel][:n_chunks * CH]
        want = np.concatenate([js_out(run_chunk(k)) for k in range(n_chunks)])
        assert np.array_equal(got, want), (name, int(np.flatnonzero(got != want)[0]))

    # ---- Osc (Osc.js:35-47), F driven by noise * 1000: negative phases, wraps, both table neighbours
    sine = fn("sineTable", "S", "\nreturn sineTable")[0](None)
    n_ch = 12
    fsig = (noise(7, n_ch * CH).astype(np.float64) * 1000.0).astype(np.float32)
    tick, _ = fn("Osc")
    S = Bag(out=F32(CH), f=None, phase=0, waveTable=sine)

    def osc_chunk(k):
        S.f = F32(fsig[k * CH:(k + 1) * CH])
        tick(S)
        return S.out.a.copy()
    check("Osc", C.Osc(C.Multiply(C.Noise(seed=7), 1000)), osc_chunk, n_ch)

    # ---- Ramp (Ramp.js:25-40) past its end
    tick, _ = fn("Ramp")
    R = Bag(out=F32(CH), t=0, playing=True, duration=5000, y0=200, y1=100)
    check("Ramp", C.Ramp(5000, 200, 100).trigger(), lambda k: (tick(R), R.out.a.copy())[1], 24)

    # ---- Timer (Timer.js:37-42)
    tick, _ = fn("Timer")
    T = Bag(out=F32(CH), t=0, samplePeriod=1 / SR)
    check("Timer", C.Timer(), lambda k: (tick(T), T.out.a.copy())[1], 8)

    # ---- Shape (Shape/index.js:28-59), decay over 0.05 s: the playing sum, the lerp, the right edge after the finish
    tick, _ = fn("Shape")
    decay = F32([1 - x / SR for x in range(SR + 1)])  # shapeTables.js:3-25: table[x] = func(x / sampleRate), func = 1 - x
    H = Bag(out=F32(CH), t=0, playing=True, duration=F32([0.05] * CH), min=F32([0] * CH), max=F32([1] * CH), leftEdge=0,
            rightEdge="shape", shapeTableData=decay, finished=False)
    H.finish = lambda: setattr(H, "finished", True)
    check("Shape", C.Shape("decay", 0.05).trigger(), lambda k: (tick(H), H.out.a.copy())[1], 12)

    # ---- Filter (Filter.js:27-51 + the LP coefficients :66-75): constant F, then F as a signal (recomputed per sample)
    lp, _ = fn("Filter.LP", "S, f")
    tick, _ = fn("Filter")
    for variant in ("constant", "signal"):
        n_ch = 6
        x = noise(9, n_ch * CH)
        if variant == "constant":
            fs = np.full(n_ch * CH, 1000.0, dtype=np.float32)
            unit = C.Filter(C.Noise(seed=9), 1000, "LP")
        else:
            fs = ((noise(11, n_ch * CH).astype(np.float64) * 300.0).astype(np.float32).astype(np.float64) + 900.0).astype(np.float32)
            unit = C.Filter(C.Noise(seed=9), C.Sum(C.Multiply(C.Noise(seed=11), 300), 900), "LP")
        nan = float("nan")
        Fl = Bag(in_=None, f=None, out=[], lastF=None, x1=JSArray(), x2=JSArray(), y1=JSArray(), y2=JSArray(),
                 a0=nan, a1=nan, a2=nan, b1=nan, b2=nan)   # `this.kind = ...` ran calculateCoefficients(undefined): NaN
        Fl.calculateCoefficients = lambda f, Fl=Fl: lp(Fl, f)

        def filt_chunk(k, Fl=Fl, x=x, fs=fs):
            Fl.in_ = [F32(x[k * CH:(k + 1) * CH])]
            Fl.f = F32(fs[k * CH:(k + 1) * CH])
            tick(Fl)
            return Fl.out[0].a.copy()
        check("Filter " + variant, unit, filt_chunk, n_ch)

    # ---- Noise (Noise.js:16-28): a draw per sample at the default F, sample-and-hold at F = 10 kHz
    for f_hz, seed in ((SR, 21), (10000.0, 22)):
        tick, env = fn("Noise")
        k_draw = [0]

        def rnd(seed=seed, k_draw=k_draw):
            k_draw[0] += 1
            return config.rand_tick(seed, k_draw[0])
        env["M_random"] = rnd
        N = Bag(out=[F32(CH)], f=[F32([f_hz] * CH)], phase=0, y=2 * config.rand_tick(seed, 0) - 1)
        check("Noise %g" % f_hz, C.Noise(f_hz, seed=seed), lambda k, N=N, tick=tick: (tick(N), N.out[0].a.copy())[1], 5)

    # ---- FixedDelay / CombFilter / AllPass (FixedDelay.js:14-20, CombFilter.js:11-17, AllPass.js:8-15)
    for name, cls in (("FixedDelay", C.FixedDelay), ("CombFilter", C.CombFilter), ("AllPass", C.AllPass)):
        n_ch = 8
        x = noise(31, n_ch * CH)
        tick, _ = fn(name)
        D = Bag(in_=None, out=F32(CH), buffer=F32(441), tBuffer=0, feedbackGain=F32([0.6] * CH))
        unit = cls(0.01) if name == "FixedDelay" else cls(0.01, 0.6)
        unit.IN = C.Noise(seed=31)

        def dly_chunk(k, D=D, x=x, tick=tick):
            D.in_ = F32(x[k * CH:(k + 1) * CH])
            tick(D)
            return D.out.a.copy()
        check(name, unit, dly_chunk, n_ch)


def test_oracle_matches_the_reference_bodies_on_the_benchmark_voice(oracle):
    """BASELINE config 2's voice -- Noise -> LP Filter -> CircleBuffer feedback delay -> Pan, the circuit `bench.py`
    times -- run by the reference's OWN code: every unit's `_tick` (Noise, Filter + LP coefficients, the two
    CircleBufferWriters, the CircleBufferReader with CircleBuffer.read / write / mix, Multiply, Sum, Pan) translated from
    the verbatim JavaScript (tests/js2py.py), wired as the reference wires them (a connected inlet reads the producing
    outlet's chunk itself, Inlet.js:53; `mono` ports expose channel 0, Piglet.js:13-17), ticked in the circuit's order
    chunk after chunk.  The oracle's render of the same blob must be bit-identical on both channels."""
    import json, os, sys
    import numpy as np
    sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
    import js2py
    from js2py import F32, Bag, JSArray
    import dusp_b200 as d
    from dusp_b200 import components as C, patches as P, config
    from dusp_b200.render import circuit_blob
    import ref_text
    ref = ref_text.reference_text()
    bodies, exprs = ref["tick_body"], ref["tick_expr"]
    CH, SR = 256, 44100

    def fn(name, args="S"):
        env = dict(js2py.ENV)
        exec(js2py.translate(bodies[name], "f", args), env)
        return env["f"], env

    # the circuit (tests/circuits.w3_voice): seeded noise -> LP -> feedback delay (pre-wiping writer, tap, gain, mixing writer) -> pan
    filt = C.Filter(C.Noise(seed=77), 1234.5, "LP")
    mt = P.MultiTapDelay(1, 1.0, filt)
    tap = mt.addFeedback(0.137, 0.55, 0)
    root = C.Pan(C.Sum(filt, tap), -0.3)
    blob, _ = circuit_blob(root)
    units = list(root.circuit.units)

    cb_read, cb_write, cb_mix = (fn("CircleBuffer." + k, a)[0] for k, a in (("read", "S, c, t"), ("write", "S, c, t, y"), ("mix", "S, c, t, y")))

    class CB:  # CircleBuffer.js:3-35
        def __init__(self, b):
            self.lengthInSamples = b.lengthInSamples
            self.channelData = JSArray(F32(b.lengthInSamples) for _ in range(b.numberOfChannels))
        read = lambda self, c, t: cb_read(self, c, t)
        write = lambda self, c, t, y: cb_write(self, c, t, y)
        mix = lambda self, c, t, y: cb_mix(self, c, t, y)
    buffers = {}
    chunk_of = {}   # outlet -> its SignalChunk's channelData (a list of F32)
    bags, ticks = [], []
    for u in units:
        for o in u.outletsOrdered:
            chunk_of[id(o)] = JSArray(F32(CH) for _ in range(o.numberOfChannels))
    lp = fn("Filter.LP", "S, f")[0]
    for u in units:
        S = Bag()
        for o in u.outletsOrdered:
            setattr(S, o.name, chunk_of[id(o)][0] if o.exposeAsMono else chunk_of[id(o)])
        for i in u.inletsOrdered:
            if i.noData:
                continue
            data = chunk_of[id(i.outlet)] if i.connected else JSArray(F32([v] * CH) for v in i.chunk)
            setattr(S, "in_" if i.name == "in" else i.name, data[0] if i.exposeAsMono else data)
        kind = type(u).__name__
        if kind == "Noise":
            tick, env = fn("Noise")
            k_draw = [0]

            def rnd(seed=u.seed, k_draw=k_draw):
                k_draw[0] += 1
                return config.rand_tick(seed, k_draw[0])
            env["M_random"] = rnd
            S.phase, S.y = u.phase, u.y
        elif kind == "Filter":
            tick = fn("Filter")[0]
            nan = float("nan")
            S.__dict__.update(lastF=None, x1=JSArray(), x2=JSArray(), y1=JSArray(), y2=JSArray(), a0=nan, a1=nan, a2=nan, b1=nan, b2=nan)
            S.calculateCoefficients = lambda f, S=S: lp(S, f)
        elif kind in ("CircleBufferWriter", "CircleBufferReader"):
            tick = fn(kind)[0]
            S._buffer = buffers.setdefault(id(u.buffer), CB(u.buffer))
            S.__dict__.update(t=u.t, numberOfChannels=u.buffer.numberOfChannels, tickInterval=CH,
                              preWipe=bool(getattr(u, "preWipe", False)), postWipe=bool(getattr(u, "postWipe", False)))
        elif kind in ("Multiply", "Sum", "Pan"):
            src = _js_statement_to_python(exprs[kind]["stmt"], exprs[kind]["aliases"], exprs[kind]["helpers"])
            env = {"jspow": _jspow, "abs": abs}
            exec(src, env)
            per_sample = env["f"]
            fields = {"compensationDB": getattr(u, "compensationDB", None)}

            def tick(S, per_sample=per_sample, fields=fields, kind=kind):
                for t in range(CH):
                    if kind == "Pan":   # Pan.js:18-24: mono in / pan, two output channels
                        l, r = per_sample({"in": S.in_[t], "pan": S.pan[t]}, fields)
                        S.out[0][t], S.out[1][t] = l, r
                    else:               # SignalCombiner: channel c of a and b (one channel each here)
                        S.out[0][t] = per_sample({"a": S.a[0][t], "b": S.b[0][t]}, fields)[0]
        else:
            raise AssertionError("unit kind without a translated body: " + kind)
        bags.append(S)
        ticks.append(tick)
    n_chunks = 70   # 0.406 s: the 0.137 s feedback tap comes round twice
    out = [[], []]
    root_chunk = chunk_of[id(root.outletsOrdered[0])]
    for k in range(n_chunks):
        for S, tick in zip(bags, ticks):
            tick(S)
        for c in range(2):
            out[c].append(root_chunk[c].a.copy())
    got = oracle.render(blob, n_chunks * CH / SR)
    for c in range(2):
        w = np.concatenate(out[c])
        w = np.where(np.isnan(w) | (w == 0), np.float32(0), w)
        g = got[c][:n_chunks * CH]
        assert np.array_equal(g, w), (c, int(np.flatnonzero(g != w)[0]))
        assert float(np.max(np.abs(w))) > 0.05


def _ref_runtime_cases():
    import circuits
    return sorted(circuits.CASES)


@pytest.mark.parametrize("name", _ref_runtime_cases())
def test_parity_cases_run_by_the_references_own_code(oracle, name):
    """Every parity case of tests/circuits.py whose units all have a translated body (tests/ref_runtime.py: the
    reference's verbatim `_tick`s run by tests/js2py.py) is rendered by that code and compared with the oracle, bit for
    bit (cases with a libm transcendental on the device keep their 1e-5 class only on the GPU side: here both sides call
    the host's libm).  Cases with events, multi-rate units or kinds without a translated body are skipped."""
    import os, sys
    import numpy as np
    sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
    import circuits, ref_runtime
    import dusp_b200 as d
    from dusp_b200.render import circuit_blob
    d.config.seed(12345)
    fac, seconds, _ = circuits.CASES[name]
    unit = fac()
    blob, _ = circuit_blob(unit)
    if "\nEPOCH" in blob or "WARMUP" in blob:
        pytest.skip("graph edits")
    events = [(int(t[1]), int(t[2]), t[3], float(t[4])) for t in (l.split() for l in blob.splitlines()) if t and t[0] == "EV"]
    # a second, untouched instance for the reference's code (writing the blob runs target.trigger() on the live objects to
    # record what it assigns)
    d.config.seed(12345)
    unit = fac()
    outlet = unit if getattr(unit, "isOutlet", False) else unit.defaultOutlet
    if outlet.unit.circuit is None:
        d.Circuit(outlet.unit)
    try:
        rt = ref_runtime.RefRuntime(outlet.unit)
    except ref_runtime.Unsupported as e:
        pytest.skip("no translated body: %s" % e)
    n_chunks = min(int(seconds * 44100) // 256 + 1, 40)
    rt.apply_initial_props(blob)
    want = rt.render(outlet, n_chunks, events)
    got = oracle.render(blob, n_chunks * 256 / 44100)
    if len(got) != len(want):
        pytest.skip("channel counts grow at run time")
    for c in range(len(want)):
        g = got[c][:n_chunks * 256]
        assert np.array_equal(g, want[c]), (name, c, int(np.flatnonzero(g != want[c])[0]))
