"""Synthetic benchmark workloads W1..W5: deterministic
generators of circuits and per-voice parameters.  All randomness comes from
the injected counter-based generator (``config.rand53``), keyed by a stated
seed and the GLOBAL voice index so any rank can build its own shard."""
import numpy as np

from . import components as C, patches as P, quick
from .config import rand53, rand_tick, _mix64

SR = 44100


def _u(seed, voice, j, lo, hi):
    return lo + (hi - lo) * rand53(seed, 8 * voice + j)


def unit_indices(blob):
    kinds = [l.split()[1] for l in blob.splitlines() if l.startswith("U ")]
    out = {}
    for i, k in enumerate(kinds):
        out.setdefault(k, []).append(i)
    return out


# ---- W1: `[Osc 440]`, 10 s mono, one circuit (BASELINE configs[0]) ----
def w1_circuit():
    from .construct import unDusp
    return unDusp("[Osc 440]")


# ---- W2: FM / envelope ring-mod patch (BASELINE configs[1]) ----
W2_SEED = 1001


def w2_literal_circuit():
    from .construct import unDusp
    return unDusp("[Multiply A:[Osc #o f:[Ramp from:200 to:100 duration:2]] B:#o]")


def w2_live_circuit(f0=300.0):
    osc = C.Osc(C.Ramp(88200, f0, f0 / 2).trigger())
    return C.Multiply(osc, osc)


def w2_params(voice0, n, indices=None):
    """f0_v ~ U[100,800): Ramp(88200, f0, f0/2).trigger().  ``indices``: explicit global voice indices."""
    idx = range(voice0, voice0 + n) if indices is None else indices
    f0 = np.array([_u(W2_SEED, i, 0, 100.0, 800.0) for i in idx])
    return {"f0": f0}


def w2_overrides(blob, params):
    ramp = unit_indices(blob)["Ramp"][0]
    return {"u%d.y0" % ramp: params["f0"], "u%d.y1" % ramp: params["f0"] / 2}


# ---- W3: seeded Noise -> LP Filter -> CircleBuffer feedback delay -> Pan (BASELINE configs[2]) ----
W3_SEED = 1002
W3_UNITS = 8  # Noise, Filter, Writer, Reader, Multiply, Writer2, Sum, Pan


def w3_circuit(f=1000.0, dly=0.2, g=0.5, pan=0.0, seed=1):
    filt = C.Filter(C.Noise(seed=seed), f, "LP")
    mt = P.MultiTapDelay(1, 1.0, filt)
    tap = mt.addFeedback(dly, g, 0)
    return C.Pan(C.Sum(filt, tap), pan)


def w3_params(voice0, n, indices=None):
    """f_v ~ U[200,4000) Hz, d_v ~ U[0.05,0.5) s, g_v ~ U[0.2,0.7), p_v ~ U[-1,1), noise stream seed per voice."""
    v = list(range(voice0, voice0 + n)) if indices is None else list(indices)
    return {
        "f": np.array([_u(W3_SEED, i, 0, 200.0, 4000.0) for i in v]),
        "dly": np.array([_u(W3_SEED, i, 1, 0.05, 0.5) for i in v]),
        "g": np.array([_u(W3_SEED, i, 2, 0.2, 0.7) for i in v]),
        "pan": np.array([_u(W3_SEED, i, 3, -1.0, 1.0) for i in v]),
        "seed": np.array([_mix64(W3_SEED * 0x10001 + i) for i in v], dtype=np.uint64),
    }


def w3_overrides(blob, params):
    u = unit_indices(blob)
    filt, pan, reader = u["Filter"][0], u["Pan"][0], u["CircleBufferReader"][0]
    mult, noise = u["Multiply"][0], u["Noise"][0]
    seeds = params["seed"]
    y0 = np.array([2 * rand_tick(int(s), 0) - 1 for s in seeds])
    return {
        "u%d.f[0]" % filt: params["f"], "u%d.offset[0]" % reader: params["dly"], "u%d.b[0]" % mult: params["g"],
        "u%d.pan[0]" % pan: params["pan"], "u%d.seed_lo" % noise: (seeds & np.uint64(0xFFFFFFFF)).astype(np.float64),
        "u%d.seed_hi" % noise: (seeds >> np.uint64(32)).astype(np.float64), "u%d.y" % noise: y0,
    }


def w3_voice_circuit(params, v):
    """The stand-alone circuit of voice v (what the reference / the oracle renders)."""
    return w3_circuit(float(params["f"][v]), float(params["dly"][v]), float(params["g"][v]),
                      float(params["pan"][v]), int(params["seed"][v]))


# ---- W3b: cyclic-graph coverage, `Pan(SimpleDelay(Filter(Noise)))` ----
def w3b_circuit(f=800.0, dly=0.25, g=0.5, pan=0.3, seed=1):
    return C.Pan(P.SimpleDelay(C.Filter(C.Noise(seed=seed), f), dly, g, 0.4), pan)


# ---- W5: 100 000 voices = W2-live (even global index) + W3 (odd), stereo mixdown only (BASELINE configs[4]) ----
def w5_split(voice0, n):
    """Global voice indices of a shard [voice0, voice0+n), by kind."""
    idx = range(voice0, voice0 + n)
    return [i for i in idx if i % 2 == 0], [i for i in idx if i % 2 == 1]


def w5_voice_circuit(i):
    """The stand-alone circuit of global voice i (what the reference / the oracle renders)."""
    if i % 2 == 0:
        return w2_live_circuit(float(w2_params(0, 0, [i])["f0"][0]))
    return w3_voice_circuit(w3_params(0, 0, [i]), 0)


# ---- W4: 10 000 random Dusp circuits, depth 4-8, Osc/Ramp/Multiply/Sum/filter/delay mix (BASELINE configs[3]) ----
W4_SEED = 1003
W4_MAX_UNITS = 64


class _W4Gen:
    """Recursive generator of circuit `index`: internal node in {Multiply 30 %, Sum 30 %, Filter-LP
    15 %, AllPass / FixedDelay 10 %, Osc with modulated f 15 %}, leaf in {Osc const f 50 %, live Ramp 20 %, Shape
    decay 15 %, Noise 15 %}; at most 64 units.  Builds the unit graph and the equivalent Dusp text side by side."""

    def __init__(self, index):
        self.index, self.k, self.units = index, 0, 0

    def u(self, lo=0.0, hi=1.0):
        self.k += 1
        return lo + (hi - lo) * rand53(W4_SEED, (self.index << 12) + self.k)

    def leaf(self):
        r = self.u()
        self.units += 1
        if r < 0.5:
            f = round(self.u(50.0, 2000.0), 3)
            return C.Osc(f), "[Osc f:%r]" % f
        if r < 0.7:
            dur, y0, y1 = int(self.u(4410, 441000)), round(self.u(0.0, 1.0), 4), round(self.u(0.0, 1.0), 4)
            return C.Ramp(dur, y0, y1).trigger(), "([Ramp %d %r %r] at 0)" % (dur, y0, y1)
        if r < 0.85:
            dur = round(self.u(0.2, 8.0), 3)
            return C.Shape("decay", dur).trigger(), "D%r" % dur
        return C.Noise(), "[Noise]"  # stream seed: the host generator's next Noise seed (config.seed, see w4_circuit)

    def node(self, depth):
        if depth <= 1 or self.units >= W4_MAX_UNITS - 4:
            return self.leaf()
        r = self.u()
        if r < 0.6:
            self.units += 1
            a, ta = self.node(depth - 1)
            b, tb = self.node(depth - 1)
            if r < 0.3:
                return C.Multiply(a, b), "(%s * %s)" % (ta, tb)
            return C.Sum(a, b), "(%s + %s)" % (ta, tb)
        if r < 0.75:
            self.units += 1
            f = round(self.u(100.0, 8000.0), 2)
            a, ta = self.node(depth - 1)
            return C.Filter(a, f, "LP"), "(%s -> LP%r)" % (ta, f)
        if r < 0.85:
            self.units += 1
            dly = round(self.u(0.002, 0.03), 5)
            a, ta = self.node(depth - 1)
            if self.u() < 0.5:
                ap = C.AllPass(dly, round(self.u(0.1, 0.8), 3))
                ap.IN = a
                return ap, "(%s -> [AllPass %r %r])" % (ta, dly, float(ap.inlets["feedbackGain"].constant))
            fd = C.FixedDelay(dly)
            fd.IN = a
            return fd, "(%s -> [FixedDelay %r])" % (ta, dly)
        self.units += 3  # Osc + the Multiply and Sum that scale the modulator
        depth_hz, centre = round(self.u(1.0, 300.0), 2), round(self.u(80.0, 1500.0), 2)
        a, ta = self.node(depth - 1)
        return C.Osc(quick.add(quick.multiply(a, depth_hz), centre)), "[Osc f:(%s * %r + %r)]" % (ta, depth_hz, centre)


def w4_circuit(index, with_text=False):
    """Random circuit `index` of W4 (deterministic in `index`): the rendered unit, optionally with its Dusp text."""
    from . import config
    config.seed(_mix64(W4_SEED * 0x20003 + index))  # Noise units take their stream seeds in construction order
    g = _W4Gen(index)
    depth = 4 + int(g.u() * 5)  # uniform in 4..8
    unit, text = g.node(depth)
    return (unit, text) if with_text else unit
