"""Test infrastructure: a circuit of the host mirror run by the REFERENCE'S OWN `_tick` code -- the verbatim JavaScript
bodies read from the reference's source at test time (tests/ref_text.py; never committed), translated by
tests/js2py.py -- wired as the reference wires units (a connected inlet reads the producing outlet's SignalChunk itself,
Inlet.js:53; `mono` ports expose channel 0, Piglet.js:13-17; constants fill their chunk, Inlet.js:76-93), ticked in the
circuit's order with the circuit's clock (Circuit.js:34-39), copied out as renderChannelData.js:42-44 does.
Raises Unsupported for unit kinds without a translated body, multi-channel uses of the stateless units, and circuits
with events."""
import json
import os

import numpy as np

import js2py
from js2py import F32, Bag, JSArray

import ref_text

CH, SR = 256, 44100


class _Lazy(dict):
    """The reference's verbatim `_tick` bodies, pulled from its source on first use (tests/ref_text.py; skips the test when
    the reference is not on this machine)."""

    def __init__(self, key):
        super().__init__()
        self.key, self.loaded = key, False

    def _load(self):
        if not self.loaded:
            self.update(ref_text.reference_text()[self.key])
            self.loaded = True

    def __getitem__(self, k):
        self._load()
        return dict.__getitem__(self, k)

    def __contains__(self, k):
        self._load()
        return dict.__contains__(self, k)

    def get(self, k, default=None):
        self._load()
        return dict.get(self, k, default)


BODIES, EXPRS = _Lazy("tick_body"), _Lazy("tick_expr")


class Unsupported(Exception):
    pass


def _fn(name, args="S"):
    env = dict(js2py.ENV)
    exec(js2py.translate(BODIES[name], "f", args), env)
    return env["f"], env


_cache = {}


def fn(name, args="S"):
    if (name, args) not in _cache:
        _cache[(name, args)] = _fn(name, args)
    return _cache[(name, args)]


def _sine_table():
    if "sine" not in _cache:
        env = dict(js2py.ENV)
        exec(js2py.translate(BODIES["sineTable"] + "\nreturn sineTable", "f", "S"), env)
        _cache["sine"] = env["f"](None)
    return _cache["sine"]


def _wave_table(name):  # waveTables.js:11-40 (the sine table comes from the reference's own loop, :3-9)
    if name in ("sin", "sine"):
        return _sine_table()
    if ("wave", name) not in _cache:
        n = SR + 1
        if name == "saw":        # :11-13  sawTable[t] = -1 + t * 2/sawTable.length, t < sampleRate (the last entry stays 0)
            t = F32([-1 + i * 2 / n for i in range(SR)] + [0])
        elif name == "square":   # :25-27  fill(1, 0, sampleRate/2); fill(-1, sampleRate/2, sampleRate+1)
            t = F32([1] * (SR // 2) + [-1] * (n - SR // 2))
        elif name == "triangle":  # :15-23
            t = F32(n)
            q = SR // 4
            for i in range(q):
                t[i] = i / SR * 4
                t[i + q] = 1 - t[i]
                t[i + 2 * q] = -t[i]
                t[i + 3 * q] = -1 + t[i]
            t[SR] = 0
        elif name == "8bit":     # :29-32  Math.round(sample * 2^7) / 2^7 of the sine table
            import math
            t = F32([math.floor(float(x) * 128 + 0.5) / 128 for x in _sine_table().a])
        else:
            raise Unsupported("waveform " + name)
        _cache[("wave", name)] = t
    return _cache[("wave", name)]


def _shape_table(name):  # shapeTables.js:3-38: table[x] = func(x / sampleRate)
    import math
    funcs = {"decay": lambda x: 1 - x, "attack": lambda x: x, "semiSine": lambda x: math.sin(math.pi * x),
             "decaySquared": lambda x: (1 - x) * (1 - x)}
    if ("shape", name) not in _cache:
        _cache[("shape", name)] = F32([funcs[name](x / SR) for x in range(SR + 1)])
    return _cache[("shape", name)]


def _jspow(a, b):
    import math
    try:
        return math.pow(a, b)
    except (ValueError, OverflowError):
        if a < 0 and b != int(b):
            return float("nan")
        return float("inf") if abs(a) > 1 else 0.0
    except ZeroDivisionError:
        return float("inf")


class RefRuntime:
    def __init__(self, root_unit, stmt_translator=None):
        from dusp_b200 import config
        circuit = root_unit.circuit
        self.units = list(circuit.units)
        self.clock = 0
        chunk_of = {}
        for u in self.units:
            if getattr(u, "tickInterval", CH) != CH:
                raise Unsupported("tick interval")
            for o in u.outletsOrdered:
                chunk_of[id(o)] = JSArray(F32(CH) for _ in range(o.numberOfChannels))
        self.chunk_of = chunk_of
        buffers = {}
        cb = [fn("CircleBuffer." + k, a)[0] for k, a in (("read", "S, c, t"), ("write", "S, c, t, y"), ("mix", "S, c, t, y"))]

        class CB:  # CircleBuffer.js:3-35
            def __init__(self, b):
                self.lengthInSamples = b.lengthInSamples
                self.channelData = JSArray(F32(b.lengthInSamples) for _ in range(b.numberOfChannels))
            read = lambda self, c, t: cb[0](self, c, t)
            write = lambda self, c, t, y: cb[1](self, c, t, y)
            mix = lambda self, c, t, y: cb[2](self, c, t, y)
        self.steps = []
        bag_of = {}

        class Target:  # what a Retriggerer holds: the target unit with its own trigger() (the translated method body)
            def __init__(self, unit):
                self.unit = unit
                kind = type(unit).__name__
                if kind + ".trigger" not in BODIES:
                    raise Unsupported("retrigger target " + kind)
                self.f = fn(kind + ".trigger")[0]
            trigger = property(lambda self: (lambda: self.f(bag_of[id(self.unit)])))
        for u in self.units:
            S = Bag()
            bag_of[id(u)] = S
            S.label = u.label
            for o in u.outletsOrdered:
                setattr(S, o.name, chunk_of[id(o)][0] if o.exposeAsMono else chunk_of[id(o)])
            nch_in = {}
            for i in u.inletsOrdered:
                if i.noData:
                    continue
                data = chunk_of[id(i.outlet)] if i.connected else JSArray(F32([v] * CH) for v in i.chunk)
                nch_in[i.name] = len(data)
                setattr(S, "in_" if i.name == "in" else i.name, data[0] if i.exposeAsMono else data)
            kind = type(u).__name__
            takes_clock = False
            if kind == "Noise":
                tick, env = _fn("Noise")  # (its own environment: the injected generator is per unit)
                k_draw = [0]

                def rnd(seed=u.seed, k_draw=k_draw):
                    k_draw[0] += 1
                    return config.rand_tick(seed, k_draw[0])
                env["M_random"] = rnd
                S.phase, S.y = u.phase, u.y
            elif kind == "Osc":
                tick = fn("Osc")[0]
                S.phase, S.waveTable = u.phase, _wave_table(u.waveform)
            elif kind == "MultiChannelOsc":
                tick, takes_clock = fn("MultiChannelOsc", "S, clock")[0], True
                S.phase, S.waveTable = JSArray(np.float64(x) for x in u.phase), _wave_table(u.waveform)
            elif kind == "Ramp":
                tick = fn("Ramp")[0]
                S.__dict__.update(t=u.t, playing=bool(u.playing), duration=u.duration, y0=u.y0, y1=u.y1)
            elif kind == "Timer":
                tick = fn("Timer")[0]
                S.__dict__.update(t=u.t, samplePeriod=1 / SR)
            elif kind == "Shape":
                tick = fn("Shape")[0]
                S.__dict__.update(t=u.t, playing=bool(u.playing), leftEdge=u.leftEdge, rightEdge=u.rightEdge, finished=False,
                                  shapeTableData=_shape_table(u.shape))
                if getattr(u, "onFinish", None) or getattr(u, "_finish_handlers", None):
                    raise Unsupported("finish handler")
                S.finish = lambda S=S: setattr(S, "finished", True)
            elif kind == "Filter":
                tick = fn("Filter")[0]
                coeff = fn("Filter." + u.kind, "S, f")[0]
                nan = float("nan")
                S.__dict__.update(lastF=None, x1=JSArray(), x2=JSArray(), y1=JSArray(), y2=JSArray(), a0=nan, a1=nan, a2=nan, b1=nan, b2=nan)
                S.calculateCoefficients = lambda f, S=S, coeff=coeff: coeff(S, f)
            elif kind in ("FixedDelay", "CombFilter", "AllPass"):
                tick = fn(kind)[0]
                S.__dict__.update(buffer=F32(int(u.delayTimeInSamples)), tBuffer=u.tBuffer)
            elif kind == "MonoDelay":
                tick, takes_clock = fn("MonoDelay", "S, clock")[0], True
                S.buffer = F32(int(u.maxDelay))
            elif kind == "Delay":
                tick, takes_clock = fn("Delay", "S, clock")[0], True
                S.__dict__.update(maxDelay=int(u.maxDelay), buffers=JSArray([F32(int(u.maxDelay))]))
            elif kind == "ReadBackDelay":
                tick, takes_clock = fn("ReadBackDelay", "S, clock")[0], True
                S.__dict__.update(buffer=JSArray(), bufferLength=int(u.bufferLength), tBuffer=u.tBuffer, tickInterval=CH)
            elif kind in ("CircleBufferWriter", "CircleBufferReader"):
                tick = fn(kind)[0]
                S._buffer = buffers.setdefault(id(u.buffer), CB(u.buffer))
                if kind == "CircleBufferReader":  # CircleBufferNode.js:19-27, `set buffer`: OUT grows to the buffer's channel count
                    while len(S.out) < u.buffer.numberOfChannels:
                        S.out.append(F32(CH))
                S.__dict__.update(t=u.t, numberOfChannels=u.buffer.numberOfChannels, tickInterval=CH,
                                  preWipe=bool(getattr(u, "preWipe", False)), postWipe=bool(getattr(u, "postWipe", False)))
            elif kind == "AHD":
                tick = fn("AHD")[0]
                S.__dict__.update(state=np.float64(u.state), playing=bool(u.playing), tickInterval=CH)
                S.t = u.t

                def stop(S=S):  # AHD.js:27-31
                    S.state = 0
                    S.playing = False
                S.stop = stop
            elif kind in ("Retriggerer", "SporadicRetriggerer"):
                tick, env = _fn(kind)
                S.__dict__.update(_target=Target(u.target) if u.target is not None else None, tickInterval=CH)
                if kind == "Retriggerer":
                    S.t = u.t
                else:
                    k_draw = [u.draw - 1]

                    def rnd(seed=u.seed, k_draw=k_draw):  # the injected generator: one draw per tick from draw 0 on
                        k_draw[0] += 1
                        return config.rand_tick(seed, k_draw[0])
                    env["M_random"] = rnd
            elif kind == "SampleRateRedux":
                tick = fn("SampleRateRedux")[0]
                S.val = JSArray([np.float64(0)])
                S.timeSinceLastUpdate = float("inf")
            elif kind in BODIES and kind not in ("Monitor", "GreaterThan", "LessThan", "MultiChannelOsc"):
                # stateless units and the rest: the whole _tick, the unit's plain fields as they are in the mirror
                args = BODIES.get(kind + ".args", "")
                takes_clock = bool(args)
                tick = fn(kind, "S, clock" if takes_clock else "S")[0]
                for k, v in vars(u).items():
                    if isinstance(v, (int, float, str)) and not k.startswith("_") and not hasattr(S, k):
                        setattr(S, k, v)
            else:
                raise Unsupported(kind)
            self.steps.append((S, tick, takes_clock))

    def apply_initial_props(self, blob):
        """State at the first tick as the blob has it (`P` records): events due at t = 0 have run by then (Circuit.js:23)."""
        ui = -1
        for t in (l.split() for l in blob.splitlines()):
            if not t:
                continue
            if t[0] == "U":
                ui += 1
            elif t[0] == "P" and 0 <= ui < len(self.steps):
                S = self.steps[ui][0]
                if hasattr(S, t[1]) and isinstance(getattr(S, t[1]), (bool, int, float, np.floating)):
                    setattr(S, t[1], bool(float(t[2])) if isinstance(getattr(S, t[1]), bool) else float(t[2]))
            elif t[0] == "END":
                break

    def render(self, root_outlet, n_chunks, events=()):
        """events: (chunk, unit index, field, value) -- the `EV` records of the blob: what the circuit's event queue
        assigns before that chunk's tick (Circuit.js:23,57-65), as found by the host mirror."""
        chunk = self.chunk_of[id(root_outlet)]
        out = []
        for k in range(n_chunks):
            for ck, ui, field, value in events:
                if ck == k:
                    S = self.steps[ui][0]
                    setattr(S, field, bool(value) if isinstance(getattr(S, field, None), bool) else value)
            for S, tick, takes_clock in self.steps:
                with np.errstate(all="ignore"):
                    tick(S, self.clock) if takes_clock else tick(S)
            self.clock += CH
            while len(out) < len(chunk):  # renderChannelData.js:38-41: a channel that appears later starts with silence
                out.append([np.zeros(CH, dtype=np.float32) for _ in range(k)])
            for c in range(len(chunk)):
                out[c].append(chunk[c].a.copy())
        res = []
        for c in range(len(out)):
            w = np.concatenate(out[c])
            res.append(np.where(np.isnan(w) | (w == 0), np.float32(0), w))  # renderChannelData.js:44: x || 0
        return res
