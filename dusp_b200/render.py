"""The sink: ``renderChannelData`` (mirror of ``src/renderChannelData.js:5-49``)
and the batch API that renders many independent circuits / voices per launch.
"""
import math

import numpy as np

from . import capi, config
from .graph import Circuit, DuspError, Outlet, UnitOrPatch
from .graphblob import circuit_to_blob


class ChannelData(list):
    """``Array<Float32Array>`` with a ``sampleRate`` property (renderChannelData.js:47)."""
    sampleRate = config.sampleRate


def _resolve_outlet(outlet):  # renderChannelData.js:12-17
    if outlet is None or outlet is False or outlet == 0:
        raise DuspError("renderAudioBuffer expects an outlet")
    if isinstance(outlet, UnitOrPatch):
        outlet = outlet.defaultOutlet
    if not isinstance(outlet, Outlet):
        raise DuspError("renderAudioBuffer expects an outlet")
    return outlet


EVENT_HORIZON_S = 600.0  # self-rescheduling events are expanded up to this render time
_IGNORED_EVENT_PROPS = {"finished"}  # assigned by trigger()/finish(); no effect without onFinish graph edits


def _fold_early_events(circuit):
    """Events the reference fires before the first tick (Circuit.js:23,57-65:
    ``runEvents(clock + 256)`` with clock 0) are applied to unit state on the host."""
    circuit.runEvents(circuit.clock + circuit.gcdTickInterval)


def collect_events(circuit, horizon_seconds=EVENT_HORIZON_S):
    """Later events (scheduleTrigger / `at`, schedule(t, f)): the reference runs an event's
    function before the tick of the chunk that contains its time (Circuit.js:23,57-65).  Here every pending
    event is run ONCE on the host mirror, in the reference's order, with all unit attribute assignments
    recorded: [(chunk, unit_index, property, value)].  The device applies them as word pokes between the
    launches of chunk-aligned segments; the oracle applies them before ticking that chunk.  Event functions
    may only ASSIGN numeric unit state (trigger / stop do); graph edits stay outside the accelerated path."""
    from . import graph
    index = {id(u): i for i, u in enumerate(circuit.units)}
    chunk = circuit.gcdTickInterval
    horizon = horizon_seconds * config.sampleRate
    out = []
    guard = 0
    while circuit.events and circuit.events[0].t < horizon:
        ev = circuit.events.pop(0)
        k = max(0, int(math.floor(ev.t / chunk)))
        with graph.record_assignments() as rec:
            follow = ev.run()
        for unit, name, val in rec:
            if name in _IGNORED_EVENT_PROPS or name.startswith("_"):
                continue
            if id(unit) not in index:
                raise capi.UnsupportedCircuit("event touches a unit outside the rendered circuit: " + unit.label)
            if name.startswith("@") or isinstance(val, bool) or not isinstance(val, (int, float)):
                if isinstance(val, bool):
                    val = 1.0 if val else 0.0
                else:
                    raise capi.UnsupportedCircuit("event at t=%g assigns %s.%s: only numeric unit state can change "
                                                  "during an accelerated render" % (ev.t, unit.label, name.lstrip("@")))
            out.append((k, index[id(unit)], name, float(val)))
        if follow is not None:
            circuit.addEvent(follow)
        guard += 1
        if guard > 200000:
            raise capi.UnsupportedCircuit("more than 200000 events inside the event horizon")
    return out


def _numeric_assignments(rec, index, what):
    out = []
    for unit, name, val in rec:
        if name in _IGNORED_EVENT_PROPS or name.startswith("_"):
            continue
        if id(unit) not in index:
            raise capi.UnsupportedCircuit(what + " touches a unit outside the rendered circuit: " + unit.label)
        if isinstance(val, bool):
            val = 1.0 if val else 0.0
        elif name.startswith("@") or not isinstance(val, (int, float)):
            raise capi.UnsupportedCircuit("%s assigns %s.%s: only numeric unit state can change during an "
                                          "accelerated render" % (what, unit.label, name.lstrip("@")))
        out.append((index[id(unit)], name, float(val)))
    return out


def collect_triggers(circuit):
    """(Sporadic)Retriggerer units (Retriggerer.js:17-18): ``target.trigger()`` is run ONCE on the
    host mirror with its assignments recorded: [(retriggerer_index, unit_index, property, value)].  The device
    and the oracle replay them whenever the retriggerer fires."""
    from . import graph
    index = {id(u): i for i, u in enumerate(circuit.units)}
    out = []
    for i, u in enumerate(circuit.units):
        if u.KIND not in ("Retriggerer", "SporadicRetriggerer"):
            continue
        target = u.target
        if target is None or not hasattr(target, "trigger"):
            continue
        with graph.record_assignments() as rec:
            target.trigger()
        out += [(i,) + a for a in _numeric_assignments(rec, index, "trigger() of " + target.label)]
    return out


def circuit_blob(outlet, horizon_seconds=EVENT_HORIZON_S):
    """(blob, circuit) for a rendered outlet -- `outlet.unit.circuit || new Circuit(outlet.unit)`.
    Events after the first chunk are appended to the blob as `EV` records (the oracle applies them) and kept
    in ``circuit.device_events`` for the device path (``render_segments``)."""
    outlet = _resolve_outlet(outlet)
    circuit = outlet.unit.circuit or Circuit(outlet.unit)  # renderChannelData.js:20
    if circuit.rendered:
        raise DuspError("this circuit has already been rendered: renders are one-shot (SURVEY section 5)")
    _fold_early_events(circuit)
    blob = circuit_to_blob(circuit, outlet)
    events = collect_events(circuit, horizon_seconds)
    circuit.device_events = events
    lines = ["EV %d %d %s %.17g" % e for e in events]
    lines += ["TG %d %d %s %.17g" % t for t in collect_triggers(circuit)]
    if lines:
        blob = blob[:-len("END\n")] + "\n".join(lines) + "\nEND\n"
    return blob, circuit


def blob_events(blob):
    """[(chunk, unit, property, value)] of a blob's EV records."""
    ev = []
    for l in blob.splitlines():
        if l.startswith("EV "):
            _, k, u, name, v = l.split()
            ev.append((int(k), int(u), name, float(v)))
    return ev


def render_segments(sess, n, events_by_program, out_mode=capi.OUT_VOICES):
    """``sess.render_host(n)`` with events: chunk-aligned segments, the events of chunk k poked into every
    voice of their program group before the segment that starts at sample 256*k.  events_by_program:
    {program_index: [(chunk, unit, property, value)]}."""
    chunk = 256
    cuts = sorted({k * chunk for evs in events_by_program.values() for (k, _, _, _) in evs if 0 < k * chunk < n})
    if not cuts:
        return sess.render_host(n, out_mode)
    pending = {pi: sorted(evs, key=lambda e: e[0]) for pi, evs in events_by_program.items()}  # stable: keeps run order
    parts, pos = [], 0
    for b in cuts + [n]:
        for pi, evs in pending.items():
            now = [e for e in evs if e[0] * chunk == pos and pos > 0]
            if now:
                sess.poke(pi, -1, ["u%d.%s" % (u, name) for (_, u, name, _) in now], [v for (_, _, _, v) in now])
        parts.append(sess.render_host(b - pos, out_mode))
        pos = b
    return np.concatenate(parts, axis=-1)


def n_samples_for(duration):
    length = duration * config.sampleRate  # renderChannelData.js:24
    return max(int(length), 0) if math.isfinite(length) else 0


def renderChannelData(outlet, duration=1, TypedArray=np.float32, device=0):
    """Render ``duration`` seconds of ``outlet`` (an Outlet, Unit or Patch) on the
    GPU.  Returns a ``ChannelData`` list with one float32 array per channel."""
    blob, circuit = circuit_blob(outlet)
    prog = capi.Program(blob)
    sess = capi.Session([prog], [1], device)
    try:
        n = n_samples_for(duration)
        out = render_segments(sess, n, {0: circuit.device_events})
    finally:
        sess.close()
    circuit.rendered = True
    res = ChannelData(np.asarray(out[0, c], dtype=TypedArray) for c in range(prog.num_channels))
    return res


class Batch:
    """Many voices of ONE circuit shape rendered in one launch.  ``overrides``
    maps parameter names ("u<unit>.<inlet>[<ch>]" / "u<unit>.<prop>", see
    ``Program.params()``) to one value per voice."""

    def __init__(self, outlet, n_voices, overrides=None, device=0):
        self.blob, self.circuit = circuit_blob(outlet)
        self.program = capi.Program(self.blob)
        self.session = capi.Session([self.program], [n_voices], device)
        self.n_voices = n_voices
        if overrides:
            names = list(overrides)
            vals = np.stack([np.asarray(overrides[k], dtype=np.float64) for k in names], axis=1)
            if vals.shape[0] != n_voices:
                raise DuspError("override arrays must hold one value per voice")
            self.session.set_params(0, names, vals)

    def render(self, duration, out_mode=capi.OUT_VOICES):
        return render_segments(self.session, n_samples_for(duration), {0: self.circuit.device_events}, out_mode)

    def close(self):
        self.session.close()


def render_batch(outlets, duration=1, device=0, out_mode=capi.OUT_VOICES):
    """Render a list of independent circuits in one session (one program group
    each).  Returns [circuit][channel][n] (zero-padded to the widest circuit)."""
    progs, events = [], {}
    for i, o in enumerate(outlets):
        blob, circuit = circuit_blob(o)
        progs.append(capi.Program(blob))
        events[i] = circuit.device_events
        circuit.rendered = True
    sess = capi.Session(progs, [1] * len(progs), device)
    try:
        return render_segments(sess, n_samples_for(duration), events, out_mode)
    finally:
        sess.close()
