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
# host-side bookkeeping an event may assign: finish flags (acted on by the host mirror itself, which runs onFinish)
# and what Circuit.add / computeProcessIndex write while an edit adds units (Circuit.js:67-107, Unit.js:171-209)
_IGNORED_EVENT_PROPS = {"finished", "onFinish", "processIndex", "circuit", "events", "promises", "nChains"}


def _fold_early_events(circuit):
    """Events the reference fires before the first tick (Circuit.js:23,57-65:
    ``runEvents(clock + 256)`` with clock 0) are applied to unit state on the host."""
    circuit.runEvents(circuit.clock + circuit.gcdTickInterval)


def _structure(circuit):
    """What a graph edit changes: the unit list and every inlet's source / constant."""
    sig = []
    for u in circuit.units:
        ins = []
        for inlet in u.inletsOrdered:
            if inlet.connected:
                src = inlet.outlet
                ins.append((inlet.name, src.unit.label, src.name))
            else:
                ins.append((inlet.name, tuple(inlet.chunk)))
        sig.append((u.label, tuple(ins)))
    return tuple(sig)


def _numeric_assignments(rec, what):
    """[(unit, property, value)] of a recorded run: numeric unit state only.  Inlet assignments are graph edits
    (picked up by the structure comparison in ``collect_timeline``) and are skipped here."""
    out = []
    for unit, name, val in rec:
        if name in _IGNORED_EVENT_PROPS or name.startswith("_") or name.startswith("@"):
            continue
        if isinstance(val, bool):
            val = 1.0 if val else 0.0
        elif not isinstance(val, (int, float)):
            raise capi.UnsupportedCircuit("%s assigns %s.%s: only numeric unit state and inlets can change during an "
                                          "accelerated render" % (what, unit.label, name))
        out.append((unit, name, float(val)))
    return out


def collect_timeline(circuit, outlet, horizon_seconds=EVENT_HORIZON_S):
    """Later events.  The reference runs an event's function before the tick of the chunk that
    contains its time (Circuit.js:23,57-65).  Here every pending event is run ONCE on the host mirror, in the
    reference's order, and what it did is recorded:

    * numeric unit-state assignments (trigger / stop / schedule(t, f)) -> ``(chunk, unit, property, value)``,
      applied as word pokes between the launches of chunk-aligned segments (the oracle: before that chunk's tick);
    * graph edits (`then` hands an inlet a new source, an event assigns an inlet; Inlet.js:40-93) -> a new EPOCH:
      the circuit as it stands after the edit, from that chunk on.  Units are matched by label across epochs:
      those that were there keep their device state, new ones start from their construction state.

    Returns [(start_chunk, graph_text, units, [(chunk, unit_index, property, value)])]."""
    from . import graph
    chunk = circuit.gcdTickInterval
    horizon = horizon_seconds * config.sampleRate
    epochs = [[0, circuit_to_blob(circuit, outlet), list(circuit.units)]]
    pending = []
    sig = _structure(circuit)
    guard = k_now = 0
    while circuit.events and circuit.events[0].t < horizon:
        ev = circuit.events.pop(0)
        # (an event queued in the past by a unit that joined the circuit late runs in the same runEvents pass)
        k = k_now = max(k_now, int(math.floor(ev.t / chunk)))
        before = {id(u) for u in circuit.units}
        with graph.record_assignments() as rec:
            follow = ev.run()
        rec = [r for r in rec if id(r[0]) in before]  # units the event constructs arrive with the snapshot
        pending += [(k,) + a for a in _numeric_assignments(rec, "event at t=%g" % ev.t)]
        now = _structure(circuit)
        if now != sig:
            sig = now
            snap = [k, circuit_to_blob(circuit, outlet), list(circuit.units)]
            if k == epochs[-1][0] and len(epochs) > 1:
                epochs[-1] = snap
            elif k == 0:
                raise capi.UnsupportedCircuit("graph edit before the first chunk")
            else:
                epochs.append(snap)
        if follow is not None:
            circuit.addEvent(follow)
        guard += 1
        if guard > 200000:
            raise capi.UnsupportedCircuit("more than 200000 events inside the event horizon")
    if len(epochs) > 1:
        for _, _, units in epochs:
            labels = [u.label for u in units]
            if len(set(labels)) != len(labels):
                raise capi.UnsupportedCircuit("graph edits need unique unit labels (units are matched by label)")
    out = []
    for e, (k0, text, units) in enumerate(epochs):
        k1 = epochs[e + 1][0] if e + 1 < len(epochs) else None
        index = {id(u): i for i, u in enumerate(units)}
        evs = []
        for (k, unit, name, val) in pending:
            if k < k0 or (k1 is not None and k >= k1):
                continue
            if id(unit) not in index:
                raise capi.UnsupportedCircuit("event touches a unit outside the rendered circuit: " + unit.label)
            evs.append((k, index[id(unit)], name, val))
        out.append((k0, text, units, evs))
    return out


def collect_triggers(units):
    """(Sporadic)Retriggerer units (Retriggerer.js:17-18): ``target.trigger()`` is run ONCE on the
    host mirror with its assignments recorded: [(retriggerer_index, unit_index, property, value)].  The device
    and the oracle replay them whenever the retriggerer fires."""
    from . import graph
    index = {id(u): i for i, u in enumerate(units)}
    out = []
    for i, u in enumerate(units):
        if u.KIND not in ("Retriggerer", "SporadicRetriggerer"):
            continue
        target = u.target
        if target is None or not hasattr(target, "trigger"):
            continue
        with graph.record_assignments() as rec:
            target.trigger()
        for unit, name, val in _numeric_assignments(rec, "trigger() of " + target.label):
            if id(unit) not in index:
                raise capi.UnsupportedCircuit("trigger() of %s touches a unit outside the rendered circuit: %s"
                                              % (target.label, unit.label))
            out.append((i, index[id(unit)], name, val))
    return out


def circuit_blob(outlet, horizon_seconds=EVENT_HORIZON_S):
    """(blob, circuit) for a rendered outlet -- `outlet.unit.circuit || new Circuit(outlet.unit)`.
    Events after the first chunk are appended as `EV` records, retriggerer targets as `TG` records, and every
    graph edit starts a new `EPOCH <chunk>` section holding the circuit after the edit (graphblob.py).  The
    epochs are also kept in ``circuit.timeline`` = [(start_chunk, text, events)]."""
    outlet = _resolve_outlet(outlet)
    circuit = outlet.unit.circuit or Circuit(outlet.unit)  # renderChannelData.js:20
    if circuit.rendered:
        raise DuspError("this circuit has already been rendered: renders are one-shot (SURVEY section 5)")
    _fold_early_events(circuit)
    parts, timeline = [], []
    epochs = collect_timeline(circuit, outlet, horizon_seconds)
    trig = [collect_triggers(units) for (_, _, units, _) in epochs]
    for e, ((k0, text, units, evs), tgs) in enumerate(zip(epochs, trig)):
        lines = ["EV %d %d %s %.17g" % ev for ev in evs] + ["TG %d %d %s %.17g" % t for t in tgs]
        if e + 1 < len(epochs):
            lines.append("KEEP")  # an edit follows: the program keeps every outlet's last chunk for its successor
        if lines:
            text = text[:-len("END\n")] + "\n".join(lines) + "\nEND\n"
        parts.append(("EPOCH %d\n" % k0 if k0 else "") + text)
        timeline.append((k0, text, evs))
    circuit.timeline = timeline
    circuit.device_events = timeline[0][2]
    _reject_mid_tick_finish(circuit, {id(u): u for (_, _, units, _) in epochs for u in units}.values())
    return "".join(parts), circuit


def _reject_mid_tick_finish(circuit, units):
    """A Shape calls ``finish()`` from inside its ``_tick`` when its envelope runs out (Shape/index.js:40-42): an
    ``onFinish`` that edits the graph or assigns unit state would then act in the middle of ``Circuit.tick``, at a
    data-dependent sample.  That is outside the accelerated path; an ``onFinish`` that only raises finish flags
    (Boop.js:15-17) is fine.  Dry run on the host mirror, after all snapshots have been taken."""
    from . import graph
    for u in units:
        if u.KIND != "Shape" or not callable(getattr(u, "onFinish", None)):
            continue
        sig = _structure(circuit)
        with graph.record_assignments() as rec:
            u.onFinish()
        if _structure(circuit) != sig or _numeric_assignments(rec, "onFinish of " + u.label):
            raise capi.UnsupportedCircuit("onFinish of %s would act from inside a tick when the envelope runs out "
                                          "(Shape/index.js:40-42): outside the accelerated render path" % u.label)


def split_epochs(blob):
    """[(start_chunk, graph_text)] of a blob (one entry unless events edit the graph)."""
    out, k, cur = [], 0, []
    for l in blob.splitlines():
        if l.startswith("EPOCH "):
            out.append((k, "\n".join(cur) + "\n"))
            k, cur = int(l.split()[1]), []
        else:
            cur.append(l)
    out.append((k, "\n".join(cur) + "\n"))
    return out


def blob_events(blob):
    """[(chunk, unit, property, value)] of a blob's EV records."""
    ev = []
    for l in blob.splitlines():
        if l.startswith("EV "):
            _, k, u, name, v = l.split()
            ev.append((int(k), int(u), name, float(v)))
    return ev


class SessionRun:
    """A render driven through its circuits' timelines.  The protocol itself -- chunk-aligned
    segments, the events of chunk k poked into every voice before the segment that starts at sample 256*k, a new
    session that adopts the old one's device state wherever an event edits a graph (``EPOCH`` sections), warm-up
    programs while channel counts still grow -- lives behind the C ABI (``dusp_run_*``, csrc/run.cc); this is the
    host-side handle.

    ``blobs``: one per program group (``circuit_blob``); ``voices``: voices per group."""

    def __init__(self, blobs, voices, device=0):
        self.run = capi.Run(blobs, voices, device)
        self.device = device
        self.voices = [int(v) for v in voices]

    @property
    def session(self):
        return self.run.session

    @property
    def programs(self):
        return self.run.programs

    @property
    def pos(self):
        return self.run.pos

    def set_params(self, program_index, names, values):
        self.run.set_params(program_index, list(names), values)

    def render(self, n, out_mode=capi.OUT_VOICES):
        """The next ``n`` samples: [voices][channels][n] or [2][n] (mix)."""
        return self.run.render_host(n, out_mode)

    def close(self):
        self.run.close()


def n_samples_for(duration):
    length = duration * config.sampleRate  # renderChannelData.js:24
    return max(int(length), 0) if math.isfinite(length) else 0


def renderChannelData(outlet, duration=1, TypedArray=np.float32, device=0):
    """Render ``duration`` seconds of ``outlet`` (an Outlet, Unit or Patch) on the
    GPU.  Returns a ``ChannelData`` list with one float32 array per channel."""
    blob, circuit = circuit_blob(outlet)
    run = SessionRun([blob], [1], device)
    try:
        out = run.render(n_samples_for(duration))
    finally:
        run.close()
    circuit.rendered = True
    res = ChannelData(np.asarray(out[0, c], dtype=TypedArray) for c in range(out.shape[1]))
    return res


class Batch:
    """Many voices of ONE circuit shape rendered in one launch.  ``overrides``
    maps parameter names ("u<unit>.<inlet>[<ch>]" / "u<unit>.<prop>", see
    ``Program.params()``) to one value per voice."""

    def __init__(self, outlet, n_voices, overrides=None, device=0):
        self.blob, self.circuit = circuit_blob(outlet)
        self.run = SessionRun([self.blob], [n_voices], device)
        self.program = self.run.programs[0]
        self.n_voices = n_voices
        if overrides:
            names = list(overrides)
            vals = np.stack([np.asarray(overrides[k], dtype=np.float64) for k in names], axis=1)
            if vals.shape[0] != n_voices:
                raise DuspError("override arrays must hold one value per voice")
            self.run.set_params(0, names, vals)

    @property
    def session(self):
        return self.run.session

    def render(self, duration, out_mode=capi.OUT_VOICES):
        return self.run.render(n_samples_for(duration), out_mode)

    def close(self):
        self.run.close()


def render_batch(outlets, duration=1, device=0, out_mode=capi.OUT_VOICES):
    """Render a list of independent circuits in one session (one program group
    each).  Returns [circuit][channel][n] (zero-padded to the widest circuit)."""
    blobs = []
    for o in outlets:
        blob, circuit = circuit_blob(o)
        blobs.append(blob)
        circuit.rendered = True
    run = SessionRun(blobs, [1] * len(blobs), device)
    try:
        return run.render(n_samples_for(duration), out_mode)
    finally:
        run.close()
