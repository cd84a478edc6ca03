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


def _with_records(text, records):
    assert text.endswith("END\n")
    return text[:-len("END\n")] + "".join(r + "\n" for r in records) + "END\n"


def expand_warmup(timeline):
    """Channel growth at the start of a render.  The reference's tick order is not topological (Unit.js:171-209), so
    a unit can tick before its producer and see that outlet with the channel count of the previous tick: a
    multi-channel signal takes S ticks to reach every unit (Space / ScaryPatch / stereo-detuned FM patches).  The
    library compiles tick k < S from a ``WARMUP k`` record (``dusp_program_warmup_ticks``); here those programs become
    one-chunk epochs in front of the steady program, with ``KEEP`` so each hands every outlet's last chunk on."""
    k0, text = timeline[0]
    try:
        capi.Program(text)
        return timeline
    except capi.UnsupportedCircuit as e:
        if "warm-up" not in str(e):
            raise
    s = capi.Program(_with_records(text, ["WARMUP 0"])).warmup_ticks
    if len(timeline) > 1 and timeline[1][0] <= s:
        raise capi.UnsupportedCircuit("graph edit while channel counts are still growing")
    keep = [] if "\nKEEP\n" in text else ["KEEP"]
    return [(k, _with_records(text, ["WARMUP %d" % k] + keep)) for k in range(s)] + \
        [(s, _with_records(text, ["WARMUP %d" % s]))] + timeline[1:]


class SessionRun:
    """A session driven through its circuits' timelines: chunk-aligned segments, the events of chunk
    k poked into every voice of their program group before the segment that starts at sample 256*k, and a new
    session that adopts the old one's device state wherever an event edits a graph (``EPOCH`` sections).

    ``blobs``: one per program group (``circuit_blob``); ``voices``: voices per group."""

    def __init__(self, blobs, voices, device=0):
        self.device = device
        self.voices = [int(v) for v in voices]
        self.timelines = [expand_warmup(split_epochs(b)) for b in blobs]
        self.epoch = [0] * len(blobs)
        self.programs = [capi.Program(tl[0][1]) for tl in self.timelines]
        self.events = [blob_events(tl[0][1]) for tl in self.timelines]
        self.session = capi.Session(self.programs, self.voices, device)
        self.pos = 0  # samples rendered so far (a multiple of 256 between renders, except after the last one)

    def set_params(self, program_index, names, values):
        self.session.set_params(program_index, names, values)

    def _cuts(self, n):
        cuts = set()
        for tl, evs in zip(self.timelines, self.events):
            for (k, _) in tl[1:]:
                cuts.add(k * 256)
            for tl_epoch in tl:
                for (k, _, _, _) in blob_events(tl_epoch[1]):
                    cuts.add(k * 256)
        return sorted(c for c in cuts if self.pos < c < self.pos + n)

    def _enter(self, pos):
        """What the reference does in runEvents before the tick of the chunk at ``pos``: graph edits, then pokes."""
        k = pos // 256
        switch = [i for i, tl in enumerate(self.timelines)
                  if self.epoch[i] + 1 < len(tl) and tl[self.epoch[i] + 1][0] <= k]
        if switch:
            for i in switch:
                while self.epoch[i] + 1 < len(self.timelines[i]) and self.timelines[i][self.epoch[i] + 1][0] <= k:
                    self.epoch[i] += 1
                text = self.timelines[i][self.epoch[i]][1]
                self.programs[i] = capi.Program(text)
                self.events[i] = blob_events(text)
            new = capi.Session(self.programs, self.voices, self.device)
            try:
                for i in range(len(self.programs)):
                    new.adopt(i, self.session, i)
            except Exception:
                new.close()
                raise
            self.session.close()
            self.session = new
        if pos > 0:
            for i, evs in enumerate(self.events):
                now = [e for e in evs if e[0] == k]
                if now:
                    self.session.poke(i, -1, ["u%d.%s" % (u, name) for (_, u, name, _) in now], [v for (_, _, _, v) in now])

    def render(self, n, out_mode=capi.OUT_VOICES):
        """The next ``n`` samples: [voices][channels][n] or [2][n] (mix)."""
        if self.pos % 256:
            raise DuspError("a render that stopped inside a chunk cannot be continued")
        parts = []
        for b in self._cuts(n) + [self.pos + n]:
            self._enter(self.pos)
            parts.append(self.session.render_host(b - self.pos, out_mode))
            self.pos = b
        if len(parts) == 1:
            return parts[0]
        if out_mode != capi.OUT_MIX:  # an edit may add output channels: earlier segments read as silence there
            nch = max(p.shape[1] for p in parts)
            parts = [p if p.shape[1] == nch else np.pad(p, ((0, 0), (0, nch - p.shape[1]), (0, 0))) for p in parts]
        return np.concatenate(parts, axis=-1)

    def close(self):
        self.session.close()


def render_segments(sess, n, events_by_program, out_mode=capi.OUT_VOICES):
    """``sess.render_host(n)`` from clock 0 with events but no graph edits (see ``SessionRun`` for the general
    driver).  events_by_program: {program_index: [(chunk, unit, property, value)]}."""
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
