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


def _fold_early_events(circuit):
    """Events the reference fires before the first tick (Circuit.js:23,57-65:
    ``runEvents(clock + 256)`` with clock 0) are applied to unit state on the
    host; later events would need run-time graph edits."""
    circuit.runEvents(circuit.clock + circuit.gcdTickInterval)
    if circuit.events:
        raise capi.UnsupportedCircuit("event scheduled at t >= one chunk (%g samples): outside the accelerated path"
                                      % circuit.events[0].t)


def circuit_blob(outlet):
    """(blob, circuit) for a rendered outlet -- `outlet.unit.circuit || new Circuit(outlet.unit)`."""
    outlet = _resolve_outlet(outlet)
    circuit = outlet.unit.circuit or Circuit(outlet.unit)  # renderChannelData.js:20
    if circuit.rendered:
        raise DuspError("this circuit has already been rendered: renders are one-shot (SURVEY section 5)")
    _fold_early_events(circuit)
    return circuit_to_blob(circuit, outlet), circuit


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
        out = sess.render_host(n, capi.OUT_VOICES)
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
        return self.session.render_host(n_samples_for(duration), out_mode)

    def close(self):
        self.session.close()


def render_batch(outlets, duration=1, device=0, out_mode=capi.OUT_VOICES):
    """Render a list of independent circuits in one session (one program group
    each).  Returns [circuit][channel][n] (zero-padded to the widest circuit)."""
    progs = []
    for o in outlets:
        blob, circuit = circuit_blob(o)
        progs.append(capi.Program(blob))
        circuit.rendered = True
    sess = capi.Session(progs, [1] * len(progs), device)
    try:
        return sess.render_host(n_samples_for(duration), out_mode)
    finally:
        sess.close()
