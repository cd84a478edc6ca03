"""Host-side graph objects: Unit / Patch / Circuit / Inlet / Outlet.

Mirror of the reference's L2 runtime *as a graph description* (SURVEY 8a rows
a2-a4).  These objects never compute audio: they record wiring, constants and
unit state exactly the way the reference's objects do, and
``renderChannelData`` hands the resulting circuit to the CUDA library.

Reference files restated here (host logic only, no DSP):
  src/Piglet.js:5-61, src/Inlet.js:16-93, src/Outlet.js:4-22,
  src/Unit.js:7-209, src/UnitOrPatch.js:9-84, src/Patch.js:6-119,
  src/Circuit.js:4-148, src/Event.js:3-30.
Errors are raised as ``DuspError(str)`` where the reference throws a string.
"""
import math

from . import config


class DuspError(Exception):
    """The reference throws plain strings; the message is kept verbatim."""


def _is_number(x):
    return isinstance(x, (int, float)) and not isinstance(x, bool)


# ----------------------------------------------------------------- ports
class Piglet:
    """src/Piglet.js:5-61.  ``chunk`` models the SignalChunk this port owns:
    a list with one constant value per channel (what the channel is filled
    with) -- outlets only use its length (channel count)."""

    isPiglet = True
    isInlet = False
    isOutlet = False

    def __init__(self, model):
        model = dict(model or {})
        self.name = model.pop("name")
        self.unit = model.pop("unit")
        self.noData = bool(model.pop("noData", False))
        self.measuredIn = model.pop("measuredIn", None)
        self.type = model.pop("type", None)
        nch = model.pop("numberOfChannels", None) or 1
        mono = model.pop("mono", False)
        self.chunkSize = model.pop("chunkSize", None) or config.standardChunkSize  # Piglet.js:10
        self.options = model
        self.sampleRate = config.sampleRate
        if nch == "mono" or mono:  # Piglet.js:13-17
            nch = 1
            self.exposeAsMono = True
        else:
            self.exposeAsMono = False
        self.numberOfChannels = nch
        self.chunk = [0.0] * nch  # new SignalChunk(numberOfChannels, chunkSize)

    @property
    def label(self):  # Piglet.js:63-65
        return self.unit.label + "." + self.name.upper()

    @property
    def circuit(self):
        return self.unit.circuit


class Outlet(Piglet):
    """src/Outlet.js:4-22"""

    isOutlet = True

    def __init__(self, model):
        super().__init__(model)
        self.connections = []

    def disconnect(self):
        for inlet in list(self.connections):
            inlet.disconnect()


class Inlet(Piglet):
    """src/Inlet.js:4-99"""

    isInlet = True

    def __init__(self, model):
        super().__init__(model)
        self.outlet = None
        self.constant = 0
        self.connected = False

    def disconnect(self):  # Inlet.js:16-24
        if self.outlet is not None:
            self.outlet.connections.remove(self)
            self.outlet = None
            self.chunk = [0.0] * self.numberOfChannels
            self.connected = False

    def set(self, val):  # Inlet.js:26-31
        if isinstance(val, (UnitOrPatch, Outlet)):
            self.connect(val)
        else:
            self.setConstant(val)

    def get(self):
        return self.outlet if self.connected else self.constant

    def connect(self, outlet, options=None):  # Inlet.js:40-74
        if outlet is None:
            raise DuspError("connecting " + self.label + " to undefined")
        if isinstance(outlet, UnitOrPatch):
            outlet = outlet.defaultOutlet
        if self.connected:
            self.disconnect()
        self.connected = True
        self.outlet = outlet
        outlet.connections.append(self)
        if self.unit.circuit and outlet.unit.circuit and self.unit.circuit is not outlet.unit.circuit:
            raise DuspError("SHIT: Circuit conflict")
        modified = None
        if self.unit.circuit:
            self.unit.circuit.add(outlet.unit)
            modified = self.unit.circuit
        elif outlet.unit.circuit:
            outlet.unit.circuit.add(self.unit)
            modified = outlet.unit.circuit
        if modified:
            self.unit.computeProcessIndex()
            outlet.unit.computeProcessIndex()
            modified.computeOrders()

    def setConstant(self, value):  # Inlet.js:76-93
        if self.outlet is not None:
            self.disconnect()
        self.constant = value
        vals = list(value) if isinstance(value, (list, tuple)) else [value]
        c = 0
        while c < len(self.chunk) or c < len(vals):
            v = vals[c % len(vals)]
            if c < len(self.chunk):
                self.chunk[c] = float(v)
            else:
                self.chunk.append(float(v))
            c += 1

    @property
    def printValue(self):
        return self.outlet.label if self.outlet else self.constant


# ---------------------------------------------------------------- events
# While an event function runs for the accelerated path (render.collect_events), every attribute a unit is
# assigned is recorded here: the assignments ARE the event as far as the device is concerned (they become
# word pokes between two launches).
_recorder = None


class record_assignments:
    def __enter__(self):
        global _recorder
        _recorder = []
        return _recorder

    def __exit__(self, *exc):
        global _recorder
        _recorder = None
        return False


class Event:
    """src/Event.js:3-30 (``time`` in seconds, ``t`` in samples)."""

    def __init__(self, time, f, unit=None, circuit=None):
        self.t = time * config.sampleRate
        self.function = f
        self.unit = unit
        self.circuit = circuit

    @property
    def time(self):
        return self.t / config.sampleRate

    def run(self):
        subject = self.unit or self.circuit
        rv = self.function(subject)
        if _is_number(rv) and rv > 0:
            return Event(self.time + rv, self.function, self.unit, self.circuit)
        return None


class UnitOrPatch:
    """src/UnitOrPatch.js:3-90 (schedule / scheduleTrigger / finish)."""

    isUnitOrPatch = True
    isUnit = False
    isPatch = False
    _times_used = {}

    def _give_label(self):
        cls = type(self)
        n = UnitOrPatch._times_used.get(cls, 0) + 1
        UnitOrPatch._times_used[cls] = n
        self.label = cls.__name__ + str(n)

    def schedule(self, time, func):  # UnitOrPatch.js:9-23
        if isinstance(time, (list, tuple)):
            for t in time:
                self.schedule(t, func)
            return None
        self.addEvent(Event(time, func, self))
        return self

    def scheduleTrigger(self, t, val=None):  # UnitOrPatch.js:25-33
        if not hasattr(self, "trigger"):
            raise DuspError(self.label + ": cannot call scheduleTrigger because trigger is undefined")
        self.schedule(t, lambda subject: subject.trigger())

    def finish(self):  # UnitOrPatch.js:77-84
        self.finished = True
        if getattr(self, "onFinish", None):
            self.onFinish()

    def scheduleFinish(self, t):
        self.schedule(t, lambda subject: subject.finish())


def _port_sugar_get(obj, name):
    ports = obj.__dict__.get("_sugar")
    if ports is not None and name in ports:
        return ports[name]
    raise AttributeError(name)


class Unit(UnitOrPatch):
    """src/Unit.js:7-304.  Subclasses declare ports in the reference's order
    and set ``KIND`` to the reference constructor name."""

    isUnit = True
    KIND = None
    sampleRate = config.sampleRate
    samplePeriod = 1 / config.sampleRate
    standardChunkSize = config.standardChunkSize

    def __init__(self):
        d = self.__dict__
        d["_sugar"] = {}      # UPPERCASE name -> Inlet / Outlet (Unit.js:56-67,82)
        d["inlets"] = {}
        d["inletsOrdered"] = []
        d["outlets"] = {}
        d["outletsOrdered"] = []
        d["events"] = []
        d["circuit"] = None
        d["tickInterval"] = config.standardChunkSize
        d["finished"] = False
        d["nChains"] = 0
        d["processIndex"] = None
        self._give_label()

    # -- `unit.F = 440` / `unit.F = otherUnit` (Unit.js:59-67) --
    def __setattr__(self, name, val):
        sugar = self.__dict__.get("_sugar")
        if sugar is not None and name in sugar and sugar[name].isInlet:
            inlet = sugar[name]
            if val is None:
                raise DuspError("Passed bad value to " + inlet.label)
            if _recorder is not None:
                _recorder.append((self, "@" + inlet.name, val))
            if _is_number(val) or isinstance(val, (list, tuple)):
                inlet.setConstant(val)
            if isinstance(val, (Outlet, UnitOrPatch)):
                inlet.connect(val)
            return
        if _recorder is not None:
            _recorder.append((self, name, val))
        object.__setattr__(self, name, val)

    def __getattr__(self, name):
        return _port_sugar_get(self, name)

    def addInlet(self, name, options=None):  # Unit.js:48-72
        options = dict(options or {})
        options["name"] = name
        options["unit"] = self
        inlet = Inlet(options)
        self.inlets[name] = inlet
        self.inletsOrdered.append(inlet)
        self._sugar[name.upper()] = inlet
        return inlet

    def addOutlet(self, name, options=None):  # Unit.js:73-86
        options = dict(options or {})
        options["name"] = name
        options["unit"] = self
        outlet = Outlet(options)
        self.outlets[name] = outlet
        self._sugar[name.upper()] = outlet
        self.outletsOrdered.append(outlet)
        return outlet

    def chainAfter(self, unit):  # Unit.js:88-98
        if not getattr(unit, "isUnit", False):
            raise DuspError("chainAfter expects a Unit")
        inlet = self.addInlet("chain" + str(self.nChains), {"noData": True})
        self.nChains += 1
        outlet = unit.addOutlet("chain" + str(unit.nChains), {"noData": True})
        unit.nChains += 1
        inlet.connect(outlet)

    chain = chainAfter

    def chainBefore(self, unit):
        if not getattr(unit, "isUnit", False):
            raise DuspError("chainBefore expects a Unit")
        return unit.chainAfter(self)

    @property
    def inputUnits(self):  # Unit.js:121-132
        out = []
        for inlet in self.inlets.values():
            if not inlet.connected:
                continue
            u = inlet.outlet.unit
            if u not in out:
                out.append(u)
        return out

    @property
    def outputUnits(self):  # Unit.js:133-143
        out = []
        for outlet in self.outlets.values():
            for inlet in outlet.connections:
                if inlet.unit not in out:
                    out.append(inlet.unit)
        return out

    def computeProcessIndex(self, history=None):  # Unit.js:171-209
        history = (history or []) + [self]
        inputs = [u for u in self.inputUnits if u not in history]
        mx = -1
        for u in inputs:
            if u.processIndex is None:
                u.computeProcessIndex(history)
            if u.processIndex > mx:
                mx = u.processIndex
        self.processIndex = mx + 1
        outputs = [u for u in self.outputUnits if u not in history]
        for u in outputs:
            if u.processIndex is None or u.processIndex <= self.processIndex:
                u.computeProcessIndex(history)
        return self.processIndex

    @property
    def defaultInlet(self):
        return self.inletsOrdered[0]

    @property
    def defaultOutlet(self):
        return self.outletsOrdered[0]

    def addEvent(self, ev):  # Unit.js:262-275
        if self.circuit:
            self.circuit.addEvent(ev)
        else:
            for i, e in enumerate(self.events):
                if ev.t < e.t:
                    self.events.insert(i, ev)
                    return
            self.events.append(ev)

    def getOrBuildCircuit(self):
        return self.circuit or Circuit(self)

    def trigger(self):  # Unit.js:293-297
        for u in self.inputUnits:
            u.trigger()

    def stop(self):  # Unit.js:298-302
        for u in self.inputUnits:
            u.stop()

    # -- serialisation hooks (graphblob.py) --
    def blob_props(self):
        """(numeric props, string props) describing unit state for the device."""
        return {}, {}


class Patch(UnitOrPatch):
    """src/Patch.js:6-119"""

    isPatch = True

    def __init__(self):
        d = self.__dict__
        d["_alias"] = {}
        d["inlets"] = {}
        d["outlets"] = {}
        d["inletsOrdered"] = []
        d["outletsOrdered"] = []
        d["units"] = []
        d["finished"] = False
        self._give_label()

    def __setattr__(self, name, val):
        alias = self.__dict__.get("_alias")
        if alias is not None and name in alias:
            piglet = alias[name]
            setattr(piglet.unit, piglet.name.upper(), val)  # Patch.js:41-43
            return
        object.__setattr__(self, name, val)

    def __getattr__(self, name):
        alias = self.__dict__.get("_alias")
        if alias is not None and name in alias:
            piglet = alias[name]
            return getattr(piglet.unit, piglet.name.upper())
        raise AttributeError(name)

    def aliasInlet(self, inlet, name=None):  # Patch.js:24-44
        if isinstance(inlet, UnitOrPatch):
            inlet = inlet.inletsOrdered[0]
        if name is None:
            name = inlet.name
            n = 0
            while name in self.inlets:
                n += 1
                name = inlet.name + str(n)
        self.inlets[name] = inlet
        self.inletsOrdered.append(inlet)
        self._alias[name.upper()] = inlet

    def aliasOutlet(self, outlet, name=None):  # Patch.js:45-65
        if isinstance(outlet, UnitOrPatch):
            outlet = outlet.defaultOutlet
        if name is None:
            name = outlet.name
            n = 0
            while name in self.outlets:
                n += 1
                name = outlet.name + str(n)
        self.outlets[name] = outlet
        self.outletsOrdered.append(outlet)
        self._alias[name.upper()] = outlet

    def alias(self, piglet, name=None):  # Patch.js:66-71
        if piglet.isInlet:
            self.aliasInlet(piglet, name)
        elif piglet.isOutlet:
            self.aliasOutlet(piglet, name)

    @property
    def defaultInlet(self):
        return self.inletsOrdered[0]

    @property
    def defaultOutlet(self):
        return self.outletsOrdered[0]

    def addUnit(self, unit):  # Patch.js:80-88
        if isinstance(unit, UnitOrPatch):
            self.units.append(unit)
            unit.ownerPatch = self

    def addUnits(self, *args):  # Patch.js:90-98
        for a in args:
            if isinstance(a, (list, tuple)):
                for u in a:
                    self.addUnit(u)
            else:
                self.addUnit(a)

    def addEvent(self, ev):  # Patch.js:102-107
        if self.units:
            self.units[0].addEvent(ev)
        else:
            raise DuspError("Could not add event as Patch posseses no units: " + self.label)

    def trigger(self):  # Patch.js:114-119
        for u in self.units:
            if hasattr(u, "trigger"):
                u.trigger()
        return self


# --------------------------------------------------------------- circuit
def v8_sort_by_process_index(units):
    """``units.sort((a,b) => a.processIndex - b.processIndex)`` as V8 runs it
    (TimSort below the 64-element min-run: CountAndMakeRun + binary insertion,
    array-sort.tq).  An undefined processIndex makes the comparator NaN, for
    which both ``order < 0`` and ``order >= 0`` are false -- reproduced here
    because Circuit.add sorts while indices are still being assigned."""
    def cmp(a, b):
        if a.processIndex is None or b.processIndex is None:
            return math.nan
        return a.processIndex - b.processIndex

    w = list(units)
    n = len(w)
    if n < 2:
        return w
    run = 2
    order = cmp(w[1], w[0])
    desc = order < 0
    prev = w[1]
    for idx in range(2, n):
        order = cmp(w[idx], prev)
        if desc:
            if order >= 0:
                break
        elif order < 0:
            break
        prev = w[idx]
        run += 1
    if desc:
        w[:run] = w[:run][::-1]
    for start in range(run, n):
        left, right, pivot = 0, start, w[start]
        while left < right:
            mid = left + ((right - left) >> 1)
            if cmp(pivot, w[mid]) < 0:
                right = mid
            else:
                left = mid + 1
        del w[start]
        w.insert(left, pivot)
    return w


class Circuit:
    """src/Circuit.js:4-148.  Holds the dependency-ordered tick list that the
    device program is compiled from; ``tick`` itself runs on the GPU."""

    def __init__(self, *units):
        self.units = []
        self.tickIntervals = []
        self.clock = 0
        self.events = []
        self.gcdTickInterval = None
        self.rendered = False
        for u in units:
            self.add(u)

    def add(self, unit):  # Circuit.js:67-107
        if unit.circuit and unit.circuit is not self:
            raise DuspError("circuit clash, oh god " + unit.label + "\n" + str(unit.circuit is self))
        if unit in self.units:
            return None
        self.units.append(unit)
        unit.__dict__["circuit"] = self
        if unit.tickInterval not in self.tickIntervals:
            self.tickIntervals.append(unit.tickInterval)
            self.tickIntervals.sort()
        if unit.events:
            for ev in unit.events:
                self.addEvent(ev)
            unit.__dict__["events"] = None
        for u in unit.inputUnits:
            self.add(u)
        for u in unit.outputUnits:
            self.add(u)
        unit.computeProcessIndex()
        self.computeOrders()
        return True

    def addEvent(self, ev):  # Circuit.js:109-121
        ev.circuit = self
        for i, e in enumerate(self.events):
            if ev.t < e.t:
                self.events.insert(i, ev)
                return
        self.events.append(ev)

    def computeOrders(self):  # Circuit.js:125-148
        self.units = v8_sort_by_process_index(self.units)
        g = self.tickIntervals[0]
        for ti in self.tickIntervals[1:]:
            g = math.gcd(g, ti)
        self.gcdTickInterval = g

    def runEvents(self, beforeT):  # Circuit.js:57-65
        while self.events and self.events[0].t < beforeT:
            follow = self.events.pop(0).run()
            if follow:
                self.addEvent(follow)

    @property
    def lastUnit(self):
        return self.units[-1]
