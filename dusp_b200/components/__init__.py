"""Hot-path units as host-side graph nodes.

Each class mirrors the reference constructor of the same name: same port
names in the same declaration order, same defaults (``x || default``), same
methods that edit state before rendering (``trigger``, ``stop``, ``waveform``,
``setSeconds`` ...).  There is no ``_tick`` here -- the per-sample arithmetic of
every unit lives in ``csrc/render_sm100.cu`` (and, for the tests, in the
oracle).  ``blob_props`` serialises the state the device kernels start from.
"""
from .. import config
from ..config import js_or
from ..graph import Unit, DuspError

SR = config.sampleRate

WAVEFORMS = ["sin", "sine", "saw", "square", "triangle", "8bit"]  # waveTables.js:33-40
SHAPES = ["decay", "attack", "semiSine", "decaySquared"]          # shapeTables.js:20-38


def _js_round(x):
    import math
    return math.floor(x + 0.5)


# ------------------------------------------------------------ oscillators
class Osc(Unit):
    """src/components/Osc/Osc.js:6-66"""
    KIND = "Osc"

    def __init__(self, f=None, waveform=None):
        super().__init__()
        self.addInlet("f", {"mono": True, "measuredIn": "Hz"})
        self.addOutlet("out", {"mono": True, "type": "audio"})
        self.F = js_or(f, 440)
        self.phase = 0
        self.waveform = js_or(waveform, "sin")

    @property
    def waveform(self):
        return self._waveform

    @waveform.setter
    def waveform(self, waveform):  # Osc.js:52-62
        if waveform == "random":
            waveform = WAVEFORMS[int(config.host_random.random() * len(WAVEFORMS))]
        if waveform not in WAVEFORMS:
            raise DuspError("waveform doesn't exist: " + str(waveform))
        object.__setattr__(self, "_waveform", waveform)

    def randomPhaseFlip(self):  # Osc.js:63-66
        if config.host_random.random() < 0.5:
            self.phase += SR / 2

    def blob_props(self):
        return {"phase": self.phase}, {"waveform": self._waveform}


class MultiChannelOsc(Osc):
    """src/components/Osc/MultiChannelOsc.js:7-62 (multi-channel f/out, phase per channel)"""
    KIND = "MultiChannelOsc"

    def __init__(self, f=None, waveform=None):
        Unit.__init__(self)
        self.addInlet("f", {"measuredIn": "Hz"})
        self.addOutlet("out", {"type": "audio"})
        self.F = js_or(f, 440)
        self.phase = []
        self.waveform = js_or(waveform, "sin")

    def resetPhase(self):  # :52-55 -- element assignments: as an event it zeroes every channel that exists by then
        from .. import graph
        object.__setattr__(self, "phase", [0 for _ in self.phase])
        if graph._recorder is not None:
            graph._recorder.append((self, "phase[*]", 0))

    def randomPhaseFlip(self):  # :57-61 -- iterates an empty phase array before the first tick
        if config.host_random.random() < 0.5:
            self.phase = [p + SR / 2 for p in self.phase]

    def blob_props(self):
        return {}, {"waveform": self._waveform}


# -------------------------------------------------------------- envelopes
class Ramp(Unit):
    """src/components/Ramp.js:3-40 (duration in SAMPLES; inert until trigger())"""
    KIND = "Ramp"

    def __init__(self, duration=None, y0=None, y1=None):
        super().__init__()
        self.addOutlet("out", {"mono": True, "type": "control"})
        self.duration = js_or(duration, self.sampleRate)
        self.y0 = js_or(y0, 1)
        self.y1 = js_or(y1, 0)
        self.t = 0
        self.playing = False

    def trigger(self):
        self.playing = True
        self.t = 0
        return self

    def blob_props(self):
        return {"duration": self.duration, "y0": self.y0, "y1": self.y1, "t": self.t,
                "playing": int(self.playing)}, {}


class Shape(Unit):
    """src/components/Shape/index.js:7-99"""
    KIND = "Shape"

    def __init__(self, shape=None, durationInSeconds=None, min=None, max=None):
        super().__init__()
        self.addInlet("duration", {"mono": True, "type": "time", "measuredIn": "s"})
        self.addInlet("min", {"mono": True, "type": "scalar"})
        self.addInlet("max", {"mono": True, "type": "scalar"})
        self.addOutlet("out", {"mono": True, "type": "control", "min": 0, "max": 1})
        self.t = 0
        self.playing = False
        self.leftEdge = 0
        self.rightEdge = "shape"
        self.shape = js_or(shape, "decay")
        self.DURATION = js_or(durationInSeconds, 1)
        self.MIN = js_or(min, 0)
        self.MAX = js_or(max, 1)

    def __setattr__(self, name, val):
        if name == "shape":  # index.js:103-109
            if val not in SHAPES:
                raise DuspError(self.label + ":\n\tinvalid shape function: " + str(val))
            object.__setattr__(self, "_shape", val)
        else:
            super().__setattr__(name, val)

    @property
    def shape(self):
        return self._shape

    dusp_flags = {"trigger": lambda self: self.trigger()}  # index.js:62-67

    @staticmethod
    def randomShapeStr():  # index.js:137-140
        return SHAPES[int(config.host_random.random() * len(SHAPES))]

    @staticmethod
    def randomInRange(maxDuration=None, minMin=None, maxMax=None):  # index.js:114-135
        maxDuration = js_or(maxDuration, 1)
        r = config.host_random.random
        nan = float("nan")
        lo = nan if minMin is None else minMin
        span = nan if (maxMax is None or minMin is None) else maxMax - minMin
        a = lo + r() * span
        b = lo + r() * span
        mn, mx = (b, a) if a > b else (a, b)
        return Shape(Shape.randomShapeStr(), r() * maxDuration, mn, mx)

    class _RandomDecay:
        """``Shape.randomDecay(maxDuration)`` (static, index.js:142-147: a NEW decaySquared
        Shape) and ``shape.randomDecay(maxDuration)`` (method, :149-154: re-randomise this one)."""

        def __get__(self, obj, cls):
            r = config.host_random.random
            if obj is None:
                return lambda maxDuration=None, *_: Shape("decaySquared", r() * js_or(maxDuration, 5))

            def method(maxDuration=None):
                obj.shape = "decay"
                obj.DURATION = r() * js_or(maxDuration, 5)
                obj.MIN = 0
                obj.MAX = 1
            return method

    randomDecay = _RandomDecay()

    def trigger(self):
        self.playing = True
        self.t = 0
        return self

    def stop(self):
        self.playing = False

    def blob_props(self):
        p = {"t": self.t, "playing": int(self.playing), "finished": int(self.finished)}
        s = {"shape": self._shape}
        for nm in ("leftEdge", "rightEdge"):
            v = getattr(self, nm)
            if isinstance(v, str):
                s[nm] = v
            else:
                p[nm] = v
        return p, s


class AHD(Unit):
    """src/components/AHD.js:6-76"""
    KIND = "AHD"

    def __init__(self, attack=None, hold=None, decay=None):
        super().__init__()
        self.addInlet("attack", {"mono": True, "type": "time", "measuredIn": "s"})
        self.addInlet("hold", {"mono": True, "type": "time", "measuredIn": "s"})
        self.addInlet("decay", {"mono": True, "type": "time", "measuredIn": "s"})
        self.addOutlet("out", {"mono": True, "type": "control", "min": 0, "max": 1})
        self.ATTACK = js_or(attack, 0)
        self.HOLD = js_or(hold, 0)
        self.DECAY = js_or(decay, 0)
        self.state = 0
        self.playing = False
        self.t = 0

    def trigger(self):
        self.state = 1
        self.playing = True
        return self

    def stop(self):
        self.state = 0
        self.playing = False
        return self

    def blob_props(self):
        return {"state": self.state, "playing": int(self.playing), "t": self.t}, {}


class Timer(Unit):
    """src/components/Timer.js:25-46"""
    KIND = "Timer"

    def __init__(self):
        super().__init__()
        self.addOutlet("out", {"mono": True})
        self.t = 0

    def trigger(self):
        self.t = 0

    def blob_props(self):
        return {"t": self.t}, {}


class Noise(Unit):
    """src/components/Noise.js:3-28.  ``Math.random`` is replaced by the
    injected counter-based generator: the unit owns a 64-bit stream ``seed``;
    draw 0 initialises ``y`` (ctor, :10), draws 1.. happen in ``_tick``."""
    KIND = "Noise"

    def __init__(self, f=None, seed=None):
        super().__init__()
        self.addInlet("f", {"measuredIn": "Hz"})
        self.addOutlet("out", {"type": "audio"})
        self.F = js_or(f, Unit.sampleRate)
        self.phase = 0
        self.seed = config.host_random.next_noise_seed() if seed is None else int(seed) & ((1 << 64) - 1)
        self.draw = 1
        self.y = config.rand_tick(self.seed, 0) * 2 - 1

    def reseed(self, seed):
        self.seed = int(seed) & ((1 << 64) - 1)
        self.draw = 1
        self.y = config.rand_tick(self.seed, 0) * 2 - 1
        return self

    def blob_props(self):
        return {"phase": self.phase, "seed_lo": self.seed & 0xFFFFFFFF, "seed_hi": self.seed >> 32,
                "draw": self.draw, "y": self.y}, {}


# ---------------------------------------------------------------- filters
class Filter(Unit):
    """src/components/Filter.js:5-104 (2nd-order Butterworth, LP/HP)"""
    KIND = "Filter"

    def __init__(self, input=None, f=None, kind=None):
        super().__init__()
        self.addInlet("in", {"type": "audio"})
        self.addInlet("f", {"mono": True, "measuredIn": "Hz"})
        self.addOutlet("out", {"type": "audio"})
        if js_or(input, None) is not None:
            self.IN = input
        if js_or(f, None) is not None:
            self.F = f
        self.kind = js_or(kind, "LP")

    def __setattr__(self, name, val):
        if name == "kind":  # Filter.js:56-64
            if val not in ("LP", "HP", "BP", "BR"):
                raise DuspError("invalid filter type: " + str(val))
            object.__setattr__(self, "_kind", val)
        else:
            super().__setattr__(name, val)

    @property
    def kind(self):
        return self._kind

    def blob_props(self):
        return {}, {"kind": self._kind}


# ------------------------------------------------------------- arithmetic
class _AB(Unit):
    DEFAULT = 0

    def __init__(self, a=None, b=None):
        super().__init__()
        self.addInlet("a")
        self.addInlet("b")
        self.addOutlet("out")
        self.A = js_or(a, self.DEFAULT)
        self.B = js_or(b, self.DEFAULT)


class Multiply(_AB):
    """src/components/Multiply.js:4-34"""
    KIND = "Multiply"
    DEFAULT = 1


class Divide(_AB):
    """src/components/Divide.js:3-27"""
    KIND = "Divide"
    DEFAULT = 1


class SignalCombiner(_AB):
    """src/components/SignalCombiner.js:3-33"""
    KIND = "Sum"


class Sum(SignalCombiner):
    """src/components/Sum.js:5-44"""
    KIND = "Sum"

    @staticmethod
    def many(inputs):  # Sum.js:18-29
        if len(inputs) == 1:
            return inputs[0]
        sums = [Sum(inputs[0], inputs[1])]
        for i in range(2, len(inputs)):
            sums.append(Sum(sums[i - 2], inputs[i]))
        return sums[-1]


class Subtract(_AB):
    """src/components/Subtract.js:4-29"""
    KIND = "Subtract"


class Pow(Unit):
    """src/components/Pow.js:3-32 (no defaults: undefined operands throw)"""
    KIND = "Pow"

    def __init__(self, a=None, b=None):
        super().__init__()
        self.addInlet("a")
        self.addInlet("b")
        self.addOutlet("out")
        self.A = a
        self.B = b


class ConcatChannels(_AB):
    """src/components/ConcatChannels.js:3-30"""
    KIND = "ConcatChannels"


class _In(Unit):
    DEFAULT = 0

    def __init__(self, input=None):
        super().__init__()
        self.addInlet("in")
        self.addOutlet("out")
        self.IN = js_or(input, self.DEFAULT)


class PolarityInvert(_In):
    """src/components/PolarityInvert.js:3-22"""
    KIND = "PolarityInvert"


class Abs(_In):
    """src/components/Abs.js:3-24"""
    KIND = "Abs"


class DecibelToScaler(Unit):
    """src/components/DecibelToScaler.js:3-19"""
    KIND = "DecibelToScaler"

    def __init__(self, input=None):
        super().__init__()
        self.addInlet("in", {"measuredIn": "dB"})
        self.addOutlet("out")
        self.IN = js_or(input, 0)


class SemitoneToRatio(_In):
    """src/components/SemitoneToRatio.js:3-22"""
    KIND = "SemitoneToRatio"
    DEFAULT = 69


class MidiToFrequency(Unit):
    """src/components/MidiToFrequency.js:3-21"""
    KIND = "MidiToFrequency"

    def __init__(self, midi=None):
        super().__init__()
        self.addInlet("midi", {"type": "midi"})
        self.addOutlet("frequency", {"measuredIn": "Hz"})
        self.MIDI = js_or(midi, 69)


class SecondsToSamples(Unit):
    """src/components/SecondsToSamples.js:4-20"""
    KIND = "SecondsToSamples"

    def __init__(self):
        super().__init__()
        self.addInlet("in", {"measuredIn": "s"})
        self.addOutlet("out", {"measuredIn": "samples"})


class Repeater(Unit):
    """src/components/Repeater.js:3-30"""
    KIND = "Repeater"

    def __init__(self, val=None, measuredIn=None):
        super().__init__()
        self.addInlet("in", {"measuredIn": measuredIn})
        self.addOutlet("out", {"measuredIn": measuredIn})
        self.measuredIn = measuredIn
        self.IN = js_or(val, 0)


class SampleRateRedux(Unit):
    """src/components/SampleRateRedux.js:3-37 (sample-and-hold every ``ammount`` samples)"""
    KIND = "SampleRateRedux"

    def __init__(self, input=None, ammount=None):
        super().__init__()
        self.addInlet("in")
        self.addInlet("ammount", {"mono": True})
        self.addOutlet("out")
        self.val = [0]
        self.timeSinceLastUpdate = float("inf")
        self.IN = js_or(input, 0)
        self.AMMOUNT = js_or(ammount, 0)

    def blob_props(self):
        return {"timeSinceLastUpdate": self.timeSinceLastUpdate}, {}


class Monitor(Unit):
    """src/components/Monitor.js:3-16 -- ``console.log(this.in)`` once per tick: no outlet, no effect on any
    signal.  Kept for graph compatibility; it cannot be rendered (no outlets) and the device ignores it."""
    KIND = "Monitor"

    def __init__(self, input=None):
        super().__init__()
        self.addInlet("in")
        self.IN = input


class FixedMultiply(Unit):
    """src/components/FixedMultiply.js:3-22"""
    KIND = "FixedMultiply"

    def __init__(self, sf=None, input=None):
        super().__init__()
        self.addInlet("in", {"mono": True})
        self.addOutlet("out", {"mono": True})
        self.sf = sf
        self.IN = js_or(input, 0)

    def blob_props(self):
        return {"sf": float("nan") if self.sf is None else self.sf}, {}


class Gain(Unit):
    """src/components/Gain.js:3-28"""
    KIND = "Gain"

    def __init__(self, gain=None):
        super().__init__()
        self.addInlet("in")
        self.addInlet("gain", {"mono": True, "measuredIn": "dB"})
        self.addOutlet("out")
        self.GAIN = js_or(gain, 0)


class Clip(Unit):
    """src/components/Clip.js:3-23 (THRESHOLD = undefined throws, as in the reference)"""
    KIND = "Clip"

    def __init__(self, threshold=None):
        super().__init__()
        self.addInlet("in")
        self.addInlet("threshold")
        self.addOutlet("out")
        self.THRESHOLD = threshold


class _InThreshold(Unit):
    def __init__(self, input=None, threshold=None):
        super().__init__()
        self.addInlet("in")
        self.addInlet("threshold")
        self.addOutlet("out")
        self.IN = js_or(input, 0)
        self.THRESHOLD = js_or(threshold, 0)


class HardClipAbove(_InThreshold):
    """src/components/HardClipAbove.js:3-26"""
    KIND = "HardClipAbove"


class HardClipBelow(_InThreshold):
    """src/components/HardClipBelow.js:3-26"""
    KIND = "HardClipBelow"


class Rescale(Unit):
    """src/components/Rescale.js:3-37"""
    KIND = "Rescale"

    def __init__(self, inLower=None, inUpper=None, outLower=None, outUpper=None):
        super().__init__()
        for nm in ("in", "inLower", "inUpper", "outLower", "outUpper"):
            self.addInlet(nm)
        self.addOutlet("out")
        self.IN = 0
        self.INLOWER = js_or(inLower, -1)
        self.INUPPER = js_or(inUpper, 1)
        self.OUTLOWER = js_or(outLower, 0)
        self.OUTUPPER = js_or(outUpper, 1)


class CrossFader(Unit):
    """src/components/CrossFader.js:3-29"""
    KIND = "CrossFader"

    def __init__(self, a=None, b=None, dial=None):
        super().__init__()
        self.addInlet("a", {"type": "audio"})
        self.addInlet("b", {"type": "audio"})
        self.addInlet("dial", {"mono": True, "min": 0, "max": 1, "zero": 0.5})
        self.addOutlet("out", {"type": "audio"})
        self.A = js_or(a, 0)
        self.B = js_or(b, 0)
        self.DIAL = js_or(dial, 0)


class Pan(Unit):
    """src/components/Pan.js:3-28"""
    KIND = "Pan"

    def __init__(self, input=None, pan=None):
        super().__init__()
        self.addInlet("in", {"mono": True, "type": "audio"})
        self.addInlet("pan", {"mono": True, "min": -1, "max": 1})
        self.addOutlet("out", {"numberOfChannels": 2, "type": "audio"})
        self.PAN = js_or(pan, 0)
        self.IN = js_or(input, 0)
        self.compensationDB = 1.5

    def blob_props(self):
        return {"compensationDB": self.compensationDB}, {}


class PickChannel(Unit):
    """src/components/PickChannel.js:3-21"""
    KIND = "PickChannel"

    def __init__(self, input=None, c=None):
        super().__init__()
        self.addInlet("in")
        self.addInlet("c", {"mono": True})
        self.addOutlet("out", {"mono": True})
        self.IN = js_or(input, 0)
        self.C = js_or(c, 0)


class VectorMagnitude(Unit):
    """src/components/VectorMagnitude.js:5-28"""
    KIND = "VectorMagnitude"

    def __init__(self):
        super().__init__()
        self.addInlet("in")
        self.addOutlet("out", {"mono": True})
        self.IN = [0, 0]


# ----------------------------------------------------------------- delays
class Delay(Unit):
    """src/components/Delay.js:6-43 (write-ahead interpolating delay, 5 s ring per channel)"""
    KIND = "Delay"

    def __init__(self, input=None, delay=None, maxDelay=None):
        super().__init__()
        self.addInlet("in")
        self.addInlet("delay", {"measuredIn": "samples"})
        self.addOutlet("out")
        self.maxDelay = js_or(maxDelay, Unit.sampleRate * 5)
        self.IN = js_or(input, 0)
        self.DELAY = js_or(delay, 4410)

    def blob_props(self):
        return {"maxDelay": int(self.maxDelay)}, {}


class MonoDelay(Unit):
    """src/components/MonoDelay.js:3-30"""
    KIND = "MonoDelay"

    def __init__(self, input=None, delay=None):
        super().__init__()
        self.addInlet("in", {"mono": True, "type": "audio"})
        self.addInlet("delay", {"mono": True, "measuredIn": "samples"})
        self.addOutlet("out", {"mono": True, "type": "audio"})
        self.maxDelay = Unit.sampleRate * 5
        self.IN = js_or(input, 0)
        self.DELAY = js_or(delay, 4410)


class ReadBackDelay(Unit):
    """src/components/ReadBackDelay.js:4-43"""
    KIND = "ReadBackDelay"

    def __init__(self, input=None, delay=None, bufferLength=None):
        super().__init__()
        self.addInlet("in")
        self.addInlet("delay", {"measuredIn": "samples"})
        self.addOutlet("out")
        self.bufferLength = js_or(bufferLength, config.sampleRate)
        self.tBuffer = 0
        self.IN = js_or(input, 0)
        self.DELAY = js_or(delay, 0)

    def blob_props(self):
        return {"bufferLength": int(self.bufferLength), "tBuffer": self.tBuffer}, {}


class FixedDelay(Unit):
    """src/components/FixedDelay.js:3-38"""
    KIND = "FixedDelay"

    def __init__(self, delayTime=None):
        super().__init__()
        self.addInlet("in", {"mono": True, "type": "audio"})
        self.addOutlet("out", {"mono": True, "type": "audio"})
        self.setSeconds(delayTime)
        self.tBuffer = 0

    def setDelayTime(self, tSamples):  # :22-28
        if tSamples is None or not tSamples or tSamples != tSamples or tSamples < 0.5:
            raise DuspError("Cannot have fixed delay of length 0 samples")
        self.delayTimeInSamples = _js_round(tSamples)
        self.delayTimeInSeconds = tSamples / self.sampleRate

    def setSeconds(self, duration):
        self.setDelayTime(None if duration is None else duration * self.sampleRate)

    def setFrequency(self, f):
        self.setSeconds(1 / f)

    def blob_props(self):
        return {"delayTimeInSamples": int(self.delayTimeInSamples), "tBuffer": self.tBuffer}, {}


class CombFilter(FixedDelay):
    """src/components/CombFilter.js:3-26"""
    KIND = "CombFilter"

    def __init__(self, delayTime=None, feedbackGain=None):
        super().__init__(delayTime)
        self.addInlet("feedbackGain", {"mono": True, "type": "scalar"})
        self.FEEDBACKGAIN = js_or(feedbackGain, 0)


class AllPass(CombFilter):
    """src/components/AllPass.js:3-51"""
    KIND = "AllPass"

    @staticmethod
    def random(maxDelayTime=None, maxFeedbackGain=None):  # :20-25
        r = config.host_random.random
        return AllPass(js_or(maxDelayTime, 1) * r(), js_or(maxFeedbackGain, 1) * r())

    @staticmethod
    def manyRandom(n, maxDelay, maxFeedback):  # :27-38 (the `delay` local is computed and unused)
        r = config.host_random.random
        out = []
        for _ in range(n):
            delay = 2 / SR + r() * (maxDelay - 2 / SR)  # `this.sampleRate` is undefined on the class: NaN, unused
            while delay == 0:
                delay = r() * maxDelay
            out.append(AllPass(r() * maxDelay, r() * maxFeedback))
        return out


# ----------------------------------------------------- CircleBuffer taps
class CircleBuffer:
    """src/CircleBuffer.js:3-35 -- device-resident ring; host object only carries its shape."""

    def __init__(self, numberOfChannels=None, lengthInSeconds=None):
        import math
        self.numberOfChannels = js_or(numberOfChannels, 1)
        self.lengthInSeconds = lengthInSeconds
        self.sampleRate = config.sampleRate
        self.lengthInSamples = math.ceil(self.lengthInSeconds * self.sampleRate)


class CircleBufferNode(Unit):
    """src/components/CircleBufferNode.js:7-31"""

    def __init__(self, buffer=None, offset=None):
        super().__init__()
        self.t = 0
        self._buffer = None
        if buffer is not None:
            self.buffer = buffer
        self.addInlet("offset", {"measuredIn": "s"})
        self.OFFSET = js_or(offset, 0)

    def __setattr__(self, name, val):
        if name == "buffer":  # set buffer(): grow OUT to the buffer's channel count (:19-27)
            out = self.__dict__.get("_sugar", {}).get("OUT")
            if out is not None and out.isOutlet:
                while len(out.chunk) < val.numberOfChannels:
                    out.chunk.append(0.0)
            object.__setattr__(self, "_buffer", val)
        else:
            super().__setattr__(name, val)

    @property
    def buffer(self):
        return self._buffer


class CircleBufferReader(CircleBufferNode):
    """src/components/CircleBufferReader.js:3-27"""
    KIND = "CircleBufferReader"

    def __init__(self, buffer=None, offset=None):
        super().__init__(None, offset)
        self.addOutlet("out")
        self.buffer = buffer
        self.postWipe = False

    def blob_props(self):
        return {"t": self.t, "postWipe": int(bool(self.postWipe))}, {}


class CircleBufferWriter(CircleBufferNode):
    """src/components/CircleBufferWriter.js:3-27"""
    KIND = "CircleBufferWriter"

    def __init__(self, buffer=None, offset=None):
        super().__init__(buffer, offset)
        self.addInlet("in")
        self.preWipe = False

    def blob_props(self):
        return {"t": self.t, "preWipe": int(bool(self.preWipe))}, {}


class Retriggerer(Unit):
    """src/components/Retriggerer.js:3-35.  No outlet: accumulates ``rate`` and calls ``target.trigger()`` whenever
    the sum crosses the sample rate; chained before its target so the trigger lands before the target's tick.
    On the device the decision is taken once per 256-sample chunk (every call inside one tick assigns the same
    constants) and ``target.trigger()``'s assignments travel with the graph as ``TG`` records (render.py)."""
    KIND = "Retriggerer"

    def __init__(self, target=None, rate=None):
        super().__init__()
        self.addInlet("rate", {"mono": True, "type": "frequency"})
        object.__setattr__(self, "_target", None)
        if target:
            self.target = target
        self.t = 0
        self.RATE = js_or(rate, 1)

    @property
    def target(self):
        return self._target

    @target.setter
    def target(self, target):
        if self._target is not None:
            raise DuspError("re-targeting a retriggerer needs Unit.unChain, which the reference does not define")
        if target:
            object.__setattr__(self, "_target", target)
            self.chainBefore(target)  # throws for a Patch, like the reference (Patch has no chainAfter)

    def blob_props(self):
        return {"t": self.t}, {}


class SporadicRetriggerer(Retriggerer):
    """src/components/SporadicRetrigger.js:3-32.  One ``Math.random()`` draw per tick, replaced by the unit's own
    counter-based stream (as Noise)."""
    KIND = "SporadicRetriggerer"

    def __init__(self, target=None, rate=None, seed=None):
        super().__init__(target, rate)
        self.seed = config.host_random.next_noise_seed() if seed is None else int(seed) & ((1 << 64) - 1)
        self.draw = 0

    def blob_props(self):
        return {"seed_lo": self.seed & 0xFFFFFFFF, "seed_hi": self.seed >> 32, "draw": self.draw}, {}


# ------------------------------------------------------------ spectral units
# SURVEY 8f-3.  Frame-rate units: their ports carry chunks of `windowSize` / `2*windowSize` / `hopSize` samples
# and they tick every `tickInterval` samples (Circuit.js:34-37).  `_tick`s live in csrc/render_sm100.cu.
def _is_pow2(n):
    return isinstance(n, int) and n > 1 and (n & (n - 1)) == 0


def _need_int(what, v):
    if not isinstance(v, int) or isinstance(v, bool) or v <= 0:
        raise DuspError(what + " must be a positive integer number of samples")
    return v


class Hopper(Unit):
    """src/components/spectral/Hopper.js:4-38 -- collects the input into overlapping frames of `frameSize`
    samples, one every `hopSize`."""
    KIND = "Hopper"

    def __init__(self, hopSize=None, frameSize=None):
        super().__init__()
        import math
        _need_int("Hopper hopSize", hopSize)
        _need_int("Hopper frameSize", frameSize)
        self.addInlet("in")
        self.addOutlet("out", {"chunkSize": frameSize})
        self.hopSize = hopSize
        self.frameSize = frameSize
        self.t = 0
        self.tickInterval = math.gcd(hopSize, self.IN.chunkSize)

    def blob_props(self):
        return {"hopSize": self.hopSize, "frameSize": self.frameSize, "t": self.t, "tickInterval": self.tickInterval}, {}


class Windower(Unit):
    """src/components/spectral/Windower.js:3-49"""
    KIND = "Windower"

    def __init__(self, windowSize=None, kind="hamming", hopSize=None):
        super().__init__()
        if not windowSize:
            raise DuspError("Windower constructor expects a windowSize")
        _need_int("Windower windowSize", windowSize)
        self.addInlet("in", {"chunkSize": windowSize})
        self.addOutlet("out", {"chunkSize": windowSize})
        self.tickInterval = _need_int("Windower hopSize (its tick interval)", hopSize)
        self.windowSize = windowSize
        self.windowKind = kind
        if kind != "hamming":  # Windower.js:31-35,38-39
            raise DuspError("Window type '" + str(kind) + "' is not defined.")

    def blob_props(self):
        return {"windowSize": self.windowSize, "tickInterval": self.tickInterval}, {"windowKind": self.windowKind}


class FFT(Unit):
    """src/components/spectral/FFT.js:4-26 (fft.js realTransform + completeSpectrum)"""
    KIND = "FFT"

    def __init__(self, windowSize=None, hopSize=None):
        super().__init__()
        if not windowSize:
            raise DuspError("FFT expects window size")
        if not _is_pow2(windowSize):  # fft.js constructor
            raise DuspError("FFT size must be a power of two and bigger than 1")
        self.windowSize = windowSize
        self.frameSize = windowSize * 2
        self.tickInterval = _need_int("FFT hopSize (its tick interval)", hopSize)
        self.addInlet("in", {"chunkSize": windowSize})
        self.addOutlet("out", {"chunkSize": self.frameSize, "type": "spectral"})

    def blob_props(self):
        return {"windowSize": self.windowSize, "tickInterval": self.tickInterval}, {}


class IFFT(Unit):
    """src/components/spectral/IFFT.js:4-34 (fft.js inverseTransform, real part kept)"""
    KIND = "IFFT"

    def __init__(self, windowSize=None, hopSize=None):
        super().__init__()
        if not windowSize:
            raise DuspError("IFFT constructor requires argument: windowSize")
        if not _is_pow2(windowSize):
            raise DuspError("FFT size must be a power of two and bigger than 1")
        self.windowSize = windowSize
        self.frameSize = windowSize * 2
        self.tickInterval = _need_int("IFFT hopSize (its tick interval)", hopSize)
        self.addInlet("in", {"type": "spectral", "chunkSize": self.frameSize})
        self.addOutlet("out", {"chunkSize": windowSize})

    def blob_props(self):
        return {"windowSize": self.windowSize, "tickInterval": self.tickInterval}, {}


class UnHopper(Unit):
    """src/components/spectral/UnHopper.js:3-45 -- overlap-add of `windowSize` frames, one hop out per tick"""
    KIND = "UnHopper"

    def __init__(self, hopSize=None, windowSize=None):
        super().__init__()
        self.windowSize = _need_int("UnHopper windowSize", windowSize)
        self.hopSize = _need_int("UnHopper hopSize", hopSize)
        self.tickInterval = hopSize
        self.addInlet("in", {"chunkSize": windowSize})
        self.addOutlet("out", {"chunkSize": hopSize})
        self.t = 0

    def blob_props(self):
        return {"hopSize": self.hopSize, "windowSize": self.windowSize, "t": self.t, "tickInterval": self.tickInterval}, {}


class ReChunk(Unit):
    """src/components/spectral/ReChunk.js:5-51 -- re-blocks a signal between two chunk sizes"""
    KIND = "ReChunk"

    def __init__(self, inputInterval=None, outputInterval=None):
        super().__init__()
        import math
        if not inputInterval or not outputInterval:
            raise DuspError("ReChunk expects 2 numeric contructor arguments")
        self.inputInterval = _need_int("ReChunk inputInterval", inputInterval)
        self.outputInterval = _need_int("ReChunk outputInterval", outputInterval)
        self.addInlet("in", {"chunkSize": inputInterval})
        self.addOutlet("out", {"chunkSize": outputInterval})
        self.tickInterval = math.gcd(inputInterval, outputInterval)
        self.bufferSize = inputInterval * outputInterval // self.tickInterval  # compute-lcm
        self.t = 0

    def blob_props(self):
        return {"inputInterval": self.inputInterval, "outputInterval": self.outputInterval, "t": self.t,
                "bufferSize": self.bufferSize, "tickInterval": self.tickInterval}, {}


class SpectralUnit(Unit):
    """src/components/spectral/SpectralUnit.js:8-33"""
    isSpectralUnit = True

    def __init__(self):
        super().__init__()
        self.windowSize = config.fft["windowSize"]
        self.frameSize = self.windowSize * 2
        self.hopInterval = config.fft["hopSize"]
        self.tickInterval = self.hopInterval

    def addSpectralInlet(self, name, options=None):
        options = dict(options or {}, type="spectral", chunkSize=self.frameSize)
        return self.addInlet(name, options)

    def addSpectralOutlet(self, name, options=None):
        options = dict(options or {}, type="spectral", chunkSize=self.frameSize)
        return self.addOutlet(name, options)

    def _frame_props(self):
        return {"windowSize": self.windowSize, "tickInterval": self.tickInterval}

    def blob_props(self):
        return self._frame_props(), {}


class HardLowPass(SpectralUnit):
    """src/components/spectral/HardLowPass.js:7-32"""
    KIND = "HardLowPass"

    def __init__(self, f=None):
        super().__init__()
        self.addSpectralInlet("in")
        self.addInlet("f", {"mono": True, "type": "frequency"})
        self.addSpectralOutlet("out")
        self.fPerBin = self.sampleRate / self.windowSize
        self.F = f


class HardHighPass(HardLowPass):
    """src/components/spectral/HardHighPass.js:7-32"""
    KIND = "HardHighPass"


class SpectralGate(SpectralUnit):
    """src/components/spectral/SpectralGate.js:3-33"""
    KIND = "SpectralGate"

    def __init__(self, threshold=None):
        super().__init__()
        self.addSpectralInlet("in")
        self.addInlet("threshold", {"mono": True})
        self.addSpectralOutlet("out")
        self.invert = True
        self.THRESHOLD = js_or(threshold, 0.5)

    def blob_props(self):
        return dict(self._frame_props(), invert=int(bool(self.invert))), {}


class BinShift(SpectralUnit):
    """src/components/spectral/BinShift.js:3-27"""
    KIND = "BinShift"

    def __init__(self, shift=None):
        super().__init__()
        self.addSpectralInlet("in")
        self.addInlet("shift", {"mono": True})
        self.addSpectralOutlet("out")
        self.SHIFT = js_or(shift, 0)


class Augment(SpectralUnit):
    """src/components/spectral/Augment.js:3-28.  ``incrementMapping`` {ratio: gain}: every bin is added, scaled,
    into bin round(bin * ratio); keys iterate in JavaScript's own-property order (integer keys ascending first)."""
    KIND = "Augment"

    def __init__(self, incrementMapping=None, windowSize=None, hopInterval=None):
        super().__init__()
        self.addSpectralInlet("in")
        self.addSpectralOutlet("out")
        self.incrementMapping = {1: 1} if incrementMapping is None else dict(incrementMapping)

    def mapping_in_js_order(self):
        def is_index(k):
            s = _js_key(k)
            return s.isdigit() and (s == "0" or not s.startswith("0")) and int(s) < 2 ** 32 - 1
        keys = list(self.incrementMapping)
        ints = sorted((k for k in keys if is_index(k)), key=lambda k: int(_js_key(k)))
        rest = [k for k in keys if not is_index(k)]
        return [(float(_js_key(k)), float(self.incrementMapping[k])) for k in ints + rest]

    def blob_props(self):
        p = self._frame_props()
        m = self.mapping_in_js_order()
        if len(m) > 8:
            raise DuspError("Augment with more than 8 mapping entries has no device implementation")
        p["nmap"] = len(m)
        for i, (r, g) in enumerate(m):
            p["ratio%d" % i] = r
            p["gain%d" % i] = g
        return p, {}


def _js_key(k):
    """Property-key string of a number used as an object key."""
    if isinstance(k, str):
        return k
    from ..dusp import js_number_string
    return js_number_string(k)


class SpectralSum(SpectralUnit):
    """src/components/spectral/SpectralSum.js:3-25"""
    KIND = "SpectralSum"

    def __init__(self, a=None, b=None, windowSize=None, hopInterval=None):
        super().__init__()
        self.addSpectralInlet("a")
        self.addSpectralInlet("b")
        self.addSpectralOutlet("out")
        self.A = a
        self.B = b


__all__ = [
    "Hopper", "Windower", "FFT", "IFFT", "UnHopper", "ReChunk", "SpectralUnit", "HardLowPass", "HardHighPass",
    "SpectralGate", "BinShift", "Augment", "SpectralSum",
    "AHD", "Abs", "AllPass", "CircleBuffer", "CircleBufferNode", "CircleBufferReader", "CircleBufferWriter",
    "Clip", "CombFilter", "ConcatChannels", "CrossFader", "DecibelToScaler", "Delay", "Divide", "Filter",
    "FixedDelay", "FixedMultiply", "Gain", "HardClipAbove", "HardClipBelow", "MidiToFrequency", "Monitor", "MonoDelay",
    "Multiply", "Noise", "MultiChannelOsc", "Osc", "Pan", "PickChannel", "PolarityInvert", "Pow", "Ramp",
    "ReadBackDelay", "Repeater", "Rescale", "Retriggerer", "SampleRateRedux", "SecondsToSamples", "SemitoneToRatio", "Shape", "SignalCombiner",
    "SporadicRetriggerer", "Subtract", "Sum", "Timer", "VectorMagnitude",
]
