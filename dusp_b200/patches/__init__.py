"""Graph-builder patches: they only wire hot-path units
together, so once flattened they need no kernel of their own.  Each class
mirrors the reference patch of the same name (file:line in the docstring)."""
from .. import config, quick
from ..config import js_or
from ..graph import Patch, DuspError
from .. import components as C


class SimpleDelay(Patch):
    """src/patches/SimpleDelay.js:9-39 -- a CYCLIC graph (Sum -> Delay -> Multiply -> Sum)."""

    def __init__(self, input=None, delay=None, feedback=None, dryWet=None):
        super().__init__()
        self.inputRepeater = C.Repeater()
        self.feedbackInputSum = C.Sum()
        self.delayer = C.Delay()
        self.mixDryWet = C.CrossFader()
        self.feedbackScaler = C.Multiply()
        self.delayScaler = C.SecondsToSamples()
        self.addUnits(self.inputRepeater, self.feedbackInputSum, self.delayer, self.mixDryWet,
                      self.feedbackScaler, self.delayScaler)
        self.feedbackInputSum.A = self.inputRepeater.OUT
        self.feedbackInputSum.B = self.feedbackScaler.OUT
        self.feedbackScaler.A = self.delayer.OUT
        self.mixDryWet.B = self.delayer.OUT
        self.mixDryWet.A = self.inputRepeater.OUT
        self.delayer.IN = self.feedbackInputSum.OUT
        self.delayer.DELAY = self.delayScaler.OUT
        self.aliasInlet(self.inputRepeater.IN)
        self.aliasInlet(self.delayScaler.IN, "delay")
        self.aliasInlet(self.feedbackScaler.B, "feedback")
        self.aliasInlet(self.mixDryWet.DIAL, "dryWet")
        self.aliasOutlet(self.mixDryWet.OUT)
        self.IN = js_or(input, 0)
        self.DELAY = js_or(delay, 4410)
        self.FEEDBACK = js_or(feedback, 0)
        self.DRYWET = js_or(dryWet, 0.4)


class MultiTapDelay(Patch):
    """src/patches/MultiTapDelay.js:8-44"""

    def __init__(self, nChannels=None, maxDelay=None, input=None):
        super().__init__()
        if not nChannels or not maxDelay:
            raise DuspError("MultiTapDelay requires constructor args (nChannels, maxDelay[, input])")
        self.buffer = C.CircleBuffer(nChannels, maxDelay)
        self.writer = C.CircleBufferWriter(self.buffer)
        self.addUnits(self.buffer, self.writer)
        self.writer.preWipe = True
        self.aliasInlet(self.writer.IN)
        self.IN = js_or(input, 0)

    def addTap(self, delay):  # :28-34
        reader = C.CircleBufferReader(self.buffer, delay)
        reader.t = self.writer.t
        self.addUnits(reader)
        reader.chain(self.writer)
        return reader

    def addFeedback(self, delay, feedbackGain, feedbackDelay=None):  # :36-44
        reader = self.addTap(delay)
        writer = C.CircleBufferWriter(self.buffer, js_or(feedbackDelay, 0))
        writer.IN = quick.multiply(reader, feedbackGain)
        writer.t = self.writer.t
        writer.chain(self.writer)
        self.addUnits(writer)
        return reader


class DelayMixer(Patch):
    """src/patches/DelayMixer.js:7-35"""

    def __init__(self, nChannels=None, maxDelay=None):
        super().__init__()
        if not nChannels or not maxDelay:
            raise DuspError("DelayMixer requires constructor arguments: (nChannels, maxDelay)")
        self.buffer = C.CircleBuffer(nChannels, maxDelay)
        self.outReader = C.CircleBufferReader(self.buffer)
        self.addUnits(self.outReader)
        self.outReader.postWipe = True
        self.aliasOutlet(self.outReader.OUT)

    def addInput(self, input, delay=None, attenuation=None):  # :23-33
        writer = C.CircleBufferWriter(self.buffer, delay)
        writer.t = self.outReader.t
        self.outReader.chain(writer)
        self.addUnits(writer)
        if attenuation:
            writer.IN = quick.multiply(input, attenuation)
        else:
            writer.IN = input


class Worm(Patch):
    """src/patches/Worm.js:7-26"""

    def __init__(self, f=1):
        super().__init__()
        self.noise = C.Noise()
        self.filter = C.Filter(self.noise, f)
        self.addUnits(self.noise, self.filter)
        self.aliasInlet(self.filter.F)
        self.aliasOutlet(self.filter.OUT)
        self.F = f

    @staticmethod
    def random(fMax=5):
        return Worm(quick.multiply(fMax, config.host_random.random()))


class Mixer(Patch):
    """src/patches/Mixer.js:7-119 (left-leaning chain of Sums behind a Repeater)"""

    def __init__(self, *inputs):
        super().__init__()
        self.sums = []
        self.inputs = []
        self.addRepeater = C.Repeater(0)
        self.addUnits(self.addRepeater)
        self.aliasOutlet(self.addRepeater.OUT)
        for i in inputs:
            self.addInput(i)

    def addInput(self, outlet):  # :26-47
        if not getattr(outlet, "isOutlet", False) and getattr(outlet, "defaultOutlet", None) is not None:
            outlet = outlet.defaultOutlet
        if len(self.inputs) == 0:
            self.addRepeater.IN = outlet
            self.inputs.append(outlet)
        elif len(self.inputs) == 1:
            newSum = C.Sum(self.addRepeater.IN.outlet, outlet)
            self.addRepeater.IN = newSum
            self.inputs.append(outlet)
            self.sums.append(newSum)
        else:
            lastSum = self.sums[-1]
            lastInput = lastSum.B.outlet
            newSum = C.Sum(lastInput, outlet)
            lastSum.B = newSum
            self.inputs.append(outlet)
            self.sums.append(newSum)
        return self

    def addMultiplied(self, outlet, sf=None):  # :49-56
        if not sf:
            return self.addInput(outlet)
        return self.addInput(C.Multiply(outlet, sf))

    def addAttenuated(self, outlet, gain=None):  # :58-65
        if not gain:
            return self.addInput(outlet)
        g = C.Gain()
        g.IN = outlet
        g.GAIN = gain
        return self.addInput(g)

    def addInputs(self, *args):
        for a in args:
            if isinstance(a, (list, tuple)):
                for x in a:
                    self.addInput(x)
            else:
                self.addInput(a)
        return self

    @property
    def numberOfInputs(self):
        return len(self.inputs)


class LFO(Patch):
    """src/patches/LFO.js:6-57"""

    def __init__(self, frequency=None, amplitude=None, origin=None, waveform=None):
        super().__init__()
        osc1 = C.Osc()
        self.alias(osc1.F)
        self.osc = osc1
        mult1 = C.Multiply(osc1.OUT)
        self.alias(mult1.B, "a")
        location = C.Sum(mult1.OUT)
        self.alias(location.B, "o")
        self.alias(location.OUT)
        self.addUnits(osc1, mult1, location)
        self.F = js_or(frequency, 1)
        self.A = js_or(amplitude, 1 / 2)
        self.O = js_or(origin, 1 / 2)
        self.osc.waveform = js_or(waveform, "sine")


class Boop(Patch):
    """src/patches/Boop.js:6-29"""

    def __init__(self, f=None, duration=None):
        super().__init__()
        self.osc = C.Osc(f)
        self.envelope = C.Shape("decay", duration).trigger()
        self.mult = C.Multiply(self.osc, self.envelope)
        self.addUnits(self.osc, self.envelope, self.mult)
        self.aliasOutlet(self.mult.OUT)

    def trigger(self):
        self.envelope.trigger()

    def stop(self):
        self.envelope.stop()


class StereoOsc(Patch):
    """src/patches/StereoOsc.js:9-51"""

    def __init__(self, p=None, gain=None, pan=None):
        super().__init__()
        sum1 = C.Sum()
        self.alias(sum1.A, "p")
        self.alias(sum1.B, "pControl")
        mToF1 = C.MidiToFrequency(sum1)
        osc1 = C.Osc()
        osc1.F = mToF1.FREQUENCY
        self.osc = osc1
        gain1 = C.Gain()
        gain1.IN = osc1
        self.alias(gain1.GAIN)
        pan1 = C.Pan()
        pan1.IN = gain1.OUT
        self.alias(pan1.PAN)
        self.alias(pan1.OUT)
        self.addUnit(sum1)  # reference calls addUnit(sum1, mToF1, ...): only the first is registered
        self.GAIN = js_or(gain, 0)
        self.PAN = js_or(pan, 0)
        self.P = js_or(p, 60)
        self.PCONTROL = 0

    def trigger(self):
        self.osc.phase = 0


class MidiOsc(Patch):
    """src/patches/MidiOsc.js:5-20"""

    def __init__(self, p=None):
        super().__init__()
        self.mToF = C.MidiToFrequency()
        self.osc = C.Osc(self.mToF.FREQUENCY)
        self.addUnits(self.mToF, self.osc)
        self.aliasInlet(self.mToF.MIDI, "P")
        self.aliasOutlet(self.osc.OUT)
        self.P = js_or(p, 69)


class BandFilter(Patch):
    """src/patches/BandFilter.js:4-21"""

    def __init__(self, input=None, fLow=None, fHigh=None):
        super().__init__()
        self.lowPass = C.Filter(input, fHigh, "LP")
        self.highPass = C.Filter(self.lowPass.OUT, fLow, "HP")
        self.addUnits(self.lowPass, self.highPass)
        self.highPass.kind = "HP"
        self.aliasInlet(self.lowPass.IN)
        self.aliasInlet(self.lowPass.F, "fHigh")
        self.aliasInlet(self.highPass.F, "fLow")
        self.aliasOutlet(self.highPass.OUT)


class APStack(Patch):
    """src/patches/APStack.js:4-32 (random all-pass chain; draws from the injected host generator)"""

    def __init__(self, n=4, maxDelay=0.1, maxFeedback=0.5):
        super().__init__()
        stack = C.AllPass.manyRandom(n, maxDelay, maxFeedback)
        for i in range(1, len(stack)):
            stack[i].IN = stack[i - 1]
        self.addUnits(stack)
        self.aliasInlet(stack[0].IN, "in")
        self.aliasOutlet(stack[-1].OUT, "out")


class FMOsc(Patch):
    """src/patches/FMOsc.js:8-58"""
    isFMOsc = True

    def __init__(self, f=None):
        super().__init__()
        self.fRepeater = C.Repeater()
        self.osc = C.MultiChannelOsc(self.fRepeater)
        self.addUnits(self.fRepeater, self.osc)
        self.osc.randomPhaseFlip()
        self.aliasInlet(self.fRepeater.IN, "f")
        self.aliasOutlet(self.osc.OUT)
        self.F = js_or(f, 440)

    def addModulator(self, modulator, ammount=None):  # :29-43
        ammount = js_or(ammount, 1)
        multiply1 = C.Multiply(modulator, ammount)
        m2f = C.SemitoneToRatio(multiply1)
        multiply2 = C.Multiply(m2f, self.osc.F.outlet)
        self.addUnits(multiply1, multiply2, m2f)
        self.osc.F = multiply2

    def addModulatorOsc(self, f, ammount=None):
        self.addModulator(FMOsc(f), ammount)

    def clearModulation(self):
        self.osc.F = self.fRepeater

    def resetPhase(self):
        self.osc.resetPhase()


class StereoDetune(Patch):
    """src/patches/StereoDetune.js:5-29"""

    def __init__(self, input=None, ammount=None):
        super().__init__()
        ammount = js_or(ammount, None)
        if ammount is None:
            ammount = 0.1 * config.host_random.random()
        ratioL = quick.semitoneToRatio(ammount)
        ratioR = quick.divide(1, ratioL)
        ratios = quick.concat(ratioL, ratioR)
        self.mult = C.Multiply(input, ratios)
        self.addUnits(self.mult)
        self.alias(self.mult.A, "in")
        self.alias(self.mult.OUT)

    @staticmethod
    def random(input, maxAmmount=None):
        maxAmmount = js_or(maxAmmount, 0.1)
        return StereoDetune(input, quick.multiply(maxAmmount, config.host_random.random()))


class FrequencyGroup(Patch):
    """src/patches/FrequencyGroup.js:5-42"""

    def __init__(self, f=None):
        super().__init__()
        self.fundamentalRepeater = C.Repeater(js_or(f, 440), "Hz")
        self.addUnits(self.fundamentalRepeater)
        self.fOuts = [self.fundamentalRepeater.OUT]
        self.alias(self.fundamentalRepeater.IN, "f")

    def addHarmonic(self, ratio):  # :24-30
        harmonic = quick.mult(self.fOuts[0], ratio)
        self.fOuts.append(harmonic)
        return harmonic


class ManyOsc(Patch):
    """src/patches/ManyOsc.js:8-45"""
    isManyOsc = True

    def __init__(self, oscs):
        super().__init__()
        mix = C.Sum.many(oscs)
        self.addUnits(mix, oscs)
        self.alias(mix.OUT, "OUT")

    @staticmethod
    def ofFrequencies(fundamental, ratios):  # :23-32
        oscs = []
        for r in ratios:
            osc = C.Osc()
            osc.F = C.Multiply(fundamental, r)
            oscs.append(osc)
        return ManyOsc(oscs)

    @staticmethod
    def random(n=None, min=None, max=None):  # :34-45
        n = js_or(n, 3)
        min = js_or(min, 20)
        max = js_or(max, 1000)
        freqs = [min + config.host_random.random() * (max - min) for _ in range(n)]
        return ManyOsc.ofFrequencies(1, freqs)


__all__ = ["APStack", "BandFilter", "Boop", "DelayMixer", "FMOsc", "FrequencyGroup", "LFO", "ManyOsc",
           "MidiOsc", "Mixer", "MultiTapDelay", "SimpleDelay", "StereoDetune", "StereoOsc", "Worm"]
