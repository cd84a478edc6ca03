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


class HardBandPass(Patch):
    """src/patches/HardBandPass.js:12-34 (spectral low-pass into spectral high-pass; the constructor hands `low` to
    the low-pass and `high` to the high-pass before the aliased LOW / HIGH setters swap them into place)"""

    def __init__(self, input=None, low=None, high=None):
        super().__init__()
        self.lp = C.HardLowPass(low)
        self.hp = C.HardHighPass(high)
        self.addUnits(self.lp, self.hp)
        self.hp.IN = self.lp.OUT
        self.aliasInlet(self.lp.IN, "in")
        self.aliasInlet(self.hp.F, "low")
        self.aliasInlet(self.lp.F, "high")
        self.aliasOutlet(self.hp.OUT)
        self.IN = js_or(input, 0)
        self.LOW = js_or(low, 0)
        self.HIGH = js_or(high, 22000)


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


class Synth(Patch):
    """src/patches/Synth.js:3-28 (note-triggered patches: trigger() fans out to the registered envelopes)"""

    def __init__(self):
        super().__init__()
        self.triggerList = []

    def trigger(self, p=None, note=None):  # :10-19
        if hasattr(self, "_trigger"):
            self._trigger(p, note)
        for env in self.triggerList:
            env.trigger()
        return self

    def addEnvelope(self, env):  # :21-26
        if getattr(env, "isOutlet", False):
            env = env.unit
        self.triggerList.append(env)
        return env


class TriggerGroup(Patch):
    """src/patches/TriggerGroup.js:4-35"""

    def __init__(self):
        super().__init__()
        self.mixer = Mixer()
        self.addUnits(self.mixer)
        self.triggers = {}
        self.aliasOutlet(self.mixer.OUT)

    def addTrigger(self, trigger, name=None):  # :18-26
        if name is None:
            name = 0
            while name in self.triggers:
                name += 1
        self.triggers[name] = trigger
        self.mixer.addInput(trigger)

    def trigger(self, which=None):  # :28-35
        t = self.triggers.get(which)
        if t:
            t.trigger()
        elif getattr(self, "handleUnknownTrigger", None):
            self.handleUnknownTrigger(which)
        # else: the reference only logs "unknown trigger"


class FMSynth(Synth):
    """src/patches/FMSynth.js:13-190.  ``seed`` = {fundamental, mod, envelopes, oscs:[{h, stereoDetune,
    modulation, mix}], resetOscsOnTrigger}; every entry goes through unDusp, and ``self.seed`` is the seed
    written back as Dusp text by the serialiser (dusp.js)."""

    def __init__(self, seed):
        super().__init__()
        from ..construct import unDusp
        from ..dusp import dusp
        self.resetOscsOnTrigger = True  # `seed.resetOscsOnTrigger || true` (:17)
        fundamental = unDusp(seed.get("fundamental"))
        globalModulation = unDusp(js_or(seed.get("mod"), 1))
        envelopes = [unDusp(e) for e in (seed.get("envelopes") or [])]
        oscSeeds = [{
            "h": unDusp(o.get("h")),
            "stereoDetune": unDusp(js_or(o.get("stereoDetune"), 0)),
            "modulation": [unDusp(a) for a in (o.get("modulation") or [])],
            "mix": unDusp(js_or(o.get("mix"), 0)),
        } for o in seed["oscs"]]

        def truthy(v):  # JS truthiness of a number / unit
            return js_or(v, None) is not None

        duspIndex = {}
        out = {"fundamental": dusp(fundamental, duspIndex), "mod": dusp(globalModulation, duspIndex), "oscs": []}
        for o in oscSeeds:  # :34-50
            e = {"h": dusp(o["h"], duspIndex)}
            if truthy(o["stereoDetune"]):
                e["stereoDetune"] = dusp(o["stereoDetune"], duspIndex)
            if truthy(o["mix"]):
                e["mix"] = dusp(o["mix"], duspIndex)
            if o["modulation"]:
                e["modulation"] = [dusp(a, duspIndex) for a in o["modulation"]]
            out["oscs"].append(e)
        out["resetOscsOnTrigger"] = self.resetOscsOnTrigger
        if envelopes:
            out["envelopes"] = [dusp(e, duspIndex) for e in envelopes]
        self.seed = out

        for e in envelopes:
            self.addEnvelope(e)
        fGroup = FrequencyGroup(fundamental)
        for o in oscSeeds:
            fGroup.addHarmonic(o["h"])
        oscs = []
        for i, o in enumerate(oscSeeds):  # :65-73
            if truthy(o["stereoDetune"]):
                oscs.append(FMOsc(StereoDetune(fGroup.fOuts[i + 1], o["stereoDetune"])))
            else:
                oscs.append(FMOsc(fGroup.fOuts[i + 1]))
        for carrier, o in enumerate(oscSeeds):  # :76-84 (`modulator` is the POSITION in the modulation list)
            for modulator, ammount in enumerate(o["modulation"]):
                if truthy(ammount):
                    src = oscs[modulator] if modulator < len(oscs) else None
                    oscs[carrier].addModulator(src, quick.multiply(ammount, globalModulation))
        mixer = Mixer()
        for i, o in enumerate(oscSeeds):
            if truthy(o["mix"]):
                mixer.addInput(quick.multiply(oscs[i], o["mix"]))
        self.oscs = oscs
        self.addUnits(fGroup, oscs, mixer)
        self.aliasOutlet(mixer.OUT, "OUT")
        self.aliasInlet(fGroup.F, "F")

    def _trigger(self, p=None, note=None):  # :98-103
        self.F = quick.pToF(p)
        if self.resetOscsOnTrigger:
            for o in self.oscs:
                o.resetPhase()

    @staticmethod
    def randomSeed(f=50, duration=1, nOscs=8, pConnection=0.1, maxModulationAmmount=6, pMix=0.5,
                   maxStereoDetune=1 / 2):  # :105-143
        nOscs = js_or(nOscs, 4)
        r = config.host_random.random
        import math
        oscs, envelopes = [], []
        for _ in range(nOscs):
            osc = {"h": math.ceil(r() * 32), "modulation": []}
            if r() < pMix:
                envelope = C.Shape.randomDecay(duration, 0, 1)
                envelopes.append(envelope)
                osc["mix"] = quick.multiply(envelope, r())
            for _j in range(nOscs):
                if r() < pConnection:
                    envelope = C.Shape.randomInRange(duration, 0, 1)
                    envelopes.append(envelope)
                    osc["modulation"] += [envelope, quick.multiply(r(), maxModulationAmmount)]
            oscs.append(osc)
        return {"fundamental": f, "oscs": oscs, "envelopes": envelopes}

    @staticmethod
    def wormSeed(f=50, nOscs=8, pConnection=0.1, maxModulationAmmount=6, pMix=0.5, maxStereoDetune=1 / 2,
                 maxHarmonic=16, maxWormFrequency=5):  # :145-188
        nOscs = js_or(nOscs, 4)
        r = config.host_random.random
        import math
        oscs, envelopes = [], []
        for _ in range(nOscs):
            osc = {"h": math.ceil(quick.multiply(r(), maxHarmonic)), "modulation": [],
                   "stereoDetune": r() * maxStereoDetune}
            if r() < pMix:
                envelope = r()
                envelopes.append(envelope)
                osc["mix"] = quick.multiply(envelope, r())
            for _j in range(nOscs):
                if r() < pConnection:
                    envelope = Worm.random(maxWormFrequency)
                    envelopes.append(envelope)
                    osc["modulation"] += [envelope, quick.multiply(r(), maxModulationAmmount)]
            oscs.append(osc)
        return {"fundamental": f, "oscs": oscs, "envelopes": envelopes}


class OldFMSynth(Patch):
    """src/patches/OldFMSynth.js:8-66 (constructor name in the reference is also ``FMSynth``)"""

    def __init__(self, f=None):
        super().__init__()
        self.mixer = Mixer()
        self.fgroup = FrequencyGroup()
        self.addUnits(self.mixer, self.fgroup)
        self.oscs = []
        self.aliasInlet(self.fgroup.F, "f")
        self.aliasOutlet(self.mixer.OUT, "out")
        self.F = js_or(f, 440)
        self.stereo = True

    def addOsc(self, fRatio=None, master=None, inAmmounts=None, outAmmounts=None):  # :28-51
        fRatio = js_or(fRatio, 1)
        f = self.fgroup.addHarmonic(fRatio)
        if self.stereo:
            f = StereoDetune(f)
        osc = FMOsc(f)
        self.oscs.append(osc)
        self.addUnits(osc)
        if js_or(master, None) is not None:
            self.mixer.addMultiplied(osc, master)
        if inAmmounts:
            for i in range(min(len(inAmmounts), len(self.oscs))):
                if js_or(inAmmounts[i], None) is not None:
                    osc.addModulator(self.oscs[i], inAmmounts[i])
        if outAmmounts:
            for i in range(min(len(outAmmounts), len(self.oscs))):
                if js_or(outAmmounts[i], None) is not None:
                    self.oscs[i].addModulator(osc, outAmmounts[i])
        return osc

    def clearModulation(self):
        for o in self.oscs:
            o.clearModulation()

    def applyModulationMatrix(self, matrix):  # :58-66
        self.clearModulation()
        for i, row in enumerate(matrix):
            for j, a in enumerate(row):
                if js_or(a, None) is not None and i < len(self.oscs) and j < len(self.oscs):
                    self.oscs[j].addModulator(self.oscs[i], a)


class SineBoop(Patch):
    """src/patches/SineBoop.js:8-43"""

    def __init__(self, p=None, duration=None):
        super().__init__()
        self.osc = MidiOsc(p)
        self.ramp = C.Shape("decay", duration)
        self.multiply = C.Multiply(self.ramp, self.osc.OUT)
        self.addUnits(self.osc, self.ramp, self.multiply)
        self.alias(self.osc.P, "p")
        self.alias(self.ramp.DURATION)
        self.alias(self.multiply.OUT)
        self.P = js_or(p, 60)
        self.DURATION = js_or(duration, 1)

    @staticmethod
    def randomTwinkle(maxDuration=None):  # :32-37
        boop = SineBoop()
        boop.P = 100 + config.host_random.random() * 37
        boop.ramp.randomDecay(js_or(maxDuration, 1))
        return boop

    def trigger(self):  # :39-43 (`this.osc` is the MidiOsc PATCH: the phase assignment never reaches the Osc unit)
        self.ramp.trigger()
        self.osc.phase = 0
        return self


class SpaceChannel(Patch):
    """src/patches/SpaceChannel.js:9-47 (one speaker: distance -> attenuation in dB + propagation delay)"""

    def __init__(self, speakerPosition=None):
        super().__init__()
        self.speakerPositionSubtracter = C.Subtract()
        self.distanceCalculator = C.VectorMagnitude()
        self.attenuationScaler = C.Multiply()
        self.delayScaler = C.Multiply()
        self.delayer = C.MonoDelay()
        self.attenuator = C.Gain()
        self.addUnits(self.speakerPositionSubtracter, self.distanceCalculator, self.attenuationScaler,
                      self.delayScaler, self.delayer, self.attenuator)
        self.distanceCalculator.IN = self.speakerPositionSubtracter.OUT
        self.attenuationScaler.A = self.distanceCalculator.OUT
        self.delayScaler.A = self.distanceCalculator.OUT
        self.attenuator.GAIN = self.attenuationScaler.OUT
        self.delayer.DELAY = self.delayScaler.OUT
        self.delayer.IN = self.attenuator.OUT
        self.aliasInlet(self.attenuator.IN)
        self.aliasInlet(self.speakerPositionSubtracter.A, "placement")
        self.aliasInlet(self.speakerPositionSubtracter.B, "speakerPosition")
        self.aliasInlet(self.attenuationScaler.B, "decibelsPerMeter")
        self.aliasInlet(self.delayScaler.B, "sampleDelayPerMeter")
        self.aliasOutlet(self.delayer.OUT)
        self.IN = 0
        self.PLACEMENT = [0, 0]
        self.SPEAKERPOSITION = js_or(speakerPosition, None) or [0, 0]
        self.DECIBELSPERMETER = -3
        self.SAMPLEDELAYPERMETER = config.sampleRate / 343


class Space(Patch):
    """src/patches/Space.js:8-64 (a SpaceChannel per speaker, concatenated into one multichannel outlet)"""

    def __init__(self, input=None, place=None):
        super().__init__()
        import math
        self.signalIn = C.Repeater()
        self.placementIn = C.Repeater()
        self.outRepeater = C.Repeater()
        self.addUnits(self.signalIn, self.placementIn, self.outRepeater)
        self.spaceChannels = []
        self.alias(self.signalIn.IN)
        self.alias(self.placementIn.IN, "placement")
        self.alias(self.outRepeater.OUT)
        self.IN = js_or(input, 0)
        self.PLACEMENT = js_or(place, None) or [0, 0]
        if config.channelFormat == "stereo":
            self.addSpeaker([-1, 0])
            self.addSpeaker([1, 0])
        elif config.channelFormat == "surround":
            for pos in ([-1, 1], [1, 1], [0, math.sqrt(2)], [0, 0], [-1, -1], [1, -1]):
                self.addSpeaker(pos)

    @staticmethod
    def stereo(input=None, place=None):  # :46-51 (on top of the constructor's own speakers)
        space = Space(input, place)
        space.addSpeaker([-1, 0])
        space.addSpeaker([1, 0])
        return space

    def addSpeaker(self, speakerPosition):  # :53-64
        chan = SpaceChannel()
        chan.SPEAKERPOSITION = speakerPosition
        chan.PLACEMENT = self.placementIn.OUT
        chan.IN = self.signalIn
        if self.outRepeater.IN.connected:
            self.outRepeater.IN = C.ConcatChannels(self.outRepeater.IN.outlet, chan)
        else:
            self.outRepeater.IN = chan
        self.spaceChannels.append(chan)
        self.addUnit(chan)


class SpaceBoop(Patch):
    """src/patches/SpaceBoop.js:11-58"""

    def __init__(self, p=None, waveform=None, d=None, decayForm=None, place=None):
        super().__init__()
        self.mToF = C.MidiToFrequency()
        self.osc = C.Osc(self.mToF)
        self.durationToRate = C.Divide(1 / config.sampleRate)
        self.envelope = C.Shape("decay", self.durationToRate)
        self.envelopeAttenuator = C.Multiply(self.osc, self.envelope)
        self.space = Space(self.envelopeAttenuator.OUT)
        self.addUnits(self.mToF, self.osc, self.durationToRate, self.envelope, self.envelopeAttenuator, self.space)
        self.aliasInlet(self.mToF.MIDI, "p")
        self.aliasInlet(self.space.PLACEMENT, "placement")
        self.aliasInlet(self.durationToRate.B, "duration")
        self.aliasOutlet(self.space.OUT)
        self.P = js_or(p, 60)
        self.PLACEMENT = js_or(place, None) or [0, 0]
        self.DURATION = js_or(d, 1)
        self.waveform = js_or(waveform, "sin")
        self.decayForm = js_or(decayForm, "decay")

    def trigger(self, pitch=None, duration=None):  # :37-44
        if js_or(pitch, None) is not None:
            self.P = pitch
        if js_or(duration, None) is not None:
            self.DURATION = duration
        self.osc.phase = 0
        self.envelope.trigger()

    @property
    def waveform(self):
        return self.osc.waveform

    @waveform.setter
    def waveform(self, w):
        self.osc.waveform = w

    @property
    def decayForm(self):
        return self.envelope.shape

    @decayForm.setter
    def decayForm(self, shape):
        self.envelope.shape = shape


class ScaryPatch(Patch):
    """src/patches/ScaryPatch.js:6-27 (the signal is its own position in space)"""

    def __init__(self, input=None, ammount=None):
        super().__init__()
        self.inRepeater = C.Repeater()
        self.ammountScaler = C.Multiply(self.inRepeater, 1)
        self.space = Space(self.inRepeater, self.ammountScaler)
        self.addUnits(self.inRepeater, self.ammountScaler, self.space)
        self.alias(self.inRepeater.IN)
        self.aliasInlet(self.ammountScaler.B, "ammount")
        self.alias(self.space.OUT)
        self.IN = js_or(input, None) or [0, 0]
        self.AMMOUNT = js_or(ammount, 1)


class AttenuationMatrix(Patch):
    """src/patches/AttenuationMatrix.js:4-41 (random Gain-weighted routing between ``nodes``; draws from the
    injected host generator in the reference's order)"""

    def __init__(self, nodes, pConnection=0.5, pMix=0.5, maxAmmount=1, minAmmount=0, maxMixAmmount=1,
                 minMixAmmount=0, allowFeedback=True):
        super().__init__()
        r = config.host_random.random
        outMixer = Mixer()
        for i in range(len(nodes)):
            mixer = Mixer()
            for j in range(len(nodes)):
                if j < i and not allowFeedback:
                    continue
                if r() < pConnection:
                    ammount = r() * (maxAmmount - minAmmount) + minAmmount
                    mixer.addAttenuated(nodes[j].OUT, ammount)
            if mixer.numberOfInputs:
                self.addUnits(mixer)
                nodes[i].IN = mixer
            if r() < pMix:
                ammount = r() * (maxMixAmmount - minMixAmmount) + minAmmount  # sic: minAmmount (:32)
                outMixer.addAttenuated(nodes[i].OUT, ammount)
        self.aliasInlet(nodes[0].IN, "in")
        self.aliasOutlet(outMixer.OUT, "out")


class APWeb(Patch):
    """src/patches/APWeb.js:5-24"""

    def __init__(self, n=4, maxDelay=0.01, maxFeedback=0.1):
        super().__init__()
        nodes = C.AllPass.manyRandom(n, maxDelay, maxFeedback)
        matrix = AttenuationMatrix(nodes=nodes, allowFeedback=False, pMix=1)
        self.addUnits(matrix)
        self.aliasInlet(matrix.IN, "in")
        self.aliasOutlet(matrix.OUT, "out")


__all__ = ["APStack", "APWeb", "AttenuationMatrix", "BandFilter", "Boop", "DelayMixer", "FMOsc", "FMSynth",
           "FrequencyGroup", "HardBandPass", "LFO", "ManyOsc", "MidiOsc", "Mixer", "MultiTapDelay", "ScaryPatch", "SimpleDelay",
           "SineBoop", "Space", "SpaceBoop", "SpaceChannel", "StereoDetune", "StereoOsc", "Synth", "TriggerGroup",
           "Worm"]
