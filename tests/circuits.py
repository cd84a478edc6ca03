"""Circuit factories shared by the CPU and GPU tests.  Each returns the thing a
user would hand to renderChannelData (a Unit, Patch or Outlet).  `exact` says
whether GPU output must equal the oracle bit for bit; circuits that evaluate
tan / pow on the device (libdevice vs glibc differ in the last ulp of double)
are held to the north-star tolerance instead (max-abs <= 1e-5 of full scale).
The flag is not a judgement call: tests/test_capi_cpu.py checks that a case is
`exact` if and only if none of its programs holds a device-transcendental op
(semi, db2s, midi2f, pow, gain, pan, filter with signal-rate inlets); constant
inlets are specialised to host-derived words and stay exact.
"""
import dusp_b200 as d
from dusp_b200 import components as C, patches as P, quick


def osc440():
    return d.unDusp("[Osc 440]")


def readme_fm_literal():  # SURVEY 3.1 trap: renders a 1 Hz sine
    return d.unDusp("[Osc f:[Ramp from:200 to:100 duration:2]]")


def w2_literal():
    return d.unDusp("[Multiply A:[Osc #o f:[Ramp from:200 to:100 duration:2]] B:#o]")


def w2_live(f0=313.7):
    osc = C.Osc(C.Ramp(88200, f0, f0 / 2).trigger())
    return C.Multiply(osc, osc)


def sum_pan():  # test/test_renderChannelData.js:4
    return d.unDusp("(O200 + O250) @ 0")


def envelope_fm():
    return d.unDusp("[Osc f:(D2 * 100 + 100)]")


def attack_env():
    return d.unDusp("A0.5 * O330.5")


def waveforms():
    return d.unDusp("Z110.25 + Sq220.5 * 0.25 + [Osc f:77.7 waveform:'triangle'] + [Osc f:55.5 waveform:'8bit']")


def osc_negative_f():
    return C.Osc(quick.add(quick.multiply(C.Osc(3.3), 900), -200))


def arithmetic():
    a, b = C.Osc(123.4), C.Osc(0.77, "saw")
    x = C.Subtract(C.Multiply(a, b), C.Divide(a, quick.add(b, 1.5)))
    y = C.CrossFader(x, C.Abs(a), quick.multiply(quick.add(b, 1), 0.5))
    z = C.HardClipAbove(C.HardClipBelow(y, -0.4), 0.6)
    clip = C.Clip(0.5)
    clip.IN = C.PolarityInvert(z)
    r = C.Rescale(-1, 1, 3, 7)
    r.IN = clip
    return C.FixedMultiply(0.37, C.Sum(r, C.Repeater(a)))


def stereo_concat():
    a, b = C.Osc(300), C.Osc(301)
    st = quick.concat(a, b)                 # 2 channels
    m = C.Multiply(st, [0.5, 0.25])         # per-channel constants
    return C.Subtract(m, C.Osc(2))          # zero-fill broadcast: channel 1 gets 0 - ... no: a[1]-b[1] with b absent


def vector_magnitude():
    v = C.VectorMagnitude()
    v.IN = quick.concat(C.Osc(100), C.Osc(150))
    return v


def transcendental():  # pow on the device: tolerance class
    g = C.Gain(-6.5)
    g.IN = C.Osc(220)
    s = C.SemitoneToRatio(C.Multiply(C.Osc(0.5), 7))
    m = C.MidiToFrequency(quick.add(C.Multiply(C.Osc(0.25), 12), 60))
    db = C.DecibelToScaler(C.Multiply(C.Osc(1.5), 20))
    return C.Sum(C.Multiply(g, s), C.Multiply(C.Osc(m), C.Pow(C.Abs(db), 0.5)))


def pan_modulated():
    return C.Pan(C.Osc(440), C.Multiply(C.Osc(0.7), 0.9))


def noise_default():
    return C.Noise()


def noise_sample_hold():
    return C.Noise(3000.5)


def filter_lp():
    return C.Filter(C.Noise(), 800)


def filter_hp():
    return d.unDusp("[Noise] -> HP2500")


def filter_lp_low():
    return C.Filter(C.Noise(), 1)  # the Worm patch default: brutal on rounding fidelity


def filter_modulated():
    return C.Filter(C.Osc(110, "saw"), quick.add(quick.multiply(C.Osc(0.8), 900), 1000))


def filter_unset_f():  # NaN coefficients -> rendered as zeros
    return C.Filter(C.Osc(100))


def filter_stereo():
    return C.Filter(quick.concat(C.Noise(), C.Osc(50, "square")), 1200)


def shape_variants():
    a = C.Shape("semiSine", 0.3, -1, 2).trigger()
    b = C.Shape("decaySquared", 0.05).trigger()
    c = C.Shape("attack", 0.2)  # never triggered: sits on its left edge
    return C.Sum(C.Sum(a, b), c)


def ahd_timer():
    e = C.AHD(0.01, 0.02, 0.05).trigger()
    return C.Sum(C.Multiply(e, C.Osc(500)), C.Multiply(C.Timer(), 0.1))


def fixed_delays():
    src = C.Multiply(C.Osc(220, "saw"), C.Shape("decay", 0.05).trigger())
    fd = C.FixedDelay(0.01)
    fd.IN = src
    cf = C.CombFilter(0.0123, 0.6)
    cf.IN = fd
    ap = C.AllPass(0.00377, 0.5)
    ap.IN = cf
    ap2 = C.AllPass(0.0001, 0.7)  # 4-sample ring: recurrence inside a sub-block
    ap2.IN = ap
    return ap2


def short_rings():
    """Rings shorter than a 128- / 256-sample tile but not tiny (100, 40 and 106 samples): the tile runs in rounds of n
    independent samples (ops.cuh c_fixdelay), W4's shortest AllPass / FixedDelay times."""
    src = C.Multiply(C.Noise(), C.Shape("decay", 0.2).trigger())
    fd = C.FixedDelay(100 / 44100)
    fd.IN = src
    cf = C.CombFilter(40 / 44100, 0.55)
    cf.IN = fd
    ap = C.AllPass(0.0024, 0.151)
    ap.IN = cf
    return ap


def delay_interp():
    src = C.Multiply(C.Noise(), C.Shape("decay", 0.02).trigger())
    return C.Delay(src, quick.add(quick.multiply(C.Osc(3), 40.5), 300.25))


def delay_short_and_mono():
    src = C.Osc(333)
    return C.Sum(C.Delay(src, 3.5), C.MonoDelay(src, 100.75))


def readback_delay():
    return C.ReadBackDelay(C.Osc(410), 64, 4096)


def simple_delay():  # cyclic graph: one edge is read a chunk stale
    return P.SimpleDelay(C.Filter(C.Noise(), 800), 0.25, 0.5, 0.4)


def simple_delay_dusp():
    return d.unDusp("[SimpleDelay IN:[Filter IN:[Noise] F:800] DELAY:0.25 FEEDBACK:0.5] @ 0.3")


def w3_voice(f=1234.5, dly=0.137, g=0.55, p=-0.3):
    filt = C.Filter(C.Noise(), f, "LP")
    mt = P.MultiTapDelay(1, 1.0, filt)
    tap = mt.addFeedback(dly, g, 0)
    return C.Pan(C.Sum(filt, tap), p)


def delay_mixer():
    dm = P.DelayMixer(2, 0.5)
    dm.addInput(quick.concat(C.Osc(200), C.Osc(300)), 0.1, 0.5)
    dm.addInput(C.Osc(5, "square"), 0.25)
    return dm


def mixer_lfo():
    m = P.Mixer(C.Osc(100), P.LFO(2, 0.3, 0.1), C.Multiply(C.Osc(150), 0.5))
    m.addAttenuated(C.Osc(75), 0)
    return m


def worm():
    return P.Worm(3)


def boop():
    return P.Boop(660, 0.1)


def many_osc():
    return P.ManyOsc.ofFrequencies(110, [1, 2.01, 3.5])


def stereo_detune():
    return C.Osc(0) if False else P.StereoDetune(C.Osc(440), 0.3)


def fm_osc():
    f = P.FMOsc(220)
    f.addModulator(C.Osc(3), 2)
    return f


def ap_stack():
    s = P.APStack(3, 0.02, 0.5)
    s.IN = C.Multiply(C.Noise(), C.Shape("decay", 0.01).trigger())
    return s


def band_filter():
    return P.BandFilter(C.Noise(), 300, 3000)


def stereo_osc():
    return P.StereoOsc(57, -3, 0.4)


def midi_osc():
    return P.MidiOsc(64)


def at_zero():
    return d.unDusp("[Osc f:([Ramp 44100 400 200] at 0)]")


def at_later():  # a trigger event after the first chunk (constructOperation.js:63-69, UnitOrPatch.js:25-33)
    return d.unDusp("[Osc f:([Ramp 44100 400 200] at 0.5)]")


def scheduled_retriggers():  # Shape envelope re-triggered by events; a Ramp whose targets an event rewrites
    env = C.Shape("decay", 0.25).trigger()
    env.scheduleTrigger([0.3, 0.61, 0.9])
    ramp = C.Ramp(22050, 300, 600).trigger()

    def retarget(unit):
        unit.y0 = 600
        unit.y1 = 150
        unit.trigger()

    ramp.schedule(0.5, retarget)
    return C.Multiply(C.Osc(ramp), env)


def periodic_event():  # an event function returning a positive number re-schedules itself (Event.js:18-30)
    env = C.Shape("decay", 0.1).trigger()

    def again(unit):
        unit.trigger()
        return 0.2

    env.schedule(0.2, again)
    return C.Multiply(C.Osc(330), env)


def bang_retrigger():  # `!` = Retriggerer (constructOperation.js:71-76, Retriggerer.js:13-22), 5 per second
    return d.unDusp("O330 * (D0.15 ! 5)")


def retrigger_modulated():  # a signal-rate RATE inlet: the trigger instants follow an LFO; an AHD and a Ramp as targets
    env = C.AHD(0.01, 0.02, 0.05)
    C.Retriggerer(env, C.Sum(C.Multiply(C.Osc(0.7), 6), 9))
    ramp = C.Ramp(4000, 500, 250)
    C.Retriggerer(ramp, 3.5)
    return C.Multiply(C.Osc(ramp), env)


def sporadic_retrigger():  # `~!` = SporadicRetriggerer (constructOperation.js:78-82): one injected-random draw per chunk
    return d.unDusp("O220 * (D0.08 ~! 12)")


def retrigger_chain():  # the default Unit.trigger walks the target's input units (Unit.js:293-297): both envelopes restart
    both = C.Multiply(C.Shape("decay", 0.2), C.Shape("attack", 0.1))
    both.trigger()
    C.Retriggerer(both, 4)
    return C.Multiply(C.Osc(440), both)


def for_then():  # `for` = scheduleFinish, `then` = onFinish hands the Repeater a new source (graph edit)
    return d.unDusp("(O440 for 0.5) then O220")


def for_then_chain():  # finish times are absolute: the second switch comes 0.2 s after the first
    return d.unDusp("((O440 for 0.2) then ((Z330 for 0.4) then O220)) * 0.5")


def then_number():  # the new source is a Number: Inlet.setConstant (Inlet.js:76-93)
    return d.unDusp("[Osc f:((100 for 0.25) then 300)]")


def then_into_filter():  # units downstream of the edit keep their state (the filter rings on across the switch)
    return d.unDusp("((Noise for 0.3) then O100) -> LP500")


def then_with_delay():  # a feedback delay patch downstream: CircleBuffer contents and the stale edge survive the edit
    src = d.unDusp("(O660 * D0.1 for 0.35) then (O220 * 0.5)")
    return P.SimpleDelay(src, 0.11, 0.5, 0.4)


def shape_then_osc():  # `a then b` where a ends BY ITSELF: Shape.finish() from inside its tick (Shape/index.js:40-42) hands the
    return d.unDusp("D0.3 then O440")  # Repeater its new source in the middle of Circuit.tick; units after the Shape see it in that tick


def attack_then_decay():  # two self-finishing envelopes in a row (the second one's own finish only raises its flag)
    return d.unDusp("(A0.1 then D0.4) * O220")


def shape_then_number_into_delay():  # the edit sets a constant; a feedback delay downstream keeps ringing across it
    src = d.unDusp("(D0.2 then 0.25) * Z330")
    return P.SimpleDelay(src, 0.07, 0.4, 0.5)


def tight_taps():  # a11: MultiTapDelay taps closer than one chunk to the writer (and to each other): whole-chunk tiles
    src = C.Multiply(C.Osc(220, "saw"), C.Shape("decay", 0.2).trigger())
    mt = P.MultiTapDelay(1, 0.5, src)
    a = mt.addTap(0.003)
    b = mt.addFeedback(0.0041, 0.5, 0)
    return C.Sum(a, b)


def short_shared_ring():  # a11: a 353-sample CircleBuffer shared by a writer and a reader (shorter than two chunks)
    mt = P.MultiTapDelay(1, 0.008, C.Osc(330))
    return mt.addTap(0.004)


def modulated_tap():  # a11: signal-rate OFFSET on a shared ring (CircleBufferReader.js:12-25, sample after sample)
    mt = P.MultiTapDelay(1, 0.5, C.Osc(440))
    return mt.addTap(quick.add(quick.multiply(C.Osc(2), 0.002), 0.01))


def noise_two_channels():  # Noise.js:17-27: one phase / held value / stream, channel after channel
    n = C.Noise(seed=99)
    n.OUT.chunk = [0.0, 0.0]
    return n


def noise_two_channels_held():  # ... with F below the sample rate: the hold carries over from channel 0 to channel 1
    n = C.Noise(2500.5, seed=7)
    n.OUT.chunk = [0.0, 0.0]
    return n


def pick_channel_signal():  # PickChannel.js:19 with a signal-rate index
    st = quick.concat(C.Osc(300), C.Osc(301.5, "saw"))
    return C.PickChannel(st, quick.add(quick.multiply(C.Osc(5, "square"), 0.5), 0.5))  # 1 / 0


def event_assigns_inlet():  # an event that assigns an inlet constant and later connects a unit (Unit.js:59-67)
    osc = C.Osc(200)
    out = C.Multiply(osc, 0.5)

    def retune(unit):
        unit.F = 310.5

    def modulate(unit):
        unit.F = C.Sum(C.Multiply(C.Osc(5), 20), 400)

    osc.schedule(0.2, retune)
    osc.schedule(0.5, modulate)
    return out


def pick_channel():  # PickChannel.js:16-21 with constant indices: `c % in.length` selects the channel
    st = quick.concat(C.Osc(300), C.Osc(301.5, "saw"))
    return C.Sum(C.PickChannel(st, 1), C.Multiply(C.PickChannel(st, 2), 0.5))  # 2 % 2 = channel 0


def delay_ring_wrap():  # rendered past the 5 s wrap of the Delay / MonoDelay rings: (clock + t) % 220500, Delay.js:27, MonoDelay.js:21
    src = C.Multiply(C.Osc(0.7, "square"), C.Osc(441))
    return C.Sum(C.Delay(src, quick.add(quick.multiply(C.Osc(0.31), 2000.25), 9000.5)), C.MonoDelay(src, 30000.75))


def osc_tiny_f():  # increments below 2^-13 Hz: `phase += f` rounds on every sample (Osc.js:40), no exact scan exists
    return C.Sum(C.Osc(1e-6, "saw"), C.Multiply(C.Osc(3.1e-5), 1000))


def osc_fm_wide():  # instantaneous f jumps between +-20000 and +-1e-5 inside a chunk (a 2^31 span) and changes sign
    gate = quick.multiply(quick.add(C.Osc(1000, "square"), 1), 10000)  # 20000 / 0
    f = C.Multiply(quick.add(gate, 1e-5), C.Osc(3, "square"))
    return C.Osc(f)


def osc_fm_through_zero():  # deep audio-rate FM through zero: sub-2^-13 increments occur only near the crossings
    return C.Osc(quick.add(quick.multiply(C.Osc(110.5), 9000), 12.5))


def osc_phase_offset():  # a start phase with bits far below 2^-36 (not reachable from phase 0 by audio-range increments)
    o = C.Osc(440.25)
    o.phase = 0.3
    return o



# name -> (factory, seconds, exact?)
def sample_rate_redux():  # SampleRateRedux.js: sample-and-hold with a modulated period, stereo input
    src = C.ConcatChannels(C.Osc(441), C.Osc(100.5, "saw"))
    return C.SampleRateRedux(src, quick.add(quick.multiply(C.Osc(0.9), 6), 8))


def sample_rate_redux_static():
    return d.unDusp("[SampleRateRedux in:O330 ammount:4.5]")


def fm_synth():  # note-triggered FM patch built from a seeded random seed (FMSynth.js:105-143)
    d.config.seed(5)
    fm = P.FMSynth(P.FMSynth.randomSeed(f=110, nOscs=4, pConnection=0.4))
    fm.trigger(57)
    return fm


def fm_synth_worm():  # Worm LFO envelopes + stereo detune: channel counts settle over the first ticks
    d.config.seed(6)
    fm = P.FMSynth(P.FMSynth.wormSeed(f=110, nOscs=3, pConnection=0.5, pMix=0))
    fm.oscs[0].OUT  # (all oscillators silent in the mix when pMix = 0: render one of them instead)
    fm.trigger(50)
    return fm.oscs[1]


def fm_synth_retriggered():  # the synth's trigger() as scheduled events: F assignment + phase resets + envelopes
    d.config.seed(11)
    fm = P.FMSynth(P.FMSynth.randomSeed(f=80, nOscs=3, pConnection=0.3, pMix=1))
    fm.trigger(45)
    fm.schedule(0.25, lambda s: s.trigger(52))
    fm.schedule(0.5, lambda s: s.trigger(40))
    return fm


def old_fm_synth():
    d.config.seed(2)
    s = P.OldFMSynth(220)
    s.addOsc(1, 0.5)
    s.addOsc(2, 0.25, [3])
    s.addOsc(3.01, None, None, [1.5, 0.5])
    return s


def sine_boop():
    b = P.SineBoop(72, 0.3)
    b.trigger()
    return b


def space_boop():
    b = P.SpaceBoop(64, "triangle", 0.4, "decaySquared", [0.5, 1.0])
    b.trigger()
    return b


def space_moving():  # a source circling between two speakers: distance -> Gain (pow) and MonoDelay
    place = C.ConcatChannels(C.Multiply(C.Osc(0.7), 3), C.Multiply(C.Osc(0.45, "triangle"), 2))
    return P.Space(C.Osc(330), place)


def scary_patch():  # the signal is its own position; 3 warm-up ticks before every unit sees both channels
    return P.ScaryPatch(C.ConcatChannels(C.Osc(220), C.Osc(1.3)), 2.5)


def trigger_group():
    g = P.TriggerGroup()
    a, b = P.SineBoop(60, 0.2), P.SineBoop(67, 0.4)
    g.addTrigger(a, "a")
    g.addTrigger(b)
    g.trigger("a")
    g.trigger(0)
    g.trigger("zzz")
    return g


def ap_web():
    d.config.seed(9)
    w = P.APWeb(4, 0.01, 0.3)
    w.IN = C.Noise(seed=3)
    return w


def dusp_roundtrip():  # dusp.js text of a circuit, parsed again, renders like the original (see test_dusp.py)
    from dusp_b200.dusp import dusp
    return d.unDusp(dusp(d.unDusp("[Filter in:(Z110.25 * D0.5) f:(O2 * 300 + 900)] @ -0.5")))


# ---- frame-rate units (components/spectral)
def _stft(src, W, H, mid=None, post_window=False):
    """src -> Hopper -> Windower -> FFT -> [mid] -> IFFT -> (Windower) -> UnHopper"""
    h = C.Hopper(H, W); h.IN = src
    w = C.Windower(W, "hamming", H); w.IN = h
    f = C.FFT(W, H); f.IN = w
    x = mid(f) if mid else f
    i = C.IFFT(W, H); i.IN = x
    y = i
    if post_window:
        y = C.Windower(W, "hamming", H); y.IN = i
    u = C.UnHopper(H, W); u.IN = y
    return u


def spectral_identity_small():  # hop = the standard chunk: the UnHopper's outlet feeds standard units directly
    u = _stft(C.Sum(C.Osc(441), C.Multiply(C.Osc(1234.5, "saw"), 0.3)), 1024, 256)
    return C.Multiply(u, 0.5)


def spectral_identity():  # the reference's default sizes (config.fft: window 4096, hop 1024), back to 256 through a ReChunk
    u = _stft(C.Sum(C.Osc(441), C.Noise(seed=5)), 4096, 1024, post_window=True)
    r = C.ReChunk(1024, 256); r.IN = u
    return C.Multiply(r, 0.25)


def spectral_unhopper_root():  # renderChannelData straight off a 1024-sample outlet (outlet.chunkSize = 1024)
    return _stft(C.Osc(330.3, "triangle"), 4096, 1024)


def spectral_lowpass():
    def mid(f):
        lp = C.HardLowPass(1000); lp.IN = f
        return lp
    u = _stft(C.Sum(C.Osc(441), C.Osc(3000.7)), 4096, 1024, mid)
    r = C.ReChunk(1024, 256); r.IN = u
    return r


def spectral_bandpass_modulated():  # HardBandPass patch, LOW from a signal: read once per 1024-sample tick
    def mid(f):
        bp = P.HardBandPass(f, 100, 5000)
        bp.LOW = quick.add(quick.multiply(C.Osc(1.5), 400), 600)
        return bp
    u = _stft(C.Noise(seed=8), 4096, 1024, mid)
    r = C.ReChunk(1024, 256); r.IN = u
    return r


def spectral_gate_shift_sum():
    def mid(f):
        gate = C.SpectralGate(40.0); gate.IN = f; gate.invert = False
        shift = C.BinShift(7); shift.IN = gate
        quiet = C.SpectralGate(3.0); quiet.IN = f
        hp = C.HardHighPass(200); hp.IN = quiet
        return C.SpectralSum(shift, hp)
    u = _stft(C.Sum(C.Osc(220.5), C.Multiply(C.Noise(seed=21), 0.05)), 4096, 1024, mid)
    r = C.ReChunk(1024, 256); r.IN = u
    return r


def spectral_augment():
    def mid(f):
        a = C.Augment({1: 0.5, 2: 0.25, 0.5: 0.125}); a.IN = f
        down = C.BinShift(-3.4); down.IN = a
        return down
    u = _stft(C.Osc(500.25, "square"), 4096, 1024, mid)
    r = C.ReChunk(1024, 256); r.IN = u
    return C.Abs(r)


def rechunk_up_down():  # a standard signal re-blocked to 512-sample chunks and back
    up = C.ReChunk(256, 512); up.IN = C.Osc(777.7)
    down = C.ReChunk(512, 256); down.IN = up
    return down


CASES = {
    "osc440": (osc440, 1.0, True),
    "readme_fm_literal": (readme_fm_literal, 1.0, True),
    "w2_literal": (w2_literal, 0.5, True),
    "w2_live": (w2_live, 2.5, True),
    "sum_pan": (sum_pan, 1.0, True),
    "envelope_fm": (envelope_fm, 2.2, True),
    "attack_env": (attack_env, 0.7, True),
    "waveforms": (waveforms, 0.5, True),
    "osc_negative_f": (osc_negative_f, 0.5, True),
    "arithmetic": (arithmetic, 0.3, True),
    "stereo_concat": (stereo_concat, 0.2, True),
    "vector_magnitude": (vector_magnitude, 0.2, True),
    "transcendental": (transcendental, 0.5, False),
    "pan_modulated": (pan_modulated, 0.5, False),
    "noise_default": (noise_default, 0.3, True),
    "noise_sample_hold": (noise_sample_hold, 0.5, True),
    "filter_lp": (filter_lp, 0.5, True),
    "filter_hp": (filter_hp, 0.5, True),
    "filter_lp_low": (filter_lp_low, 1.0, True),
    "filter_modulated": (filter_modulated, 0.5, False),
    "filter_unset_f": (filter_unset_f, 0.1, True),
    "filter_stereo": (filter_stereo, 0.3, True),
    "shape_variants": (shape_variants, 0.5, True),
    "ahd_timer": (ahd_timer, 0.2, True),
    "fixed_delays": (fixed_delays, 0.3, True),
    "short_rings": (short_rings, 0.3, True),
    "delay_interp": (delay_interp, 0.3, True),
    "delay_short_and_mono": (delay_short_and_mono, 0.2, True),
    "readback_delay": (readback_delay, 0.2, True),
    "simple_delay": (simple_delay, 1.0, True),
    "simple_delay_dusp": (simple_delay_dusp, 1.0, True),
    "w3_voice": (w3_voice, 1.5, True),  # constant F: the biquad coefficients come from the host libm on both sides
    "delay_mixer": (delay_mixer, 0.8, True),
    "mixer_lfo": (mixer_lfo, 0.5, True),
    "worm": (worm, 0.5, True),
    "boop": (boop, 0.2, True),
    "many_osc": (many_osc, 0.3, True),
    "stereo_detune": (stereo_detune, 0.2, True),
    "fm_osc": (fm_osc, 0.3, False),
    "ap_stack": (ap_stack, 0.3, True),
    "band_filter": (band_filter, 0.3, True),
    "stereo_osc": (stereo_osc, 0.3, False),
    "midi_osc": (midi_osc, 0.2, False),
    "at_zero": (at_zero, 1.2, True),
    "at_later": (at_later, 1.2, True),
    "scheduled_retriggers": (scheduled_retriggers, 1.1, True),
    "periodic_event": (periodic_event, 0.9, True),
    "bang_retrigger": (bang_retrigger, 1.3, True),
    "retrigger_modulated": (retrigger_modulated, 1.5, True),
    "sporadic_retrigger": (sporadic_retrigger, 1.5, True),
    "retrigger_chain": (retrigger_chain, 1.1, True),
    "for_then": (for_then, 1.0, True),
    "for_then_chain": (for_then_chain, 0.8, True),
    "then_number": (then_number, 0.6, True),
    "then_into_filter": (then_into_filter, 0.7, True),
    "then_with_delay": (then_with_delay, 0.9, True),
    "event_assigns_inlet": (event_assigns_inlet, 0.8, True),
    "tight_taps": (tight_taps, 0.5, True),
    "short_shared_ring": (short_shared_ring, 0.3, True),
    "modulated_tap": (modulated_tap, 0.6, True),
    "noise_two_channels": (noise_two_channels, 0.2, True),
    "noise_two_channels_held": (noise_two_channels_held, 0.3, True),
    "pick_channel_signal": (pick_channel_signal, 0.5, True),
    "shape_then_osc": (shape_then_osc, 0.6, True),
    "attack_then_decay": (attack_then_decay, 0.7, True),
    "shape_then_number_into_delay": (shape_then_number_into_delay, 0.6, True),
    "sample_rate_redux": (sample_rate_redux, 0.4, True),
    "sample_rate_redux_static": (sample_rate_redux_static, 0.2, True),
    "fm_synth": (fm_synth, 0.6, False),
    "fm_synth_worm": (fm_synth_worm, 0.5, False),
    "fm_synth_retriggered": (fm_synth_retriggered, 0.8, False),
    "old_fm_synth": (old_fm_synth, 0.4, False),
    "sine_boop": (sine_boop, 0.4, False),
    "space_boop": (space_boop, 0.5, False),
    "space_moving": (space_moving, 0.6, False),
    "scary_patch": (scary_patch, 0.4, False),
    "trigger_group": (trigger_group, 0.5, False),
    "ap_web": (ap_web, 0.4, True),
    "dusp_roundtrip": (dusp_roundtrip, 0.6, False),
    "spectral_identity_small": (spectral_identity_small, 0.3, True),
    "spectral_identity": (spectral_identity, 0.5, True),
    "spectral_unhopper_root": (spectral_unhopper_root, 0.4, True),
    "spectral_lowpass": (spectral_lowpass, 0.5, True),
    "spectral_bandpass_modulated": (spectral_bandpass_modulated, 0.5, True),
    "spectral_gate_shift_sum": (spectral_gate_shift_sum, 0.5, True),
    "spectral_augment": (spectral_augment, 0.4, True),
    "rechunk_up_down": (rechunk_up_down, 0.2, True),
    "pick_channel": (pick_channel, 0.2, True),
    "delay_ring_wrap": (delay_ring_wrap, 5.3, True),
    "osc_tiny_f": (osc_tiny_f, 1.0, True),
    "osc_fm_wide": (osc_fm_wide, 1.0, True),
    "osc_fm_through_zero": (osc_fm_through_zero, 1.0, True),
    "osc_phase_offset": (osc_phase_offset, 0.5, True),
}
