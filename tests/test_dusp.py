"""The serialiser: mirror of src/dusp.js.  Expected strings are derived by hand from the source
(dusp.js:4-120, Unit.js:144-149, the per-class `dusp` hooks), quirks included; round trips are checked by rendering
the original and the re-parsed circuit with the oracle."""
import numpy as np
import pytest

import dusp_b200 as d
from dusp_b200 import components as C, patches as P
from dusp_b200.dusp import dusp, js_number_string, numberOfOutgoingConnections
from dusp_b200.render import circuit_blob


@pytest.mark.parametrize("x,s", [
    (440, "440"), (100.0, "100"), (0.1, "0.1"), (-2.5, "-2.5"), (1 / 3, "0.3333333333333333"),
    (1e21, "1e+21"), (1e-7, "1e-7"), (1.5e-7, "1.5e-7"), (0.000001, "0.000001"), (123456789012345680000, "123456789012345680000"),
    (float("nan"), "NaN"), (float("inf"), "Infinity"), (-float("inf"), "-Infinity"), (0.0, "0"), (-0.0, "0"), (5e-324, "5e-324"),
    (44100 / 343, "128.57142857142858"),
])
def test_number_to_string_follows_ecmascript(x, s):
    assert js_number_string(x) == s


def test_outgoing_connection_count_is_a_string():
    # Unit.js:144-149 adds the `connections` ARRAYS to a number
    o = C.Osc(1)
    assert numberOfOutgoingConnections(o) == "0"
    C.Multiply(o, 2)
    assert numberOfOutgoingConnections(o) == "0[object Object]"


def test_passthrough_values():
    assert dusp(None) is None and dusp(0) == 0 and dusp(1.5) == 1.5
    assert dusp("sin") == '"sin"'            # duspString, dusp.js:118-120
    assert dusp(P.Boop()) is None            # patches are not handled (dusp.js:92-94)


def test_osc_shorthand_only_at_the_root():
    assert dusp(d.unDusp("[Osc 440]")) == "O440"                       # Osc.js:26-32
    s = d.unDusp("O200 + O250")
    a, b = s.A.outlet.unit, s.B.outlet.unit        # one outgoing connection each: written in full
    assert dusp(s) == "([Osc #%s F:200] + [Osc #%s F:250])" % (a.label, b.label)   # Sum.js:13-15


def test_labels_and_references():
    m = d.unDusp("[Multiply A:[Osc #o f:[Ramp from:200 to:100 duration:2]] B:#o]")
    ramp = m.A.outlet.unit.F.outlet.unit
    assert dusp(m) == "([Osc #o F:[Ramp #%s]] * #o)" % ramp.label


def test_extra_properties_args_and_constants():
    sh = C.Shape("attack", 0.5, 0, 2).trigger()
    assert dusp(sh) == '[Shape DURATION:0.5 MIN:0 MAX:2 shape:"attack" trigger]'      # Shape/index.js:61-77
    rep = C.Repeater(C.Osc(3, "saw"), "Hz")
    osc = rep.IN.outlet.unit
    assert dusp(rep) == '[Repeater IN:[Osc #%s F:3 waveform:"saw"] "Hz"]' % osc.label  # Osc.js:22-25, Repeater.js:15-22
    sub = C.Subtract(0, [1, 0.5])
    assert dusp(sub) == "[Subtract A:0 B:1,0.5]"    # plain concatenation of an array constant (dusp.js:56-59)
    assert dusp(sub.B) == "(1,0.5)"                 # duspInlet (dusp.js:108-116)
    pan = C.Pan(C.Osc(5), 0.25)
    assert dusp(pan) == "[Pan IN:[Osc #%s F:5] PAN:0.25]" % pan.IN.outlet.unit.label
    index = {}
    assert dusp(pan.OUT, index) == dusp(pan)       # the default outlet stands for its unit (dusp.js:100-103)
    assert dusp(pan, index) == "#" + pan.label     # second mention inside one index: a reference (dusp.js:19-20)
def test_non_default_outlet_gets_a_suffix():
    m = C.MidiToFrequency(60)
    o = C.Osc(m.FREQUENCY)
    # `frequency` is MidiToFrequency's only (default) outlet: no suffix (dusp.js:100-103)
    assert dusp(o) == "[Osc F:[MidiToFrequency #%s MIDI:60]]" % m.label


ROUNDTRIP = [
    "[Osc 440]",
    "(O200 + O250) @ 0",
    "[Osc f:(D2 * 100 + 100)]",
    "[Filter in:(Z110.25 * D0.5) f:(O2 * 300 + 900)] @ -0.5",
    "[Osc f:77.7 waveform:'triangle'] * A0.1",
    "[SimpleDelay IN:[Filter IN:[Noise] F:800] DELAY:0.25 FEEDBACK:0.5] @ 0.3",
]


@pytest.mark.parametrize("src", ROUNDTRIP)
def test_roundtrip_renders_the_same(oracle, src):
    d.config.seed(77)
    a = d.unDusp(src)
    text = dusp(a)
    assert isinstance(text, str)
    d.config.seed(77)
    b = d.unDusp(text)
    # a patch serialises to null in the reference; only unit-rooted expressions round-trip
    ra = oracle.render(circuit_blob(a)[0], 0.3)
    rb = oracle.render(circuit_blob(b)[0], 0.3)
    assert len(ra) == len(rb)
    for x, y in zip(ra, rb):
        assert np.array_equal(x, y, equal_nan=True)
    d.config.seed(77)
    assert dusp(d.unDusp(text)) is not None


def test_shared_reference_under_a_root_operator_does_not_reparse():
    """The reference's own limitation (FMSynth.js:55-56 warns about it): the root Multiply is written as the
    shorthand `(a * #o)`, and unDusp gives each operand of a top-level operator its own index
    (unDusp.js:10, constructObject.js:4), so `#o` is undeclared on the way back in."""
    text = dusp(d.unDusp("[Multiply A:[Osc #o f:[Ramp from:200 to:100 duration:2]] B:#o]"))
    with pytest.raises(d.DuspError, match="not been declared: #o"):
        d.unDusp(text)


def test_fmsynth_seed_is_written_back_as_dusp():
    d.config.seed(5)
    fm = P.FMSynth(P.FMSynth.randomSeed(f=110, nOscs=4, pConnection=0.4))
    s = fm.seed
    assert s["fundamental"] == 110 and s["mod"] == 1 and s["resetOscsOnTrigger"] is True
    assert len(s["oscs"]) == 4 and all(isinstance(o["h"], int) for o in s["oscs"])
    envs = s["envelopes"]
    assert envs and all(e.startswith("#Shape") for e in envs)   # already indexed while the oscs were written
    mixes = [o["mix"] for o in s["oscs"] if "mix" in o]
    assert mixes and all(m.startswith("([Shape #Shape") and 'shape:"decaySquared"' in m for m in mixes)
    # the written seed builds the same synth again (FMSynth.js:20-31 runs every entry through unDusp)
    d.config.seed(123)
    again = P.FMSynth({k: v for k, v in s.items() if k != "envelopes"})
    assert len(again.oscs) == 4
