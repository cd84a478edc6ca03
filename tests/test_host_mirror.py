"""Host logic of the reference-facing API: parser, constructor, setter sugar,
patches, errors."""
import math

import pytest

import dusp_b200 as d
from dusp_b200 import components as C, patches as P, parse, quick
from dusp_b200.graph import DuspError


def test_number_token_quirk():
    """getNumber.js:1-18: a number is a run of [0-9.-] => `200-100` is ONE token -> 200."""
    assert parse.getNumber("200-100")["n"] == 200
    assert parse.getNumber("-3.5x")["n"] == -3.5
    assert math.isnan(parse.getNumber("-")["n"])
    assert parse.getNumber("abc") is None


def test_operator_binding_order():
    t = parse.getExpression("1 + 2 * 3")
    assert t["operator"] == "+" and t["b"]["operator"] == "*"
    assert d.unDusp("1 + 2 * 3") == 7
    assert d.unDusp("(1 + 2) * 3") == 9
    assert d.unDusp("2 ^ 3") == 8


def test_object_parse_and_ids():
    t = parse.getExpression("[Osc #myOsc f:330]")
    assert t["type"] == "object" and t["constructor"] == "Osc" and t["id"] == "myOsc"
    assert t["attributes"][0]["property"] == "f"
    o = d.unDusp("[Multiply A:[Osc #yourOsc f:100] B:#yourOsc]")
    assert o.A.outlet.unit is o.B.outlet.unit and o.A.outlet.unit.label == "yourOsc"


def test_readme_ramp_trap():
    """Ramp has no inlets, so from/to/duration become plain properties."""
    osc = d.unDusp("[Osc f:[Ramp from:200 to:100 duration:2]]")
    ramp = osc.F.outlet.unit
    assert ramp.duration == 2 and ramp.y0 == 1 and ramp.y1 == 0 and not ramp.playing
    assert getattr(ramp, "from") == 200 and ramp.to == 100


def test_shorthands():
    o = d.unDusp("O200")
    assert isinstance(o, C.Osc) and o.F.constant == 200 and o.waveform == "sin"
    assert d.unDusp("Z10").waveform == "saw" and d.unDusp("Sq10").waveform == "square"
    e = d.unDusp("D2")
    assert isinstance(e, C.Shape) and e.playing and e.shape == "decay" and e.DURATION.constant == 2
    assert d.unDusp("A0.5").shape == "attack"
    f = d.unDusp("LP300")
    assert isinstance(f, C.Filter) and f.kind == "LP" and f.F.constant == 300
    ap = d.unDusp("AP0.01,0.5")
    assert isinstance(ap, C.AllPass) and ap.delayTimeInSamples == 441
    assert isinstance(d.unDusp("Noise"), C.Noise)


def test_operators_build_units():
    s = d.unDusp("(O200 + O250) @ 0")
    assert isinstance(s, C.Pan) and isinstance(s.IN.outlet.unit, C.Sum)
    assert isinstance(d.unDusp("O1 , O2"), C.ConcatChannels)
    assert isinstance(d.unDusp("O1 / 2"), C.Divide) and isinstance(d.unDusp("O1 - 2"), C.Subtract)
    f = d.unDusp("[Noise] -> LP200")
    assert isinstance(f, C.Filter) and isinstance(f.IN.outlet.unit, C.Noise)
    assert isinstance(d.unDusp("O1 >| 0.5"), C.HardClipAbove) and isinstance(d.unDusp("0.1 |< O1"), C.HardClipBelow)


def test_flags_and_string_attributes():
    s = d.unDusp("[Shape trigger]")
    assert s.playing
    o = d.unDusp("[Osc f:50 waveform:'saw']")
    assert o.waveform == "saw"
    with pytest.raises(DuspError):
        d.unDusp("[Osc waveform:'nope']")


def test_unknown_constructor_and_bad_reference():
    with pytest.raises(DuspError, match="Unknown object constructor"):
        d.unDusp("[Nope 1]")
    with pytest.raises(DuspError, match="not been declared"):
        d.unDusp("[Multiply A:#ghost]")


def test_for_then_operators():  # constructOperation.js:36-60
    a = d.unDusp("O1 for 2")
    assert isinstance(a, C.Osc) and len(a.events) == 1 and a.events[0].t == 2 * 44100
    r = d.unDusp("7 for 1")
    assert isinstance(r, C.Repeater) and r.IN.constant == 7  # a Number is wrapped in a Repeater first
    out = d.unDusp("(O1 for 2) then O2")
    first = out.IN.outlet.unit
    assert isinstance(out, C.Repeater) and first.F.constant == 1 and callable(first.onFinish)
    first.finish()  # what the scheduled event does: finished = true, then onFinish hands the Repeater its new source
    assert first.finished and out.IN.outlet.unit.F.constant == 2 and first.OUT.connections == []
    with pytest.raises(DuspError, match="scheduleFinish"):
        d.unDusp("'abc' for 2")


def test_retrigger_operators_build_chained_units():  # constructOperation.js:71-82
    env = d.unDusp("D1 ! 2")
    assert env.playing and isinstance(env, C.Shape)
    rt = [i.outlet.unit for i in env.inlets.values() if i.connected and i.noData]
    assert len(rt) == 1 and isinstance(rt[0], C.Retriggerer) and rt[0].target is env and rt[0].RATE.constant == 2
    env = d.unDusp("[Shape] ~! 3")
    assert not env.playing  # `~!` does not trigger at construction (constructOperation.js:78-82)
    rt = [i.outlet.unit for i in env.inlets.values() if i.connected and i.noData]
    assert isinstance(rt[0], C.SporadicRetriggerer) and rt[0].RATE.constant == 3
    with pytest.raises(DuspError):
        d.unDusp("5 ! 2")  # a Number has no trigger


def test_setter_sugar_and_constants():
    o = C.Osc()
    assert o.F.constant == 440 and o.F.chunk == [440.0]
    o.F = [100, 200]
    assert o.F.chunk == [100.0, 200.0]
    o.F = 5                      # Inlet.js:82-92: existing channels are all refilled
    assert o.F.chunk == [5.0, 5.0]
    m = C.Multiply(o, 0.5)
    assert m.A.connected and m.A.outlet is o.OUT and m.B.constant == 0.5
    m.A = 3
    assert not m.A.connected and o.OUT.connections == []
    with pytest.raises(DuspError, match="Passed bad value"):
        m.A = None


def test_js_or_defaults():
    assert C.Osc(0).F.constant == 440        # `f || 440`
    assert C.Multiply(0, 0).A.constant == 1  # `a || 1`
    assert C.Ramp(0, 0, 0).duration == 44100 and C.Ramp(0, 0, 0).y0 == 1


def test_patch_alias():
    lfo = P.LFO(2, 0.3, 0.1)
    assert lfo.F.constant == 2 and lfo.A.constant == 0.3 and lfo.O.constant == 0.1
    lfo.F = 7
    assert lfo.osc.F.constant == 7
    assert lfo.defaultOutlet.unit.KIND == "Sum"
    sd = P.SimpleDelay()
    assert sd.DELAY.constant == 4410 and sd.DRYWET.constant == 0.4


def test_fixed_delay_validation():
    with pytest.raises(DuspError, match="length 0"):
        C.FixedDelay(0)
    assert C.FixedDelay(0.01).delayTimeInSamples == 441


def test_quick_constant_folding():
    assert quick.add(1, 2) == 3 and quick.multiply(3, 4) == 12 and quick.multiply(1, 5) == 5
    o = C.Osc()
    assert quick.multiply(o, 1) is o and isinstance(quick.multiply(o, 2), C.Multiply)
    assert quick.concat(1, [2, 3]) == [1, 2, 3]


def test_circuit_conflict_and_bad_outlet():
    from dusp_b200.render import circuit_blob
    with pytest.raises(DuspError, match="expects an outlet"):
        circuit_blob(None)
    with pytest.raises(DuspError, match="expects an outlet"):
        circuit_blob(42)
    a, b = C.Osc(), C.Osc()
    d.Circuit(a)
    d.Circuit(b)
    with pytest.raises(DuspError):
        C.Multiply(a, b)


def test_incremental_circuit_growth_keeps_order_valid():
    a = C.Osc(100)
    circ = d.Circuit(a)
    m = C.Multiply(a, 0.5)  # Inlet.connect adds the new unit to a's circuit (Inlet.js:60-73)
    assert m.circuit is circ and circ.units.index(a) < circ.units.index(m)


def test_cases_json_is_current():
    """bindings/node/cases.json (what a Node host renders on the untouched reference, ref_dump.js) covers every case of
    tests/circuits.py and says how each one is rebuilt."""
    import json, os
    import circuits
    path = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "bindings", "node", "cases.json")
    cases = json.load(open(path))
    assert set(cases) == set(circuits.CASES)
    for name, c in cases.items():
        assert c["seconds"] == circuits.CASES[name][1] and c["exact"] == bool(circuits.CASES[name][2])
        assert sum(k in c for k in ("dusp", "blob", "skip")) == 1, name
    assert sum("skip" in c for c in cases.values()) <= 6


# ---- structure of every unit class against the reference's own source text (tests/golden/ref_structure.json, written by
#      tests/golden/make_ref_structure.py in the build container, where /root/reference exists)
_CTOR_ARGS = {  # classes whose constructor refuses to run without arguments (as the reference's does)
    "AllPass": (0.01, 0.5), "CombFilter": (0.01, 0.5), "FixedDelay": (0.01,), "Clip": (0.5,), "Pow": (1, 2), "Monitor": (0,),
    "FFT": (4096, 1024), "IFFT": (4096, 1024), "Hopper": (1024, 4096), "UnHopper": (1024, 4096), "Windower": (4096, "hamming", 1024),
    "ReChunk": (256, 4096),
    "HardLowPass": (1000,), "HardHighPass": (1000,), "SpectralSum": (0, 0),
}


def _ref_structure():
    import json, os
    return json.load(open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "ref_structure.json")))


def _chain(ref, name, key):
    r = ref.get(name)
    if r is None:
        return []
    return _chain(ref, r["base"], key) + r[key]


def test_unit_classes_have_the_references_ports_and_defaults():
    """Port names, their order, `mono` flags, literal channel counts, constructor parameter order and `param || number`
    defaults of every component class, against what the reference's source declares.  (The oracle and the device are
    fed by this mirror: a wrong port or default here would be identically wrong on both sides of every parity test.)"""
    import inspect
    from dusp_b200 import components as C
    ref = _ref_structure()["components"]
    checked = fields_checked = 0
    for name, r in sorted(ref.items()):
        cls = getattr(C, name, None)
        if cls is None:
            assert name in ("CircularMotion", "LinearMotion", "GreaterThan", "LessThan", "SpectralUnit", "SignalCombiner"), name
            continue
        if name in ("CircleBufferReader", "CircleBufferWriter", "CircleBufferNode"):
            u = cls(C.CircleBuffer(1, 1.0))
        else:
            u = cls(*_CTOR_ARGS.get(name, ()))
        want_in = [(i["name"], bool(i.get("mono", False))) for i in _chain(ref, name, "inlets")]
        got_in = [(i.name, i.exposeAsMono) for i in u.inletsOrdered if not i.noData]
        assert got_in == want_in, (name, got_in, want_in)
        want_out = _chain(ref, name, "outlets")
        got_out = [o for o in u.outletsOrdered if not o.noData]
        assert [o.name for o in got_out] == [o["name"] for o in want_out], (name, [o.name for o in got_out], want_out)
        for o, w in zip(got_out, want_out):
            if "numberOfChannels" in w:
                assert o.numberOfChannels == w["numberOfChannels"], (name, o.name)
            if w.get("mono"):
                assert o.numberOfChannels == 1, (name, o.name)
        params = [p for p in inspect.signature(cls.__init__).parameters if p != "self"]
        assert len(params) >= len(r["params"]), (name, params, r["params"])
        if name not in _CTOR_ARGS:
            defaults = {}
            k = name
            while k in ref:  # a subclass inherits its base's `this.A = a || 0`
                for port, dv in ref[k]["defaults"].items():
                    defaults.setdefault(port, dv)
                k = ref[k]["base"]
            for port, dv in defaults.items():
                inlet = next(i for i in u.inletsOrdered if i.name.upper() == port)
                assert not inlet.connected and float(inlet.constant) == dv["default"], (name, port, inlet.constant, dv)
        for fld, val in r["fields"].items():  # state the constructor sets to a literal: this.phase = 0, this.compensationDB = 1.5 ...
            if hasattr(u, fld) and isinstance(getattr(u, fld), (int, float, bool)) and name not in _CTOR_ARGS:
                assert float(getattr(u, fld)) == val, (name, fld, getattr(u, fld), val)
                fields_checked += 1
        checked += 1
    assert checked >= 50 and fields_checked >= 12


def test_patches_declare_the_references_aliases():
    """Every patch of the mirror exposes the inlets / outlets the reference's constructor aliases, under the same names
    in the same order (src/patches/*.js, Patch.js:24-60)."""
    import inspect
    from dusp_b200 import patches as P
    ref = _ref_structure()["patches"]
    args = {"MultiTapDelay": (1, 1.0), "DelayMixer": (1, 1.0), "APStack": (2,), "Mixer": (), "HardBandPass": (), "SpaceChannel": ()}
    checked = 0
    for name, r in sorted(ref.items()):
        cls = getattr(P, name, None)
        if cls is None or not r["aliases"]:
            continue
        try:
            u = cls(*args.get(name, ()))
        except Exception:
            continue  # needs constructor arguments this table does not have
        want_in = [a["name"] for a in r["aliases"] if a["kind"] == "inlet"]
        want_out = [a["name"] for a in r["aliases"] if a["kind"] == "outlet"]
        got_in, got_out = list(u.inlets), list(u.outlets)
        assert [x.lower() for x in got_in][:len(want_in)] == [x.lower() for x in want_in] or set(x.lower() for x in want_in) <= set(x.lower() for x in got_in), (name, got_in, want_in)
        assert set(x.lower() for x in want_out) <= set(x.lower() for x in got_out), (name, got_out, want_out)
        params = [p for p in inspect.signature(cls.__init__).parameters if p != "self"]
        assert len(params) >= len(r["params"]), (name, params, r["params"])
        checked += 1
    assert checked >= 12, checked


def test_front_door_tables_match_the_reference():
    """Shorthand constructors (shorthandConstructors.js), binary operators (constructOperation.js) and the names an
    object expression `[Name ...]` can construct (components/index.js, patches/index.js), from the reference's text."""
    import inspect
    from dusp_b200 import construct, components as C, patches as P
    fd = _ref_structure()["front_door"]
    assert set(fd["shorthands"]) == set(construct.SHORTHAND_CONSTRUCTORS)
    for k, params in fd["shorthands"].items():
        assert len(inspect.signature(construct.SHORTHAND_CONSTRUCTORS[k]).parameters) == len(params), k
    src = inspect.getsource(construct)
    for op in fd["operators"]:
        assert 'op == "%s"' % op in src, op
    not_mirrored = {"GreaterThan", "LessThan", "CircularMotion", "LinearMotion"}  # vector units / units that throw at construction
    missing_c = [n for n in fd["exported"]["components"] if not hasattr(C, n) and n not in not_mirrored]
    assert missing_c in ([], ["vector", "spectral"]) or all(n in ("vector", "spectral") for n in missing_c), missing_c
    missing_p = [n for n in fd["exported"]["patches"] if not hasattr(P, n)]
    assert set(missing_p) <= {"ComplexOrbit", "OrbittySine", "SineCloud"}, missing_p  # patches built on the vector units


def test_config_constants_match_the_reference():
    """src/config.js:5-16 against dusp_b200/config.py and the constants compiled into the library."""
    from dusp_b200 import config
    c = _ref_structure()["config"]
    assert config.standardChunkSize == c["standardChunkSize"] == 256 and config.sampleRate == c["sampleRate"] == 44100
    assert config.fft["windowSize"] == c["windowSize"] and c["hopSize_expr"] == "4096/4" and config.fft["hopSize"] == 1024
    assert config.channelFormat == c["channelFormat"] and config.fft["windowKind"] == c["windowKind"]
