"""Extracts, from the REFERENCE'S OWN SOURCE TEXT (/root/reference/src, present only in the build container), the
structural facts the host mirror must share with it: per unit class the constructor's parameter names, the inlets in
declaration order with their `mono` flag, the outlets with their channel count where it is a literal, the defaults
of the form `this.NAME = param || <number>`, and the state fields the constructor sets to a literal (`this.phase = 0`).  Writes tests/golden/ref_structure.json (committed);
tests/test_host_mirror.py compares the Python mirror against it, so a wrong port name, port order or default in the
mirror -- which the oracle, fed by the same mirror, could not see -- fails a test.

The same pass also pulls the verbatim per-sample statements / `_tick` bodies out of the source; those are NOT committed:
with a second argument the script writes everything to that scratch path, which is how tests/ref_text.py hands the
reference's own text to the tests that execute it (only where /root/reference exists).

usage: python tests/golden/make_ref_structure.py [/root/reference [scratch.json]]"""
import json, os, re, sys

ref = sys.argv[1] if len(sys.argv) > 1 else "/root/reference"
src = os.path.join(ref, "src")
out = {}


def parse_opts(text):
    o = {}
    if text is None:
        return o
    m = re.search(r"mono\s*:\s*(true|false)", text)
    if m:
        o["mono"] = m.group(1) == "true"
    m = re.search(r"numberOfChannels\s*:\s*(\d+)", text)
    if m:
        o["numberOfChannels"] = int(m.group(1))
    return o


def scan(path, rel):
    text = open(path).read()
    m = re.search(r"^function\s+([A-Z]\w*)\s*\(([^)]*)\)\s*\{", text, re.M)
    cls_style = False
    if not m:
        mc = re.search(r"^class\s+([A-Z]\w*)\s+extends\s+(\w+)", text, re.M)
        m2 = re.search(r"constructor\s*\(([^)]*)\)\s*\{", text)
        if not mc:
            return
        name, base, params = mc.group(1), mc.group(2), (m2.group(1) if m2 else "")
        cls_style = True
    else:
        name, params = m.group(1), m.group(2)
        mb = re.search(r"^\s*(\w+)\.call\(this", text, re.M)
        base = mb.group(1) if mb else None
    # constructor body: up to the first prototype assignment / the first method of the class
    start = (m2.end() if cls_style and m2 else m.end() if m else 0)
    if cls_style:
        depth, i = 1, start
        while i < len(text) and depth:
            depth += text[i] == "{"
            depth -= text[i] == "}"
            i += 1
        body = text[start:i]
    else:
        end = re.search(r"^\}", text[start:], re.M)
        body = text[start:start + end.start()] if end else text[start:]
    inlets = [{"name": a, **parse_opts(b)} for a, b in re.findall(r"this\.add(?:Spectral)?Inlet\(\s*[\"'](\w+)[\"']\s*(?:,\s*\{([^}]*)\})?", body)]
    outlets = [{"name": a, **parse_opts(b)} for a, b in re.findall(r"this\.add(?:Spectral)?Outlet\(\s*[\"'](\w+)[\"']\s*(?:,\s*\{([^}]*)\})?", body)]
    defaults = {}
    for port, param, dflt in re.findall(r"this\.([A-Z][A-Z0-9_]*)\s*=\s*(\w+)\s*\|\|\s*(-?[0-9.]+(?:e-?\d+)?)\s*$", body, re.M):
        defaults[port] = {"param": param, "default": float(dflt)}
    fields = {}
    for fld, val in re.findall(r"^\s*this\.([a-z]\w*)\s*=\s*(-?[0-9.]+(?:e-?\d+)?|true|false)\s*;?\s*(?://.*)?$", body, re.M):
        fields[fld] = {"true": 1.0, "false": 0.0}.get(val, None) if val in ("true", "false") else float(val)
    out[name] = {"fields": fields, "file": rel, "base": base, "params": [p.strip().split("=")[0].strip() for p in params.split(",") if p.strip()],
                 "inlets": inlets, "outlets": outlets, "defaults": defaults}


for root, _, files in os.walk(os.path.join(src, "components")):
    for f in sorted(files):
        if f.endswith(".js") and f != "index.js":
            scan(os.path.join(root, f), os.path.relpath(os.path.join(root, f), ref))
# patches: the aliases a patch's constructor declares, in order (Patch.js:24-60: the name defaults to the aliased port's)
patches = {}
for f in sorted(os.listdir(os.path.join(src, "patches"))):
    if not f.endswith(".js") or f == "index.js":
        continue
    text = open(os.path.join(src, "patches", f)).read()
    m = re.search(r"^function\s+([A-Z]\w*)\s*\(([^)]*)\)", text, re.M) or re.search(r"^class\s+([A-Z]\w*)\s+extends", text, re.M)
    if not m:
        continue
    mc = re.search(r"constructor\s*\(([^)]*)\)", text)
    params = m.group(2) if m.lastindex and m.lastindex >= 2 else (mc.group(1) if mc else "")
    al = []
    for kind, port, nm in re.findall(r"this\.alias(Inlet|Outlet)\(\s*[\w.\[\]\-+ ]*?\.([A-Z][A-Z0-9]*)\s*(?:,\s*[\"'](\w+)[\"'])?\s*\)", text):
        al.append({"kind": kind.lower(), "name": nm or port.lower(), "explicit": bool(nm)})
    patches[m.group(1)] = {"file": "src/patches/" + f, "aliases": al,
                           "params": [q.strip().split("=")[0].strip() for q in params.split(",") if q.strip()]}
# the Dusp front door: shorthand constructors, binary operators, the names `[Name ...]` can construct
text = open(os.path.join(src, "construct", "shorthandConstructors.js")).read()
shorthands = re.findall(r"^\s{2}(\w+)\s*:\s*function\s*\(([^)]*)\)", text, re.M)
text = open(os.path.join(src, "construct", "constructOperation.js")).read()
operators = re.findall(r"case\s+\"([^\"]+)\"\s*:", text)
exported = {}
for pkg in ("components", "patches"):
    text = open(os.path.join(src, pkg, "index.js")).read()
    exported[pkg] = re.findall(r"^\s*(\w+)\s*:\s*require\(", text, re.M)
front = {"shorthands": {k: [q.strip() for q in v.split(",") if q.strip()] for k, v in shorthands}, "operators": operators, "exported": exported}
text = open(os.path.join(src, "config.js")).read()
cfg = {k: float(v) for k, v in re.findall(r"^\s*(standardChunkSize|sampleRate|windowSize)\s*:\s*([0-9.]+)", text, re.M)}
cfg["hopSize_expr"] = re.search(r"hopSize\s*:\s*([^,\n]+)", text).group(1).strip()
cfg["channelFormat"] = re.search(r"channelFormat\s*:\s*\"(\w+)\"", text).group(1)
cfg["windowKind"] = re.search(r"windowKind\s*:\s*\"(\w+)\"", text).group(1)
# the per-sample statement of the stateless units' _tick, verbatim, with the local names that alias a port's channel:
# tests/test_oracle_kat.py translates these statements (JavaScript arithmetic is Python arithmetic) and runs them
# against the oracle on known inputs
tick_expr = {}
for u in ["Multiply", "Divide", "Subtract", "Sum", "Gain", "Clip", "HardClipAbove", "HardClipBelow", "Pow", "Rescale", "CrossFader",
          "PolarityInvert", "Abs", "DecibelToScaler", "SemitoneToRatio", "MidiToFrequency", "SecondsToSamples", "FixedMultiply", "Pan"]:
    text = open(os.path.join(src, "components", u + ".js")).read()
    m = re.search(r"_tick\s*(?:=\s*function)?\s*\([^)]*\)\s*\{", text)
    body = text[m.end():]
    depth, i = 1, 0
    while i < len(body) and depth:
        depth += body[i] == "{"
        depth -= body[i] == "}"
        i += 1
    body = body[:i - 1]
    aliases = dict(re.findall(r"var\s+(\w+)\s*=\s*this\.(\w+)\[", body))
    inner = body[body.rfind("for(var t"):]
    hdr = re.match(r"for\(var t[^)]*\)\s*\{?", inner)
    stmt = inner[hdr.end():].strip().rstrip("}").strip()
    helpers = re.findall(r"^function\s+(\w+)\s*\(([^)]*)\)\s*\{[^\n]*\n\s*return\s+([^\n;]+)", text, re.M)
    tick_expr[u] = {"aliases": aliases, "stmt": stmt, "helpers": {h: {"params": p, "expr": e.strip()} for h, p, e in helpers}}
# whole _tick bodies of the stateful hot-path units, verbatim: tests/js2py.py
# translates them line by line and tests/test_oracle_kat.py runs them against the oracle
def body_after(text, pattern):
    m = re.search(pattern, text, re.M)
    b = text[m.end():]
    depth, i = 1, 0
    while depth:
        depth += b[i] == "{"
        depth -= b[i] == "}"
        i += 1
    return b[:i - 1]


comp = os.path.join(src, "components")
tick_body = {}
for u, f, pat in (("Osc", "Osc/Osc.js", r"_tick = function\(clock\) \{"), ("Ramp", "Ramp.js", r"_tick = function\(\) \{"),
                  ("Filter", "Filter.js", r"_tick = function\(\) \{"), ("Noise", "Noise.js", r"_tick = function\(\) \{"),
                  ("FixedDelay", "FixedDelay.js", r"_tick\(\) \{"), ("CombFilter", "CombFilter.js", r"_tick\(\) \{"),
                  ("AllPass", "AllPass.js", r"_tick\(\) \{"), ("Timer", "Timer.js", r"_tick = function\(\) \{"),
                  ("Shape", "Shape/index.js", r"_tick = function\(\) \{")):
    tick_body[u] = body_after(open(os.path.join(comp, f)).read(), pat)
# ... and every other non-spectral component's whole _tick, keyed by class name (the first nine above win)
for f in sorted(os.listdir(comp)):
    if not f.endswith(".js") or f == "index.js":
        continue
    text = open(os.path.join(comp, f)).read()
    mname = re.search(r"^function\s+([A-Z]\w*)\s*\(|^class\s+([A-Z]\w*)\s+extends", text, re.M)
    mt = re.search(r"_tick\s*(?:=\s*function)?\s*\(([^)]*)\)\s*\{", text)
    if mname and mt:
        nm = mname.group(1) or mname.group(2)
        tick_body.setdefault(nm, body_after(text, r"_tick\s*(?:=\s*function)?\s*\([^)]*\)\s*\{"))
        tick_body[nm + ".args"] = mt.group(1).strip()
# trigger() / stop() of the units a Retriggerer can target
for u, f in (("Ramp", "Ramp.js"), ("Shape", "Shape/index.js"), ("AHD", "AHD.js"), ("Timer", "Timer.js")):
    text = open(os.path.join(comp, f)).read()
    for meth in ("trigger", "stop"):
        m = re.search(r"(?:\.prototype\.%s\s*=\s*function\s*\(\)|^\s*%s\(\))\s*\{" % (meth, meth), text, re.M)
        if m:
            tick_body[u + "." + meth] = body_after(text, r"(?:\.prototype\.%s\s*=\s*function\s*\(\)|^\s*%s\(\))\s*\{" % (meth, meth))
text = open(os.path.join(comp, "SporadicRetrigger.js")).read()
tick_body["SporadicRetriggerer"] = body_after(text, r"_tick\(\) \{")
text = open(os.path.join(comp, "Osc", "MultiChannelOsc.js")).read()
tick_body["MultiChannelOsc"] = body_after(text, r"_tick\s*(?:=\s*function)?\s*\([^)]*\)\s*\{")
text = open(os.path.join(comp, "Filter.js")).read()
tick_body["Filter.LP"] = body_after(text, r"LP: function\(f\) \{")
tick_body["Filter.HP"] = body_after(text, r"HP: function\(f\) \{")
text = open(os.path.join(comp, "Osc", "waveTables.js")).read()
tick_body["sineTable"] = text[text.index("const PHI"):text.index("var sawTable")]
tick_body["CircleBufferWriter"] = body_after(open(os.path.join(comp, "CircleBufferWriter.js")).read(), r"_tick\(\) \{")
tick_body["CircleBufferReader"] = body_after(open(os.path.join(comp, "CircleBufferReader.js")).read(), r"_tick\(\) \{")
tick_body["MonoDelay"] = body_after(open(os.path.join(comp, "MonoDelay.js")).read(), r"_tick = function\(clock\) \{")
tick_body["Delay"] = body_after(open(os.path.join(comp, "Delay.js")).read(), r"_tick\(clock\) \{")
tick_body["ReadBackDelay"] = body_after(open(os.path.join(comp, "ReadBackDelay.js")).read(), r"_tick = function\(\) \{")
text = open(os.path.join(src, "CircleBuffer.js")).read()
tick_body["CircleBuffer.read"] = body_after(text, r"read\(c, t\) \{")
tick_body["CircleBuffer.write"] = body_after(text, r"write\(c, t, y\) \{")
tick_body["CircleBuffer.mix"] = body_after(text, r"mix\(c, t, y\) \{")
out = {"components": out, "patches": patches, "front_door": front, "config": cfg, "tick_expr": tick_expr, "tick_body": tick_body}
if len(sys.argv) > 2:
    # everything, the verbatim statements included: a scratch file for the tests that execute the reference's text
    # (tests/ref_text.py runs this script against /root/reference at test time; that text is never committed)
    json.dump(out, open(sys.argv[2], "w"))
else:
    # the committed fixture: structural FACTS only (names, orders, flags, numbers) -- no source text
    dst = os.path.join(os.path.dirname(os.path.abspath(__file__)), "ref_structure.json")
    json.dump({k: v for k, v in out.items() if k not in ("tick_expr", "tick_body")}, open(dst, "w"), indent=1, sort_keys=True)
    print("wrote %d unit classes and %d patches to %s" % (len(out["components"]), len(out["patches"]), dst))
