"""Writes bindings/node/cases.json: every parity case of tests/circuits.py in a form a Node host can rebuild on the
UNTOUCHED reference (bindings/node/ref_dump.js), so that pinning the oracle is one command there.
Per case: seconds, exact, host seed, and
  "dusp": the Dusp text, when the factory is nothing but `return d.unDusp("...")`, else
  "blob": the DUSPGRAPH text of the circuit (single-section blobs only: units, properties, inlets, EV records), else
  "skip": why ref_dump.js cannot rebuild it (graph edits or Retriggerers built through the Python API).
usage: python tests/golden/make_cases_json.py"""
import inspect, json, os, re, sys

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import dusp_b200
import circuits
from dusp_b200.render import circuit_blob

out = {}
for name in sorted(circuits.CASES):
    fac, seconds, exact = circuits.CASES[name]
    src = inspect.getsource(fac)
    body = [l.strip() for l in src.split("\n")[1:] if l.strip() and not l.strip().startswith("#")]
    entry = {"seconds": seconds, "exact": bool(exact), "seed": 12345}
    m = re.match(r'''return d\.unDusp\((["'])(.*)\1\)\s*(#.*)?$''', body[0]) if len(body) == 1 else None
    if m:
        entry["dusp"] = m.group(2)
    else:
        dusp_b200.config.seed(12345)
        blob, _ = circuit_blob(fac())
        if "\nEPOCH " in blob:
            entry["skip"] = "graph edit built through the unit API (no Dusp text)"
        elif "Retriggerer" in blob:
            entry["skip"] = "Retriggerer built through the unit API (its target is a JS object reference)"
        else:
            entry["blob"] = blob
    out[name] = entry
path = os.path.join(ROOT, "bindings", "node", "cases.json")
json.dump(out, open(path, "w"), indent=0, sort_keys=True)
kinds = {k: sum(1 for e in out.values() if k in e) for k in ("dusp", "blob", "skip")}
print("wrote %d cases to %s: %s" % (len(out), path, kinds))
print("skipped:", sorted(n for n, e in out.items() if "skip" in e))
