"""Turns the output directory of bindings/node/ref_dump.js (renders of the UNTOUCHED reference, made on a
Node-equipped host) into tests/golden/golden_ref.npz; tests/test_golden.py then also checks the oracle against it.
usage: python tests/golden/import_ref.py out_dir"""
import json, os, sys
import numpy as np

src = sys.argv[1]
index = json.load(open(os.path.join(src, "index.json")))
if "cases" in index:  # ref_dump.js of round 2: {"cases": {...}, "failed": {...}}
    for name, why in index.get("failed", {}).items():
        print("reference render FAILED for", name, "--", why.splitlines()[0])
    index = {k: v for k, v in index["cases"].items() if "skipped" not in v}
out = {}
for name, meta in index.items():
    a = np.fromfile(os.path.join(src, name + ".f32"), dtype="<f4").reshape(meta["channels"], meta["samples"])
    out[name] = a
np.savez_compressed(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden_ref.npz"), **out)
print("imported", sorted(out))
