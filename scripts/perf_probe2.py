"""Tile-shape sweep: device-timed W3 / W2 renders under DUSP_TILE overrides."""
import sys, os, json, subprocess
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
if len(sys.argv) > 1 and sys.argv[1] == "child":
    import scripts.perf_probe as pp
    from dusp_b200 import capi
    from dusp_b200.render import circuit_blob
    sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
    import circuits
    b_w2, _ = circuit_blob(circuits.w2_live())
    b_w3, _ = circuit_blob(circuits.w3_voice())
    for spec in sys.argv[2:]:
        name, nv, ns, mode = spec.split(":")
        pp.probe(name + "@" + os.environ.get("DUSP_TILE", "auto"), b_w3 if name == "w3" else b_w2, int(nv), int(ns), int(mode), reps=2)
else:
    tiles = sys.argv[1].split(";")
    for t in tiles:
        env = dict(os.environ)
        if t != "auto":
            env["DUSP_TILE"] = t
        subprocess.call([sys.executable, __file__, "child"] + sys.argv[2:], env=env)
