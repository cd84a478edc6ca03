"""Sweep of the specialised kernel's tile shape and residency: DUSP_TILE="B,T" x DUSP_FUSED_WPS=W.
usage: perf_probe3.py "8,4,8;8,8,17;4,8,17" w3:10000:88200:0 [w2:...]   (spec = workload:voices:samples:mode)"""
import sys, os, subprocess
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
if len(sys.argv) > 1 and sys.argv[1] == "child":
    import scripts.perf_probe as pp
    from dusp_b200 import workloads as W
    from dusp_b200.render import circuit_blob
    blobs = {"w3": circuit_blob(W.w3_circuit())[0], "w2": circuit_blob(W.w2_live_circuit())[0],
             "w3b": circuit_blob(W.w3b_circuit())[0]}
    tag = os.environ.get("DUSP_TILE", "auto") + "/w" + os.environ.get("DUSP_FUSED_WPS", "auto") + "/" + os.environ.get("DUSP_FUSED_DEFS", "")
    for spec in sys.argv[2:]:
        name, nv, ns, mode = spec.split(":")
        pp.probe(name + "@" + tag, blobs[name], int(nv), int(ns), int(mode), reps=2)
else:
    for t in sys.argv[1].split(";"):
        env = dict(os.environ)
        if t != "auto":
            f = t.split(",")
            env["DUSP_TILE"] = f[0] + "," + f[1]
            if len(f) > 2:
                env["DUSP_FUSED_WPS"] = f[2]
            if len(f) > 3:
                env["DUSP_FUSED_DEFS"] = ",".join(f[3:])
        subprocess.call([sys.executable, __file__, "child"] + sys.argv[2:], env=env)
