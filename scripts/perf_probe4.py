"""Pipelined-kernel probe: DUSP_PIPED x voices-per-CTA, against the tile kernel.
usage: perf_probe4.py "p24;p32;p16;t" w3:10000:88064:0 ...   (pNN = pipelined with NN voices/CTA, t = tile kernel;
optional ",DEF1,DEF2" suffix = DUSP_FUSED_DEFS)"""
import sys, os, subprocess
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
if len(sys.argv) > 1 and sys.argv[1] == "child":
    import scripts.perf_probe as pp
    from dusp_b200 import workloads as W
    from dusp_b200.render import circuit_blob
    blobs = {"w3": circuit_blob(W.w3_circuit())[0], "w2": circuit_blob(W.w2_live_circuit())[0],
             "w3b": circuit_blob(W.w3b_circuit())[0]}
    for spec in sys.argv[2:]:
        name, nv, ns, mode = spec.split(":")
        pp.probe(name + "@" + os.environ["PROBE_TAG"], blobs[name], int(nv), int(ns), int(mode), reps=2)
else:
    for t in sys.argv[1].split(";"):
        env = dict(os.environ)
        f = t.split(",")
        env["PROBE_TAG"] = t
        if f[0].startswith("p"):
            env["DUSP_PIPED"] = "1"
            env["DUSP_PIPED_VC"] = f[0][1:]
        else:
            env["DUSP_PIPED"] = "0"
        if len(f) > 1:
            env["DUSP_FUSED_DEFS"] = ",".join(f[1:])
        subprocess.call([sys.executable, __file__, "child"] + sys.argv[2:], env=env)
