"""Development aid (GPU box): which tolerance-class cases of tests/circuits.py are in fact bit-exact on every kernel path?"""
import os, sys, subprocess, json
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
if len(sys.argv) > 1:
    import numpy as np
    import circuits
    from dusp_b200.render import circuit_blob, SessionRun, n_samples_for
    from oracle import oracle
    import dusp_b200
    res = {}
    for name, (fac, secs, exact) in sorted(circuits.CASES.items()):
        if exact:
            continue
        dusp_b200.config.seed(12345)
        blob, _ = circuit_blob(fac())
        want = oracle.render(blob, secs)
        run = SessionRun([blob], [1], 0)
        got = run.render(n_samples_for(secs))[0]
        run.close()
        res[name] = int(sum(int(np.sum(got[c] != w)) for c, w in enumerate(want)))
    print(json.dumps({sys.argv[1]: res}))
else:
    for mode, env in (("interpreter", {"DUSP_FUSED": "0", "DUSP_PIPED": "0"}), ("specialised", {"DUSP_FUSED": "1", "DUSP_PIPED": "0"}),
                      ("pipelined", {"DUSP_FUSED": "1", "DUSP_PIPED": "1"})):
        e = dict(os.environ); e.update(env)
        subprocess.call([sys.executable, __file__, mode], env=e)
