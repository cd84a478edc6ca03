"""Regenerates tests/golden/golden.npz: for every circuit of tests/circuits.py the first 4096 samples per channel
of the ORACLE's render (float32) and a SHA-256 of the whole render.

Status: these pin the oracle (and, in the -m gpu tests, the CUDA path) against regressions; they are NOT outputs of
the reference itself -- it is JavaScript and cannot run in this image (no Node).  `bindings/node/ref_dump.js` writes
the same cases from the untouched reference on a Node-equipped host; drop its golden_ref.npz next to this file and
tests/test_golden.py will compare against it too.

usage: python tests/golden/make_golden.py"""
import hashlib, os, sys
import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import dusp_b200
import circuits
from dusp_b200.render import circuit_blob
from oracle import oracle

HEAD = 4096


def render_case(name):
    dusp_b200.config.seed(12345)  # the injected generator's seed (tests/conftest.py)
    fac, seconds, _ = circuits.CASES[name]
    blob, _ = circuit_blob(fac())
    return oracle.render(blob, seconds)


def main():
    oracle.build()
    out = {}
    for name in sorted(circuits.CASES):
        chans = render_case(name)
        arr = np.stack(chans).astype(np.float32)
        out[name + ".head"] = arr[:, :HEAD].copy()
        out[name + ".sha256"] = np.frombuffer(hashlib.sha256(arr.tobytes()).digest(), dtype=np.uint8)
        out[name + ".shape"] = np.array(arr.shape, dtype=np.int64)
    np.savez_compressed(os.path.join(HERE, "golden.npz"), **out)
    print("wrote %d cases" % len(circuits.CASES))


if __name__ == "__main__":
    main()
