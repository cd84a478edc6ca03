"""Committed fixtures (tests/golden/golden.npz, made by tests/golden/make_golden.py): the oracle must keep
producing them (CPU), and the CUDA path must reproduce them through the C ABI (-m gpu).  Bit-exact cases are
compared by SHA-256 of the whole render, tolerance-class cases on the stored head."""
import hashlib
import os

import numpy as np
import pytest

import circuits
from dusp_b200.render import circuit_blob, n_samples_for

GOLD = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "golden.npz"))
TOL = 1e-5


def check(name, arr, exact):
    assert tuple(GOLD[name + ".shape"]) == arr.shape
    head = GOLD[name + ".head"]
    if exact:
        assert np.array_equal(arr[:, :head.shape[1]], head), name
        assert hashlib.sha256(arr.tobytes()).digest() == GOLD[name + ".sha256"].tobytes(), name
    else:
        full = max(1.0, float(np.max(np.abs(head))) if head.size else 1.0)
        err = float(np.max(np.abs(arr[:, :head.shape[1]].astype(np.float64) - head))) if head.size else 0.0
        assert err <= TOL * full, "%s: %g" % (name, err)


def test_fixture_covers_every_case():
    assert {k[:-5] for k in GOLD.files if k.endswith(".head")} == set(circuits.CASES)


@pytest.mark.parametrize("name", sorted(circuits.CASES))
def test_oracle_reproduces_golden(oracle, name):
    fac, seconds, _ = circuits.CASES[name]
    blob, _ = circuit_blob(fac())
    check(name, np.stack(oracle.render(blob, seconds)).astype(np.float32), True)


@pytest.mark.gpu
@pytest.mark.parametrize("name", sorted(circuits.CASES))
def test_gpu_reproduces_golden(lib, name):
    from dusp_b200 import capi
    fac, seconds, exact = circuits.CASES[name]
    blob, _ = circuit_blob(fac())
    from dusp_b200.render import SessionRun
    run = SessionRun([blob], [1], 0)
    try:
        got = run.render(n_samples_for(seconds))[0]
    finally:
        run.close()
    check(name, got, exact)


REF = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "golden_ref.npz")


@pytest.mark.skipif(not os.path.exists(REF), reason="no renders of the reference itself (needs a Node host: bindings/node/ref_dump.js)")
def test_oracle_matches_reference_renders(oracle):
    """Present only after tests/golden/import_ref.py: pins the oracle to the untouched reference."""
    ref = np.load(REF)
    for name in ref.files:
        fac, seconds, exact = circuits.CASES[name]
        blob, _ = circuit_blob(fac())
        got = np.stack(oracle.render(blob, seconds)).astype(np.float32)
        assert got.shape == ref[name].shape
        err = float(np.max(np.abs(got.astype(np.float64) - ref[name])))
        assert err <= (0.0 if exact else TOL), "%s: %g" % (name, err)
