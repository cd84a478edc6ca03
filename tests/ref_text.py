"""Test infrastructure: the reference's own source text for the tests that execute it (tests/js2py.py, tests/ref_runtime.py).
Nothing of that text is committed: it is extracted at test time from /root/reference, which exists in the build container
only -- elsewhere these tests skip.  (The `-m gpu` tests, smoke() and bench.py never come here.)"""
import json
import os
import subprocess
import sys
import tempfile

REFERENCE = os.environ.get("DUSP_REFERENCE_DIR", "/root/reference")
_cache = {}


def available():
    return os.path.isdir(os.path.join(REFERENCE, "src", "components"))


def reference_text():
    """{"tick_expr": ..., "tick_body": ...} as tests/golden/make_ref_structure.py extracts them; pytest.skip without the reference."""
    if "text" not in _cache:
        if not available():
            import pytest
            pytest.skip("the reference's source is not on this machine (%s)" % REFERENCE)
        here = os.path.dirname(os.path.abspath(__file__))
        with tempfile.TemporaryDirectory() as tmp:
            path = os.path.join(tmp, "ref_text.json")
            subprocess.check_call([sys.executable, os.path.join(here, "golden", "make_ref_structure.py"), REFERENCE, path])
            _cache["text"] = json.load(open(path))
    return _cache["text"]
