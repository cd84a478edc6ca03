"""The streaming sink (dusp_b200.stream.RenderStream, mirror of src/RenderStream.js:6-80): interleaved 256-frame
chunks, running peak auto-normalise, "can't record NaN".  The literal per-sample restatement below is the checker."""
import numpy as np
import pytest

import dusp_b200 as d
from dusp_b200 import components as C
from dusp_b200.stream import RenderStream


def literal_on_tick(chunk, n_ch, nf):
    """RenderStream.js:26-61 word for word: returns (buffer, normaliseFactor)."""
    buf = np.zeros(n_ch * 256, dtype=np.float32)
    for c in range(n_ch):
        for t in range(256):
            val = float(chunk[c][t]) * nf
            if abs(val) > 1:
                sf = abs(1 / val)
                val *= sf
                nf *= sf
            if val != val:
                raise d.DuspError("can't record NaN")
            buf[t * n_ch + c] = val
    return buf, nf


def bare_stream(n_ch):
    s = RenderStream.__new__(RenderStream)
    s.numberOfChannels, s.normaliseFactor = n_ch, 1.0
    return s


def test_normaliser_matches_the_literal_scan():
    rng = np.random.default_rng(3)
    s, nf = bare_stream(2), 1.0
    for k in range(6):
        chunk = (rng.standard_normal((2, 256)) * (0.2 + 0.5 * k)).astype(np.float32)  # louder and louder: trips repeatedly
        want, nf = literal_on_tick(chunk, 2, nf)
        got = s._normalise_chunk(chunk)
        assert np.array_equal(got, want) and s.normaliseFactor == nf
    assert nf < 1.0
    quiet = (rng.standard_normal((2, 256)) * 0.01).astype(np.float32)
    want, nf2 = literal_on_tick(quiet, 2, nf)
    assert np.array_equal(s._normalise_chunk(quiet), want) and nf2 == nf  # pure-scaling fast path, factor unchanged


def test_nan_is_refused_and_arguments_are_checked():
    s = bare_stream(1)
    bad = np.zeros((1, 256), dtype=np.float32)
    bad[0, 17] = np.nan
    with pytest.raises(d.DuspError, match="can't record NaN"):
        s._normalise_chunk(bad)
    with pytest.raises(d.DuspError, match="requires an outlet"):
        RenderStream(None)
    with pytest.raises(d.DuspError, match="expects an outlet"):
        RenderStream(42)


@pytest.mark.gpu
def test_stream_chunks_equal_reference_semantics(lib, oracle):
    """40 pushed chunks of a mono circuit that clips (two oscillators summed) and of a stereo one, against the oracle's
    raw chunks run through the literal onTick restatement; events inside the stream are honoured."""
    from dusp_b200.render import circuit_blob
    for make, n_ch in ((lambda: C.Sum(C.Osc(200), C.Osc(250.5)), 1),
                       (lambda: C.ConcatChannels(C.Sum(C.Osc(100), C.Osc(301)), C.Osc(2)), 2),
                       (lambda: d.unDusp("[Osc f:([Ramp 4410 900 100] at 0.1)]"), 1)):
        rs = RenderStream(make(), n_ch, block_chunks=16)
        got = [rs.read() for _ in range(40)]
        rs.stop()
        assert rs.read() is None
        blob, _ = circuit_blob(make())
        raw = oracle.render(blob, 40 * 256 / 44100 + 1e-9, raw=True)
        nf = 1.0
        for k in range(40):
            want, nf = literal_on_tick([ch[k * 256:(k + 1) * 256] for ch in raw], n_ch, nf)
            assert got[k].dtype == np.float32 and np.array_equal(got[k], want), "chunk %d" % k
        assert rs.normaliseFactor == nf
        assert rs.format == {"channels": n_ch, "bitDepth": 32, "sampleRate": 44100, "endianness": "LE"}
