"""The streaming sink: mirror of ``src/RenderStream.js:6-80`` on top of the GPU render path.

The reference's RenderStream is a Node Readable in object mode: every tick of the outlet pushes one
``Float32Array(numberOfChannels * 256)`` of interleaved frames ``buffer[t*C + c]``, after a running
peak auto-normalise (``normaliseFactor``), and throws the string "can't record NaN" on a NaN sample.
Here the circuit is rendered on the device in blocks of many chunks (``DUSP_OUT_VOICES_RAW``: the chunk
exactly as the outlet holds it) and the sink's own logic -- the normaliser's sequential scan, the
interleave, the NaN check -- runs on the host per 256-sample chunk: the sink is PCIe/host-bound by
construction.  ``read()`` / iteration return the reference's chunks one by one.
"""
import numpy as np

from . import capi, config
from .graph import DuspError, Outlet, UnitOrPatch
from .render import circuit_blob, SessionRun

CHUNK = config.standardChunkSize if hasattr(config, "standardChunkSize") else 256


class RenderStream:
    def __init__(self, outlet, numberOfChannels=1, device=0, block_chunks=64):
        if outlet is None or outlet is False:
            raise DuspError("RenderStream requires an outlet argument")  # RenderStream.js:9-10
        if isinstance(outlet, UnitOrPatch):
            outlet = outlet.defaultOutlet
        if not isinstance(outlet, Outlet):
            raise DuspError("RenderStream expects an outlet")            # :14-15
        self.numberOfChannels = int(numberOfChannels)
        self.outlet = outlet
        self.sampleRate = outlet.sampleRate
        self.normaliseFactor = 1.0
        self.tickClock = 0
        self.format = {"channels": self.numberOfChannels, "bitDepth": 32, "sampleRate": self.sampleRate,
                       "endianness": "LE"}                               # :63-68
        self.blob, self.circuit = circuit_blob(outlet)
        self.run = SessionRun([self.blob], [1], device)
        self.program = self.run.programs[0]
        if self.program.num_channels < self.numberOfChannels:
            self.run.close()
            # the reference dereferences channelData[c] === undefined here (TypeError in onTick, :38)
            raise DuspError("RenderStream: the outlet has %d channel(s), %d requested"
                            % (self.program.num_channels, self.numberOfChannels))
        self.block = int(block_chunks) * CHUNK
        self._pending = []   # chunks of the current block, ready to hand out
        self._pos = 0        # samples rendered so far
        self._stopped = False

    # -- the reference's onTick body (RenderStream.js:26-61) for one 256-sample chunk [C][256] -> [256*C]
    def _normalise_chunk(self, x):
        C, nf = self.numberOfChannels, self.normaliseFactor
        x = x[:C].astype(np.float64)
        if np.isnan(x).any():
            # exact position semantics: the reference throws at the first NaN in (c, t) order, after
            # normalising what precedes it; the chunk is lost either way
            raise DuspError("can't record NaN")
        buf = np.empty(CHUNK * C, dtype=np.float32)
        if float(np.max(np.abs(x))) * nf <= 1.0:           # no sample can trip the normaliser: pure scaling
            buf.reshape(CHUNK, C)[:] = (x * nf).T
            return buf
        for c in range(C):                                   # the literal scan, channel-major like the reference
            for t in range(CHUNK):
                val = x[c, t] * nf
                if abs(val) > 1:
                    sf = abs(1 / val)
                    val *= sf
                    nf *= sf
                buf[t * C + c] = val
        self.normaliseFactor = nf
        return buf

    def _render_block(self):
        out = self.run.render(self.block, capi.OUT_VOICES_RAW)[0]  # events and graph edits at their chunks
        self._pos += self.block
        for i in range(0, self.block, CHUNK):
            self._pending.append(out[:, i:i + CHUNK])

    def read(self):
        """The next pushed object: a float32 array of 256 interleaved frames, or None after stop()."""
        if self._stopped:
            return None
        if not self._pending:
            self._render_block()
        self.tickClock += CHUNK
        return self._normalise_chunk(self._pending.pop(0))

    def __iter__(self):
        return self

    def __next__(self):
        b = self.read()
        if b is None:
            raise StopIteration
        return b

    def stop(self):  # RenderStream.js:76-79
        self._stopped = True
        self.run.close()
