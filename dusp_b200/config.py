"""Host-side constants -- mirror of the reference's ``src/config.js:5-17``.

The device kernels are compiled for exactly these values (``DUSP_SR`` /
``DUSP_CHUNK`` in ``csrc/vm.h``); ``capi.load()`` asserts they agree.

``Math.random`` replacement: the reference draws from the global
``Math.random()`` both at graph-construction time (``Osc`` waveform "random",
``AllPass.random`` ...) and inside ``Noise._tick`` (``Noise.js:10,23``).  Per
BASELINE.json's north_star the stochastic units are driven by an injected,
seeded generator.  Two functions: ``rand_tick`` is what replaces ``Math.random``
inside a tick (counter-based, keyed by (unit seed, draw index); the oracle and
the CUDA kernels carry their own restatements and the tests check all three
agree); ``rand53`` (splitmix64) serves the construction-time draws of the host
mirror and the workload generators -- it never runs on the device.
"""
import math

standardChunkSize = 256  # src/config.js:6
sampleRate = 44100       # src/config.js:7
channelFormat = "stereo"
fft = {"windowSize": 4096, "hopSize": 4096 // 4, "windowKind": "hamming"}  # src/config.js:10-14
useDuspShorthands = True
sampleInterval = 1 / sampleRate

_MASK = (1 << 64) - 1
_GOLD = 0x9E3779B97F4A7C15


def _mix64(z):
    z &= _MASK
    z = ((z ^ (z >> 30)) * 0xBF58476D1CE4E5B9) & _MASK
    z = ((z ^ (z >> 27)) * 0x94D049BB133111EB) & _MASK
    return z ^ (z >> 31)


def rand53(seed, k):
    """Uniform double in [0,1): draw ``k`` of the stream ``seed``."""
    return (_mix64(seed + (k + 1) * _GOLD) >> 11) * (1.0 / 9007199254740992.0)


def rand_tick(seed, k):
    """The generator injected for ``Math.random`` INSIDE a tick (Noise.js:10,23, SporadicRetrigger.js:13): draw ``k``
    of the stream ``seed``, uniform in [0,1) with 32 bits of resolution.  Host copy of ``dusp_rand``
    (include/dusp_b200.h) / ``oracle_rand``; the tests check all three agree."""
    m = 0xFFFFFFFF
    x = ((k & m) + (seed & m)) & m
    x ^= x >> 16
    x = (x * 0x7FEB352D) & m
    x ^= x >> 15
    x = (x * 0x846CA68B) & m
    x ^= x >> 16
    x ^= (seed >> 32) & m
    return x * (1.0 / 4294967296.0)


class _HostRandom:
    """Seedable stand-in for ``Math.random`` at construction time and the
    allocator of per-Noise-unit stream seeds."""

    def __init__(self, seed=0):
        self.seed(seed)

    def seed(self, seed):
        self._seed = seed & _MASK
        self._k = 0
        self._noise_ordinal = 0

    def random(self):
        r = rand53(self._seed, self._k)
        self._k += 1
        return r

    def next_noise_seed(self):
        s = _mix64((self._seed ^ 0xD1B54A32D192ED03) + self._noise_ordinal * _GOLD)
        self._noise_ordinal += 1
        return s


host_random = _HostRandom(0)


def seed(s):
    """Seed the injected generator (both construction-time draws and the
    seeds handed to subsequently constructed Noise units)."""
    host_random.seed(s)


def js_or(value, default):
    """JavaScript ``value || default`` for the value kinds constructors see."""
    if value is None or value is False:
        return default
    if isinstance(value, (int, float)) and not isinstance(value, bool):
        if value == 0 or (isinstance(value, float) and math.isnan(value)):
            return default
    if isinstance(value, str) and value == "":
        return default
    return value
