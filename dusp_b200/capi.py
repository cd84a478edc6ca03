"""ctypes binding of ``libdusp_b200.so`` (C ABI: ``include/dusp_b200.h``).

The library is built in-tree by ``__graft_entry__.build()`` /
``make -C dusp_b200/csrc``.  There is no fallback: if the shared object is
missing ``load()`` raises, and without a CUDA device every session call fails.
"""
import ctypes
import os

from . import config
from .graph import DuspError

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libdusp_b200.so")
_lib = None

OUT_VOICES = 0
OUT_MIX = 1
OUT_VOICES_RAW = 2  # per-voice channel data exactly as the outlet's chunk holds it (NaN / -0 kept)
E_UNSUPPORTED = -2

# every symbol include/dusp_b200.h declares (tests check the .so exports them all)
SYMBOLS = [
    "dusp_abi_version", "dusp_last_error", "dusp_program_create", "dusp_program_destroy",
    "dusp_program_num_channels", "dusp_program_num_units", "dusp_program_num_instructions",
    "dusp_program_num_slots", "dusp_program_bytes_per_voice_sample", "dusp_program_num_params",
    "dusp_program_param_name", "dusp_program_param_index", "dusp_program_param_default",
    "dusp_program_disassemble", "dusp_session_create", "dusp_session_destroy", "dusp_session_num_voices",
    "dusp_session_num_channels", "dusp_session_unit_count", "dusp_session_set_params", "dusp_session_reset",
    "dusp_session_render", "dusp_session_launch_count", "dusp_session_launch_info", "dusp_rand",
    "dusp_session_tile_info", "dusp_session_warning", "dusp_program_specialise", "dusp_session_poke",
    "dusp_session_adopt", "dusp_program_warmup_ticks",
    "dusp_mix_acc_create", "dusp_mix_acc_open", "dusp_mix_acc_close", "dusp_mix_acc_clear", "dusp_mix_push", "dusp_mix_wait",
    "dusp_mix_push_static",
    "dusp_session_clock", "dusp_session_state_size", "dusp_session_state_export", "dusp_session_state_import", "dusp_run_create", "dusp_run_destroy", "dusp_run_num_voices", "dusp_run_num_channels",
    "dusp_run_channels_rendered", "dusp_run_position", "dusp_run_session", "dusp_run_program", "dusp_run_num_epochs",
    "dusp_run_epoch_program", "dusp_run_set_params", "dusp_run_render",
]


class UnsupportedCircuit(DuspError):
    """The circuit uses a unit or configuration outside the accelerated path."""


def load():
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError("libdusp_b200.so is not built (run `python -c 'import __graft_entry__ as g; g.build()'` "
                           "or `make -C dusp_b200/csrc`); there is no CPU fallback")
    L = ctypes.CDLL(LIB_PATH)
    c = ctypes
    L.dusp_last_error.restype = c.c_char_p
    L.dusp_program_create.argtypes = [c.c_char_p, c.c_size_t, c.POINTER(c.c_void_p)]
    L.dusp_program_destroy.argtypes = [c.c_void_p]
    L.dusp_program_destroy.restype = None
    for nm in ("num_channels", "num_units", "num_instructions", "num_slots", "num_params", "warmup_ticks"):
        getattr(L, "dusp_program_" + nm).argtypes = [c.c_void_p]
    L.dusp_program_bytes_per_voice_sample.argtypes = [c.c_void_p, c.c_int]
    L.dusp_program_bytes_per_voice_sample.restype = c.c_double
    L.dusp_program_param_name.argtypes = [c.c_void_p, c.c_int]
    L.dusp_program_param_name.restype = c.c_char_p
    L.dusp_program_param_index.argtypes = [c.c_void_p, c.c_char_p]
    L.dusp_program_param_default.argtypes = [c.c_void_p, c.c_int]
    L.dusp_program_param_default.restype = c.c_double
    L.dusp_program_disassemble.argtypes = [c.c_void_p, c.c_char_p, c.c_size_t]
    L.dusp_program_disassemble.restype = c.c_size_t
    L.dusp_session_create.argtypes = [c.POINTER(c.c_void_p), c.POINTER(c.c_int), c.c_int, c.c_int, c.POINTER(c.c_void_p)]
    L.dusp_session_destroy.argtypes = [c.c_void_p]
    L.dusp_session_destroy.restype = None
    L.dusp_session_num_voices.argtypes = [c.c_void_p]
    L.dusp_session_num_channels.argtypes = [c.c_void_p]
    L.dusp_session_unit_count.argtypes = [c.c_void_p]
    L.dusp_session_unit_count.restype = c.c_longlong
    L.dusp_session_set_params.argtypes = [c.c_void_p, c.c_int, c.POINTER(c.c_int), c.c_int, c.POINTER(c.c_double)]
    L.dusp_session_reset.argtypes = [c.c_void_p]
    L.dusp_session_poke.argtypes = [c.c_void_p, c.c_int, c.c_int, c.c_int, c.POINTER(c.c_int), c.POINTER(c.c_double), c.c_void_p]
    L.dusp_session_adopt.argtypes = [c.c_void_p, c.c_int, c.c_void_p, c.c_int, c.c_void_p]
    L.dusp_session_render.argtypes = [c.c_void_p, c.c_longlong, c.c_int, c.c_void_p, c.c_void_p, c.c_void_p]
    L.dusp_session_launch_count.argtypes = [c.c_void_p]
    L.dusp_session_launch_count.restype = c.c_longlong
    L.dusp_session_launch_info.argtypes = [c.c_void_p, c.POINTER(c.c_int), c.POINTER(c.c_int), c.POINTER(c.c_int)]
    L.dusp_session_tile_info.argtypes = [c.c_void_p, c.POINTER(c.c_int), c.POINTER(c.c_int), c.POINTER(c.c_int)]
    L.dusp_session_warning.argtypes = [c.c_void_p]
    L.dusp_session_warning.restype = c.c_char_p
    L.dusp_program_specialise.argtypes = [c.c_void_p, c.c_int, c.c_int, c.c_char_p, c.c_size_t]
    L.dusp_program_specialise.restype = c.c_size_t
    L.dusp_rand.argtypes = [c.c_uint64, c.c_uint64]
    L.dusp_rand.restype = c.c_double
    L.dusp_mix_acc_create.argtypes = [c.c_int, c.c_size_t, c.c_size_t, c.POINTER(c.c_void_p), c.POINTER(c.c_void_p), c.c_char_p]
    L.dusp_mix_acc_open.argtypes = [c.c_int, c.c_char_p, c.c_size_t, c.POINTER(c.c_void_p), c.POINTER(c.c_void_p)]
    L.dusp_mix_acc_close.argtypes = [c.c_void_p, c.c_int]
    L.dusp_mix_acc_clear.argtypes = [c.c_void_p, c.c_size_t, c.c_size_t, c.c_void_p]
    L.dusp_mix_push.argtypes = [c.c_void_p, c.c_void_p, c.c_size_t, c.c_void_p, c.c_void_p, c.c_void_p]
    L.dusp_mix_wait.argtypes = [c.c_void_p, c.c_uint, c.c_void_p, c.c_void_p]
    L.dusp_mix_push_static.argtypes = [c.c_int, c.c_void_p, c.c_void_p, c.c_size_t, c.c_void_p, c.c_void_p]
    L.dusp_session_clock.argtypes = [c.c_void_p]
    L.dusp_session_clock.restype = c.c_longlong
    L.dusp_session_state_size.argtypes = [c.c_void_p, c.POINTER(c.c_size_t)]
    L.dusp_session_state_export.argtypes = [c.c_void_p, c.c_void_p, c.c_size_t, c.c_void_p]
    L.dusp_session_state_import.argtypes = [c.c_void_p, c.c_void_p, c.c_size_t, c.c_void_p]
    L.dusp_run_create.argtypes = [c.POINTER(c.c_char_p), c.POINTER(c.c_size_t), c.POINTER(c.c_int), c.c_int, c.c_int, c.POINTER(c.c_void_p)]
    L.dusp_run_destroy.argtypes = [c.c_void_p]
    L.dusp_run_destroy.restype = None
    for nm in ("num_voices", "num_channels", "channels_rendered"):
        getattr(L, "dusp_run_" + nm).argtypes = [c.c_void_p]
    L.dusp_run_position.argtypes = [c.c_void_p]
    L.dusp_run_position.restype = c.c_longlong
    L.dusp_run_session.argtypes = [c.c_void_p]
    L.dusp_run_session.restype = c.c_void_p
    L.dusp_run_program.argtypes = [c.c_void_p, c.c_int]
    L.dusp_run_program.restype = c.c_void_p
    L.dusp_run_num_epochs.argtypes = [c.c_void_p, c.c_int]
    L.dusp_run_epoch_program.argtypes = [c.c_void_p, c.c_int, c.c_int, c.POINTER(c.c_longlong)]
    L.dusp_run_epoch_program.restype = c.c_void_p
    L.dusp_run_set_params.argtypes = [c.c_void_p, c.c_int, c.POINTER(c.c_char_p), c.c_int, c.POINTER(c.c_double)]
    L.dusp_run_render.argtypes = [c.c_void_p, c.c_longlong, c.c_int, c.c_void_p, c.c_void_p, c.c_void_p]
    if L.dusp_abi_version() != 1:
        raise RuntimeError("libdusp_b200.so ABI version mismatch")
    assert config.sampleRate == 44100 and config.standardChunkSize == 256  # compiled into the kernels
    _lib = L
    return L


def check(rc):
    if rc == 0:
        return
    msg = load().dusp_last_error().decode()
    if rc == E_UNSUPPORTED:
        raise UnsupportedCircuit(msg)
    raise DuspError(msg)


class Program:
    """A circuit compiled to a flat op program (``dusp_program``)."""

    def __init__(self, blob, _borrowed=None):
        L = load()
        self.blob = blob
        self._owned = _borrowed is None
        if self._owned:
            h = ctypes.c_void_p()
            data = blob.encode()
            check(L.dusp_program_create(data, len(data), ctypes.byref(h)))
        else:
            h = ctypes.c_void_p(_borrowed)  # owned by a dusp_run
        self._h = h
        self.num_channels = L.dusp_program_num_channels(h)
        self.num_units = L.dusp_program_num_units(h)
        self.num_instructions = L.dusp_program_num_instructions(h)
        self.num_slots = L.dusp_program_num_slots(h)
        self.warmup_ticks = L.dusp_program_warmup_ticks(h)

    def __del__(self):
        if getattr(self, "_h", None) and _lib is not None and getattr(self, "_owned", False):
            _lib.dusp_program_destroy(self._h)
            self._h = None

    def params(self):
        L = load()
        n = L.dusp_program_num_params(self._h)
        return [(L.dusp_program_param_name(self._h, i).decode(), L.dusp_program_param_default(self._h, i))
                for i in range(n)]

    def param_index(self, name):
        i = load().dusp_program_param_index(self._h, name.encode())
        if i < 0:
            raise DuspError("no such parameter: " + name)
        return i

    def bytes_per_voice_sample(self, out_mode=OUT_VOICES):
        return load().dusp_program_bytes_per_voice_sample(self._h, out_mode)

    def specialise(self, sub_block=8, time_lanes=4):
        """(cubin size, generated CUDA source) of the kernel specialised for this circuit shape."""
        L = load()
        buf = ctypes.create_string_buffer(1 << 20)
        n = L.dusp_program_specialise(self._h, sub_block, time_lanes, buf, len(buf))
        if not n:
            raise DuspError(L.dusp_last_error().decode())
        return n, buf.value.decode()

    def disassemble(self):
        L = load()
        n = L.dusp_program_disassemble(self._h, None, 0)
        buf = ctypes.create_string_buffer(n)
        L.dusp_program_disassemble(self._h, buf, n)
        return buf.value.decode()


class Session:
    """n voices of one or more programs resident on one GPU (``dusp_session``)."""

    def __init__(self, programs, voices, device=0, _borrowed=None):
        L = load()
        self.programs = list(programs)
        self._owned = _borrowed is None
        if self._owned:
            n = len(self.programs)
            ph = (ctypes.c_void_p * n)(*[p._h for p in self.programs])
            vc = (ctypes.c_int * n)(*[int(v) for v in voices])
            h = ctypes.c_void_p()
            check(L.dusp_session_create(ph, vc, n, int(device), ctypes.byref(h)))
        else:
            h = ctypes.c_void_p(_borrowed)  # owned by a dusp_run
        self._h = h
        self.device = device
        self.num_voices = L.dusp_session_num_voices(h)
        self.num_channels = L.dusp_session_num_channels(h)
        self.unit_count = L.dusp_session_unit_count(h)

    def close(self):
        if getattr(self, "_h", None) and _lib is not None and getattr(self, "_owned", False):
            _lib.dusp_session_destroy(self._h)
        self._h = None

    __del__ = close

    @property
    def clock(self):
        return load().dusp_session_clock(self._h)

    def set_params(self, program_index, names, values):
        """values: array [n_voices_of_program][len(names)] of float64."""
        import numpy as np
        L = load()
        prog = self.programs[program_index]
        ids = (ctypes.c_int * len(names))(*[prog.param_index(n) if isinstance(n, str) else int(n) for n in names])
        vals = np.ascontiguousarray(values, dtype=np.float64)
        check(L.dusp_session_set_params(self._h, program_index, ids, len(names),
                                        vals.ctypes.data_as(ctypes.POINTER(ctypes.c_double))))

    def export_state(self, stream=None):
        """The session's state block (clock, words, rings, stale chunks) as a numpy uint8 array: checkpoint."""
        import numpy as np
        L = load()
        n = ctypes.c_size_t(0)
        check(L.dusp_session_state_size(self._h, ctypes.byref(n)))
        buf = np.empty(n.value, dtype=np.uint8)
        check(L.dusp_session_state_export(self._h, buf.ctypes.data_as(ctypes.c_void_p), n.value, stream))
        return buf

    def import_state(self, block, stream=None):
        """Resume from a block exported by a session of the same programs and voice counts."""
        import numpy as np
        block = np.ascontiguousarray(block, dtype=np.uint8)
        check(load().dusp_session_state_import(self._h, block.ctypes.data_as(ctypes.c_void_p), block.size, stream))

    def reset(self):
        check(load().dusp_session_reset(self._h))

    def poke(self, program_index, voice, names, values):
        """Assign unit state / parameter words on the device between two renders (events):
        ``voice`` = index within the program group, or -1 for every voice of the group."""
        L = load()
        prog = self.programs[program_index]
        if any(isinstance(n, str) and n.endswith("[*]") for n in names):  # every channel the program holds
            all_names = [nm for nm, _ in prog.params()]
            ex_n, ex_v = [], []
            for n, v in zip(names, values):
                if isinstance(n, str) and n.endswith("[*]"):
                    hits = [a for a in all_names if a.startswith(n[:-2])]
                    ex_n += hits
                    ex_v += [v] * len(hits)
                else:
                    ex_n.append(n)
                    ex_v.append(v)
            names, values = ex_n, ex_v
            if not names:
                return
        try:
            ids = (ctypes.c_int * len(names))(*[prog.param_index(n) if isinstance(n, str) else int(n) for n in names])
        except DuspError as e:
            raise UnsupportedCircuit("event assigns state the compiled program does not hold: " + str(e))
        vals = (ctypes.c_double * len(names))(*[float(v) for v in values])
        check(L.dusp_session_poke(self._h, program_index, int(voice), len(names), ids, vals, None))

    def adopt(self, program_index, src, src_program_index):
        """Take over the device state of a program group of session ``src`` (a graph edit: this session's
        program is the circuit after the edit; units are matched by label).  SURVEY 8f-1."""
        check(load().dusp_session_adopt(self._h, int(program_index), src._h, int(src_program_index), None))

    def render_host(self, n_samples, out_mode=OUT_VOICES):
        """Render into a fresh host array: [voices][channels][n] or [2][n] (mix)."""
        import numpy as np
        shape = (self.num_voices, self.num_channels, n_samples) if out_mode != OUT_MIX else (2, n_samples)
        out = np.empty(shape, dtype=np.float32)
        check(load().dusp_session_render(self._h, int(n_samples), out_mode, out.ctypes.data_as(ctypes.c_void_p), None, None))
        return out

    def render_into_host(self, out, n_samples, out_mode=OUT_VOICES):
        check(load().dusp_session_render(self._h, int(n_samples), out_mode, ctypes.c_void_p(out), None, None))

    def render_device(self, n_samples, device_ptr, out_mode=OUT_VOICES, stream=None):
        """Enqueue a render into caller-owned device memory (no synchronisation)."""
        check(load().dusp_session_render(self._h, int(n_samples), out_mode, None, ctypes.c_void_p(device_ptr),
                                         ctypes.c_void_p(stream) if stream else None))

    @property
    def launch_count(self):
        return load().dusp_session_launch_count(self._h)

    def launch_info(self):
        g, b, s = ctypes.c_int(), ctypes.c_int(), ctypes.c_int()
        load().dusp_session_launch_info(self._h, ctypes.byref(g), ctypes.byref(b), ctypes.byref(s))
        B, T, f = ctypes.c_int(), ctypes.c_int(), ctypes.c_int()
        load().dusp_session_tile_info(self._h, ctypes.byref(B), ctypes.byref(T), ctypes.byref(f))
        return {"grid": g.value, "block": b.value, "smem_bytes": s.value, "sub_block": B.value, "time_lanes": T.value,
                "specialised": bool(f.value), "warning": load().dusp_session_warning(self._h).decode()}


class Run:
    """A render driven through its circuits' timelines (``dusp_run``, csrc/run.cc): the blobs' EV / EPOCH / KEEP /
    WARMUP records become chunk-aligned segments, pokes and adopted sessions inside the library."""

    def __init__(self, blobs, voices, device=0):
        L = load()
        n = len(blobs)
        data = [b.encode() for b in blobs]
        arr = (ctypes.c_char_p * n)(*data)
        lens = (ctypes.c_size_t * n)(*[len(x) for x in data])
        vc = (ctypes.c_int * n)(*[int(v) for v in voices])
        h = ctypes.c_void_p()
        check(L.dusp_run_create(arr, lens, vc, n, int(device), ctypes.byref(h)))
        self._h = h
        self.blobs, self.voices, self.device = list(blobs), [int(v) for v in voices], device
        self.num_voices = L.dusp_run_num_voices(h)
        self.num_channels = L.dusp_run_num_channels(h)

    def close(self):
        if getattr(self, "_h", None) and _lib is not None:
            _lib.dusp_run_destroy(self._h)
        self._h = None

    __del__ = close

    @property
    def pos(self):
        return load().dusp_run_position(self._h)

    @property
    def channels_rendered(self):
        return load().dusp_run_channels_rendered(self._h)

    @property
    def session(self):
        """The session rendering right now (borrowed: replaced when an epoch begins; created on first use)."""
        h = load().dusp_run_session(self._h)
        if not h:
            msg = load().dusp_last_error().decode()
            raise (UnsupportedCircuit if "unsupported" in msg.lower() else DuspError)(msg)
        return Session(self.programs, self.voices, self.device, _borrowed=h)

    @property
    def programs(self):
        """The programs rendering right now, one per group (borrowed)."""
        L = load()
        return [Program(None, _borrowed=L.dusp_run_program(self._h, i)) for i in range(len(self.blobs))]

    def epochs(self, program_index=0):
        """[(first chunk, Program)] of a group's timeline: warm-up ticks, then one program per EPOCH section."""
        L = load()
        out = []
        for e in range(L.dusp_run_num_epochs(self._h, program_index)):
            k0 = ctypes.c_longlong()
            h = L.dusp_run_epoch_program(self._h, program_index, e, ctypes.byref(k0))
            out.append((k0.value, Program(None, _borrowed=h)))
        return out

    def set_params(self, program_index, names, values):
        """values: array [n_voices_of_program][len(names)] of float64 (before the first render)."""
        import numpy as np
        nm = (ctypes.c_char_p * len(names))(*[n.encode() for n in names])
        vals = np.ascontiguousarray(values, dtype=np.float64)
        check(load().dusp_run_set_params(self._h, int(program_index), nm, len(names),
                                         vals.ctypes.data_as(ctypes.POINTER(ctypes.c_double))))

    def render_host(self, n_samples, out_mode=OUT_VOICES):
        """The next n samples: [voices][channels][n] (channels = what the reference's result would hold so far) or
        [2][n] (mix)."""
        import numpy as np
        n_samples = int(n_samples)
        shape = (self.num_voices, self.num_channels, n_samples) if out_mode != OUT_MIX else (2, n_samples)
        out = np.empty(shape, dtype=np.float32)
        check(load().dusp_run_render(self._h, n_samples, out_mode, out.ctypes.data_as(ctypes.c_void_p), None, None))
        if out_mode != OUT_MIX and n_samples and self.channels_rendered < self.num_channels:
            out = out[:, :self.channels_rendered]
        return out

    def render_device(self, n_samples, device_ptr, out_mode=OUT_VOICES, stream=None):
        check(load().dusp_run_render(self._h, int(n_samples), out_mode, None, ctypes.c_void_p(device_ptr),
                                     ctypes.c_void_p(stream) if stream else None))


class MixAccumulator:
    """The root GPU's mixdown accumulator shared through CUDA IPC (csrc/mixpush.cu): ``n_floats`` f32 sums and
    ``n_counters`` u32 arrival counters.  The root creates it (``handle`` is the 64 bytes to send to the other
    processes); peers open it from the handle.  Pointers are raw device addresses valid in this process."""

    def __init__(self, device, n_floats, n_counters, handle=None):
        L = load()
        self.n_floats, self.n_counters, self.owner = int(n_floats), int(n_counters), handle is None
        acc, ctr = ctypes.c_void_p(), ctypes.c_void_p()
        if self.owner:
            buf = ctypes.create_string_buffer(64)
            check(L.dusp_mix_acc_create(device, self.n_floats, self.n_counters, ctypes.byref(acc), ctypes.byref(ctr), buf))
            self.handle = buf.raw
        else:
            self.handle = bytes(handle)
            check(L.dusp_mix_acc_open(device, self.handle, self.n_floats, ctypes.byref(acc), ctypes.byref(ctr)))
        self.acc, self.counters = acc.value, ctr.value

    def clear(self, stream=None):
        check(load().dusp_mix_acc_clear(self.acc, self.n_floats, self.n_counters, stream))

    def push(self, local_ptr, offset_floats, n_floats, counter_index, ticket_ptr, stream=None):
        """acc[offset .. offset+n) += local[0..n) (system-scope reductions over NVLink), then counter[index] += 1."""
        check(load().dusp_mix_push(local_ptr, self.acc + 4 * int(offset_floats), int(n_floats),
                                   self.counters + 4 * int(counter_index), ticket_ptr, stream))

    def wait(self, counter_index, target, stream=None):
        """Holds ``stream`` until counter[index] >= target; a wait that gives up (~30 s) bumps the spare counter
        behind the last one, which ``timed_out_ptr`` addresses (the host reads it after the render)."""
        check(load().dusp_mix_wait(self.counters + 4 * int(counter_index), int(target), self.timed_out_ptr, stream))

    @property
    def timed_out_ptr(self):
        return self.counters + 4 * self.n_counters

    def close(self):
        if self.acc:
            load().dusp_mix_acc_close(self.acc, 1 if self.owner else 0)
            self.acc = None
