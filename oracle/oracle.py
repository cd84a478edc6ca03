"""ctypes wrapper of oracle/liboracle.so -- TEST INFRASTRUCTURE ONLY.

Importable only from tests/, __graft_entry__.smoke() and bench.py's CPU
baseline legs.  The product package (dusp_b200/) never imports this module.
"""
import ctypes
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None


def build(force=False):
    so = os.path.join(_HERE, "liboracle.so")
    src = os.path.join(_HERE, "dusp_oracle.c")
    if force or not os.path.exists(so) or os.path.getmtime(so) < os.path.getmtime(src):
        subprocess.check_call(["make", "-C", _HERE, "-s", "-B", "liboracle.so"])
    return so


def lib():
    global _LIB
    if _LIB is None:
        L = ctypes.CDLL(build())
        L.oracle_load.restype = ctypes.c_void_p
        L.oracle_load.argtypes = [ctypes.c_char_p, ctypes.c_char_p, ctypes.c_int]
        L.oracle_free.argtypes = [ctypes.c_void_p]
        L.oracle_schedule.argtypes = [ctypes.c_void_p, ctypes.POINTER(ctypes.c_int)]
        L.oracle_render.argtypes = [ctypes.c_void_p, ctypes.c_double, ctypes.POINTER(ctypes.POINTER(ctypes.c_float)),
                                    ctypes.c_int, ctypes.POINTER(ctypes.c_int), ctypes.POINTER(ctypes.c_long)]
        L.oracle_free_output.argtypes = [ctypes.POINTER(ctypes.POINTER(ctypes.c_float)), ctypes.c_int]
        L.oracle_error.restype = ctypes.c_char_p
        L.oracle_error.argtypes = [ctypes.c_void_p]
        L.oracle_num_units.argtypes = [ctypes.c_void_p]
        L.oracle_rand.restype = ctypes.c_double
        L.oracle_rand.argtypes = [ctypes.c_uint64, ctypes.c_uint64]
        L.oracle_wave_table.restype = ctypes.POINTER(ctypes.c_float)
        L.oracle_wave_table.argtypes = [ctypes.c_char_p]
        L.oracle_shape_table.restype = ctypes.POINTER(ctypes.c_float)
        L.oracle_shape_table.argtypes = [ctypes.c_char_p]
        L.oracle_fft.argtypes = [ctypes.c_long, ctypes.c_int, ctypes.POINTER(ctypes.c_double),
                                 ctypes.POINTER(ctypes.c_double), ctypes.c_int]
        _LIB = L
    return _LIB


class OracleError(Exception):
    pass


def _load(blob):
    err = ctypes.create_string_buffer(256)
    h = lib().oracle_load(blob.encode(), err, 256)
    if not h:
        raise OracleError(err.value.decode())
    return h


def schedule(blob):
    """Tick order of `new Circuit(rootUnit)` per the oracle's own scheduler."""
    h = _load(blob)
    try:
        n = lib().oracle_num_units(h)
        order = (ctypes.c_int * (n + 1))()
        k = lib().oracle_schedule(h, order)
        return list(order[:k])
    finally:
        lib().oracle_free(h)


def render(blob, seconds, raw=False):
    """renderChannelData on the CPU: list of float32 arrays (one per channel).  raw=True: the outlet's chunks as
    they are (no `x || 0`), what the streaming sink reads (RenderStream.js:38)."""
    h = _load(blob)
    lib().oracle_set_raw(1 if raw else 0)
    try:
        outs = (ctypes.POINTER(ctypes.c_float) * 16)()
        nch = ctypes.c_int(0)
        ns = ctypes.c_long(0)
        rc = lib().oracle_render(h, float(seconds), outs, 16, ctypes.byref(nch), ctypes.byref(ns))
        if rc != 0:
            raise OracleError(lib().oracle_error(h).decode())
        res = [np.ctypeslib.as_array(outs[c], shape=(max(ns.value, 1),))[:ns.value].copy() for c in range(nch.value)]
        lib().oracle_free_output(outs, nch.value)
        return res
    finally:
        lib().oracle_set_raw(0)
        lib().oracle_free(h)


def wave_table(name):
    p = lib().oracle_wave_table(name.encode())
    return np.ctypeslib.as_array(p, shape=(44101,)).copy() if p else None


def shape_table(name):
    p = lib().oracle_shape_table(name.encode())
    return np.ctypeslib.as_array(p, shape=(44101,)).copy() if p else None


def rand(seed, k):
    return lib().oracle_rand(seed, k)


def fft(x, mode="real", r32=False):
    """The oracle's restatement of fft.js on one array.  mode "real": realTransform + completeSpectrum of N reals;
    "complex" / "inverse": transform / inverseTransform of N interleaved complex values.  Returns 2N doubles."""
    x = np.ascontiguousarray(x, dtype=np.float64)
    n = len(x) if mode == "real" else len(x) // 2
    out = np.zeros(2 * n, dtype=np.float64)
    dp = ctypes.POINTER(ctypes.c_double)
    rc = lib().oracle_fft(n, {"real": 0, "complex": 1, "inverse": 2}[mode], x.ctypes.data_as(dp), out.ctypes.data_as(dp), int(r32))
    if rc != 0:
        raise OracleError("FFT size must be a power of two and bigger than 1")
    return out
