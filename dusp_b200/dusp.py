"""Graph -> Dusp text.  Mirror of ``src/dusp.js:4-120``.

The reference serialiser is reproduced as written, quirks included, because
``FMSynth`` (``patches/FMSynth.js:33-53``) round-trips its seed through it and a
maintainer comparing strings expects the same text:

* ``numberOfOutgoingConnections`` (``Unit.js:144-149``) adds *arrays* to a number,
  so it is the string ``"0"`` for a unit nobody reads and ``"0[object Object]…"``
  otherwise; ``<= 1`` is therefore true only for units with NO outgoing
  connection -- shorthands (``O440``, ``(a * b)``, ``(a + b)``) are used for the
  root of an expression only, every other unit is written ``[Kind #label …]``.
* An unconnected inlet inside ``[...]`` is written by plain concatenation
  (``dusp.js:56-59``): array constants come out as ``0,0`` without parentheses;
  through ``duspInlet`` (``:105-116``, the shorthand path) they are ``(0,0)``.
* Patches are not handled (``dusp.js:92-94``): ``console.warn`` and ``null``.

Per-class hooks (``X.prototype.dusp`` in the reference) live in ``_HOOKS`` below.
"""
import math

from . import config
from .graph import Unit, Outlet, Inlet, Patch, DuspError


def js_number_string(x):
    """ECMAScript Number::toString(10) for a Python int / float."""
    if isinstance(x, bool):
        return "true" if x else "false"
    x = float(x)
    if x != x:
        return "NaN"
    if x == 0:
        return "0"
    if math.isinf(x):
        return "Infinity" if x > 0 else "-Infinity"
    sign = "-" if x < 0 else ""
    r = repr(abs(x))
    mant, _, exp = r.partition("e")
    ip, _, fp = mant.partition(".")
    e10 = int(exp) if exp else 0
    digits = (ip + fp).lstrip("0")
    # value = 0.digits * 10^n  with n = position of the decimal point
    n = len(ip) + e10 - (len(ip + fp) - len((ip + fp).lstrip("0")))
    digits = digits.rstrip("0") or "0"
    k = len(digits)
    if k <= n <= 21:
        return sign + digits + "0" * (n - k)
    if 0 < n <= 21:
        return sign + digits[:n] + "." + digits[n:]
    if -6 < n <= 0:
        return sign + "0." + "0" * (-n) + digits
    e = n - 1
    es = ("+" if e >= 0 else "-") + str(abs(e))
    if k == 1:
        return sign + digits + "e" + es
    return sign + digits[0] + "." + digits[1:] + "e" + es


def _js_concat(v):
    """``"" + v`` for the values an inlet constant can hold."""
    if v is None:
        return "undefined"
    if isinstance(v, (list, tuple)):
        return ",".join("" if e is None else _js_concat(e) for e in v)
    if isinstance(v, str):
        return v
    return js_number_string(v)


def numberOfOutgoingConnections(unit):
    """``Unit.js:144-149`` as JavaScript evaluates it: a string."""
    n = 0
    for outlet in unit.outlets.values():
        s = ",".join("[object Object]" for _ in outlet.connections)
        n = (js_number_string(n) if not isinstance(n, str) else n) + s
    return n


def _js_le_one(n):
    """``n <= 1`` with n the value above (string -> ToNumber)."""
    if isinstance(n, str):
        try:
            return float(n) <= 1
        except ValueError:
            return False  # NaN
    return n <= 1


# ---- per-class hooks: X.prototype.dusp -------------------------------------
def _osc_shorthand(o, index):  # Osc.js:26-32
    if o.waveform == "sin" and not o.F.connected:
        return "O" + _js_concat(o.F.constant)
    return None


def _binary(sym):
    def shorthand(o, index):  # Multiply.js:18-20, Sum.js:13-15
        return "(" + _str(dusp(o.A, index)) + " " + sym + " " + _str(dusp(o.B, index)) + ")"
    return shorthand


_HOOKS = {
    # Osc.js:22-33.  MultiChannelOsc inherits Osc.prototype.
    "Osc": {"extraProperties": {"waveform": "sin"}, "shorthand": _osc_shorthand},
    "Multiply": {"shorthand": _binary("*")},
    "Sum": {"shorthand": _binary("+")},
    # Shape/index.js:61-77
    "Shape": {"extraProperties": ["shape"],
              "extraArgs": lambda o: ["trigger"] if o.playing else []},
    # Repeater.js:15-22
    "Repeater": {"extraArgs": lambda o: ['"' + o.measuredIn + '"'] if o.measuredIn else None},
}


def _hooks(o):
    for cls in type(o).__mro__:
        h = _HOOKS.get(cls.__name__)
        if h is not None:
            return h
    return None


def _str(v):
    if v is None:
        return "null"
    if isinstance(v, str):
        return v
    return _js_concat(v)


def dusp(o, index=None):
    """``dusp(o, index)`` -- dusp.js:4-95.  Returns a str, a number (numbers pass
    through), or None (``null`` / ``undefined`` / unserialisable)."""
    if index is None:
        index = {}
    if o is None:
        return None
    if isinstance(o, bool):
        return None
    if isinstance(o, (int, float)):
        return o
    if isinstance(o, str):
        return '"' + o + '"'  # duspString, dusp.js:118-120
    if getattr(o, "isUnit", False):
        if o.label in index:
            return "#" + o.label
        index[o.label] = o
        use_shorthand = _js_le_one(numberOfOutgoingConnections(o)) if config.useDuspShorthands else False
        h = _hooks(o)
        if use_shorthand and h and h.get("shorthand"):
            s = h["shorthand"](o, index)
            if s:
                return s
        args = [getattr(o, "KIND", None) or type(o).__name__]
        if not use_shorthand:
            args.append("#" + o.label)
        for name, inlet in o.inlets.items():
            if inlet.outlet is not None:
                value = _str(_dusp_outlet(inlet.outlet, index))
            else:
                value = _js_concat(inlet.constant)
            args.append(name.upper() + ":" + value)
        if h:
            ep = h.get("extraProperties")
            if isinstance(ep, list):
                for prop in ep:
                    args.append(prop + ":" + _str(dusp(getattr(o, prop, None))))
            elif isinstance(ep, dict):
                for prop, default in ep.items():
                    if getattr(o, prop, None) != default:
                        args.append(prop + ":" + _str(dusp(getattr(o, prop, None))))
            if h.get("extraArgs"):
                extra = h["extraArgs"](o)
                if extra:
                    args = args + list(extra)
        return "[" + " ".join(args) + "]"
    if isinstance(o, Outlet):
        return _dusp_outlet(o, index)
    if isinstance(o, Inlet):
        return _dusp_inlet(o, index)
    return None  # dusp.js:92-94 (patches land here)


def _dusp_outlet(o, index):  # dusp.js:100-106
    if o is o.unit.defaultOutlet:
        return dusp(o.unit, index)
    return _str(dusp(o.unit, index)) + "." + o.name.upper()


def _dusp_inlet(inlet, index):  # dusp.js:108-118
    if inlet.connected:
        return dusp(inlet.outlet, index)
    c = inlet.constant
    if isinstance(c, (int, float)) and not isinstance(c, bool):
        return c
    if isinstance(c, (list, tuple)):
        return "(" + ",".join(_js_concat(e) for e in c) + ")"
    raise DuspError("strange constant: " + _js_concat(c))


dusp.usingShorthands = config.useDuspShorthands
