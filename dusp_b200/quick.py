"""Constant-folding operator builders -- mirror of ``src/quick.js:15-110``."""
import math

from . import components as C
from .graph import UnitOrPatch, Outlet


def _num(x):
    return isinstance(x, (int, float)) and not isinstance(x, bool)


def _sig(x):
    return isinstance(x, (UnitOrPatch, Outlet))


def add(a, b):  # quick.js:15-20
    if _num(a) and _num(b):
        return a + b
    return C.Sum(a, b)


def subtract(a, b):  # :22-27
    if _num(a) and _num(b):
        return a - b
    return C.Subtract(a, b)


def mult(a, b):  # :29-39
    if a is None or (_num(a) and a == 1):
        return b
    if b is None or (_num(b) and b == 1):
        return a
    if _num(a) and _num(b):
        return a * b
    return C.Multiply(a, b)


multiply = mult


def divide(a, b):  # :42-47
    if _num(a) and _num(b):
        return a / b if b != 0 else (math.nan if a == 0 or a != a else math.copysign(math.inf, a) * math.copysign(1, b))
    return C.Divide(a, b)


def invert(a):  # :49-54
    if _num(a):
        return -a
    return C.PolarityInvert(a)


def semitoneToRatio(p):  # :56-61
    if _num(p):
        return math.pow(2, p / 12)
    return C.SemitoneToRatio(p)


def pToF(p):  # :62-67
    if _num(p):
        return math.pow(2, (p - 69) / 12) * 440
    raise Exception("quick.pToF(non number) has not been implemented")


def concat(a, b):  # :69-74
    if _sig(a) or _sig(b):
        return C.ConcatChannels(a, b)
    la = list(a) if isinstance(a, (list, tuple)) else [a]
    lb = list(b) if isinstance(b, (list, tuple)) else [b]
    return la + lb


def pow(a, b):  # :76-81
    if _sig(a) or _sig(b):
        return C.Pow(a, b)
    return math.pow(a, b)


def clipAbove(input, threshold):  # :83-92
    if _sig(input) or _sig(threshold):
        return C.HardClipAbove(input, threshold)
    return threshold if input > threshold else input


def clipBelow(input, threshold):  # :94-103
    if _sig(input) or _sig(threshold):
        return C.HardClipBelow(input, threshold)
    return threshold if input < threshold else input
