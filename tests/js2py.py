"""Test infrastructure: a line-oriented translator for the small JavaScript function bodies of the reference's `_tick`s
(verbatim text read from the reference's source at test time by tests/ref_text.py -- never committed) into
Python, so that the reference's OWN statements can be executed against the oracle without a JavaScript engine.

What makes that sound: JavaScript's Number arithmetic is IEEE binary64, as Python's float arithmetic; `%` on Numbers is
C's fmod; a store into a Float32Array rounds to binary32 (F32 below does), a read widens.  The translator only changes
the surface: `this.` -> `S.`, braces -> indentation, `x++`, `a % b`, `(x || 0)`, `Math.*`, `for(var ...)`.  It covers
the constructs these bodies use and raises on anything else."""
import math
import re

import numpy as np


class F32:
    """A Float32Array: stores round to binary32, reads return Python floats, out-of-range reads are `undefined` (None)."""

    def __init__(self, n_or_values):
        self.a = np.zeros(int(n_or_values), dtype=np.float32) if np.isscalar(n_or_values) else np.array(n_or_values, dtype=np.float32)

    def __len__(self):
        return len(self.a)

    def __getitem__(self, i):
        i = int(i)  # out of range: `undefined`, which every use in these bodies turns into NaN (arithmetic) or 0 (`|| 0`)
        return np.float64(self.a[i]) if 0 <= i < len(self.a) else np.float64("nan")

    def __setitem__(self, i, v):
        i = int(i)
        if 0 <= i < len(self.a):  # a store past the end of a typed array is ignored
            with np.errstate(over="ignore", invalid="ignore"):
                self.a[i] = np.float32(float("nan") if v is None else v)


class JSArray(list):
    """A plain JavaScript array of Numbers: reading past the end gives `undefined`, writing extends."""

    def __getitem__(self, i):
        i = int(i)
        return list.__getitem__(self, i) if 0 <= i < len(self) else None

    def __setitem__(self, i, v):
        i = int(i)
        while len(self) <= i:
            self.append(None)
        list.__setitem__(self, i, v)


class Bag:
    """`this`: attribute bag; `in` is spelled `in_`."""

    def __init__(self, **kw):
        for k, v in kw.items():
            setattr(self, k, v)

    def __setattr__(self, k, v):  # Numbers are IEEE doubles with IEEE division (x / 0 = Infinity, 0 / 0 = NaN): numpy's, not Python's
        if isinstance(v, (int, float)) and not isinstance(v, bool):
            v = np.float64(v)
        object.__setattr__(self, k, v)


def js_or(a, b):  # a || b on Numbers / undefined
    return b if (a is None or a == 0 or a != a) else a


def js_mod(a, b):
    if b == 0 or a != a or b != b or math.isinf(a):
        return np.float64("nan")
    return np.float64(math.fmod(a, b))


def _np(f):
    return lambda x: np.float64(f(x)) if x == x and not math.isinf(x) else np.float64("nan")


def js_pow(a, b):  # Math.pow
    try:
        return np.float64(math.pow(a, b))
    except (ValueError, OverflowError):
        if a < 0 and b != int(b):
            return np.float64("nan")
        return np.float64("inf") if abs(a) > 1 else np.float64(0.0)
    except ZeroDivisionError:
        return np.float64("inf")


ENV = {"M_pow": js_pow, "M_sqrt": lambda x: np.float64(math.sqrt(x)) if x >= 0 else np.float64("nan"),
       "dB": lambda db: js_pow(10, db / 20),   # Gain.js:26-28 / Pan.js:26-28: the files' own helper, Math.pow(10, db/20)
       "js_or": js_or, "js_mod": js_mod, "M_floor": math.floor, "M_ceil": math.ceil, "M_tan": _np(math.tan), "M_cos": _np(math.cos),
       "M_sin": _np(math.sin), "M_PI": np.float64(math.pi), "M_abs": abs, "M_round": lambda x: math.floor(x + 0.5), "F32": F32, "len": len,
       "range": range, "True": True, "False": False, "isinstance": isinstance, "int": int, "float": float,
       "max": max, "Exception": Exception, "JSArray": JSArray, "Infinity": np.float64("inf"), "zeroChannel": F32(256), "zeroChunk": F32(256),
       "samplePeriod": np.float64(1 / 44100), "console": type("Console", (), {"log": staticmethod(lambda *a: None)})}


def _expr(e):
    e = e.strip().rstrip(";")
    e = re.sub(r"\bthis\.in\b", "S.in_", e)
    e = re.sub(r"\bthis\.", "S.", e)
    e = re.sub(r"\bthis\b", "S", e)
    e = re.sub(r"\b(?:Unit|config)\.sampleRate\b", "44100", e)
    e = re.sub(r"\bS\.sampleRate\b", "44100", e)
    e = re.sub(r"\bS\.(?:IN|OUT)\.chunkSize\b", "256", e)
    e = re.sub(r"\bMath\.(\w+)", r"M_\1", e)
    e = re.sub(r"new Float32Array\(([^()]*)\)", r"F32(\1)", e)
    e = re.sub(r"\.push\(", ".append(", e)
    e = re.sub(r"([\w.]+(?:\[[^\]]*\])*)\s*\|\|\s*(F32\([^()]*\))", r"(\1 if \1 is not None else \2)", e)   # x || new Float32Array(n)
    e = re.sub(r"([\w.]+(?:\[[^\]]*\])*)\s*\|\|\s*(zeroChannel|zeroChunk)\b", r"(\1 if \1 is not None else \2)", e)   # a missing channel reads as silence
    e = re.sub(r"^throw\s+(.*)$", r"raise Exception(\1)", e)
    e = re.sub(r"([\w.]+(?:\[[^\]]*\])*)\.length\b", r"len(\1)", e)
    e = re.sub(r"([\w.]+)\.constructor\s*==\s*Number", r"isinstance(\1, (int, float))", e)
    e = re.sub(r"\s*==\s*undefined\b", " is None", e)
    e = re.sub(r"\s*!=\s*undefined\b", " is not None", e)
    e = re.sub(r"=\s*\[\s*\]\s*$", "= JSArray()", e)
    e = re.sub(r"\b(?:Unit|config)\.standardChunkSize\b", "256", e)
    e = re.sub(r"\btrue\b", "True", e)
    e = re.sub(r"\bfalse\b", "False", e)
    e = re.sub(r"&&", " and ", e)
    e = re.sub(r"!(?!=)", " not ", e)
    # (x || 0) and x || 0 with a simple left operand
    e = re.sub(r"([\w.]+(?:\[[^\]]*\])*)\s*\|\|\s*0\b", r"js_or(\1, 0)", e)
    if "||" in e:
        raise ValueError("unsupported ||: " + e)
    # a % b with simple operands (an identifier / member / index chain, a parenthesised sum, or a number)
    opnd = r"(\((?:[^()]|\([^()]*\))*\)|[\w.]+(?:\[(?:[^\[\]]|\[[^\[\]]*\])*\]|\([^()]*\))*)"
    for _ in range(4):
        e2 = re.sub(opnd + r"\s*%\s*(?!=)" + opnd, r"js_mod(\1, \2)", e)
        if e2 == e:
            break
        e = e2
    return e


def _stmt(s):
    s = s.strip().rstrip(";")
    if not s:
        return []
    s = re.sub(r"^(?:var|const|let)\s+", "", s)
    m = re.match(r"^([\w.\[\]]+)\+\+$", s)
    if m:
        return [_expr(m.group(1)) + " += 1"]
    m = re.match(r"^([\w.\[\]]+)--$", s)
    if m:
        return [_expr(m.group(1)) + " -= 1"]
    m = re.match(r"^(.+?)\s*%=\s*(.+)$", s)
    if m:
        return ["%s = js_mod(%s, %s)" % (_expr(m.group(1)), _expr(m.group(1)), _expr(m.group(2)))]
    m = re.match(r"^(.+?[^=!<>])=\s*([^=].*?)\s\?\s(.*?)\s:\s(.*)$", s)   # lhs = cond ? a : b
    if m:
        return ["%s= (%s) if (%s) else (%s)" % (_expr(m.group(1)), _expr(m.group(3)), _expr(m.group(2)), _expr(m.group(4)))]
    if re.match(r"^\w+$", s):  # `var fraction`
        return [s + " = None"]
    # chained assignment a = b = expr (outData[c] = outData[c] || new ...): not needed by the covered bodies
    return [_expr(s)]


def translate(body, name="f", args="S"):
    body = re.sub(r"//[^\n]*", "", body)   # (a line comment hides a `/*` inside it: Filter.js:38 relies on that)
    body = re.sub(r"/\*.*?\*/", "", body, flags=re.S)
    raw = [l.rstrip() for l in body.splitlines() if l.strip()]
    lines = []
    for l in raw:  # continuation lines: the previous one ends, or this one starts, with a binary operator
        t = l.strip()
        if lines and (re.search(r"[+\-*/=(,]$", lines[-1].strip()) and not lines[-1].strip().endswith("++") and not lines[-1].strip().endswith("--")
                      or re.match(r"^[+\-*/)?:=]", t) or lines[-1].count("(") > lines[-1].count(")")):
            lines[-1] = lines[-1] + " " + t
        else:
            lines.append(t)
    out = ["def %s(%s):" % (name, args)]
    indent = 1
    once = []      # indentation levels opened by a brace-less header: closed after one statement
    loops = {}     # body indentation level -> the `i += 1` of a counting for loop (a while loop here: JavaScript leaves the
                   # counter at its final value, which ConcatChannels.js:24 relies on)
    switches = []  # [level of the switch's brace, a case body is open]
    n_sw = [0]

    def emit(text):
        out.append("    " * indent + text)

    def dedent():
        nonlocal indent
        if indent in loops:
            emit(loops.pop(indent))
        indent -= 1

    def close_once(next_is_else=False):
        # a brace-less body ends after one statement; when an `else` follows, only the `if` body it belongs to ends
        # here -- enclosing brace-less bodies stay open until the else branch is done
        while once and once[-1] == indent:
            once.pop()
            dedent()
            if next_is_else:
                return

    for li, l in enumerate(lines):
        nxt_else = li + 1 < len(lines) and lines[li + 1].strip().startswith("else")
        while l.startswith("}"):
            if switches and switches[-1][1] and indent == switches[-1][0] + 1:   # the brace that ends a switch: close the open case first
                dedent()
                switches.pop()
            elif switches and not switches[-1][1] and indent == switches[-1][0]:
                switches.pop()
            dedent()
            l = l[1:].strip()
            if not (l.startswith("else")):
                close_once()
        if not l:
            continue
        m = re.match(r"^switch\s*\((.*)\)\s*\{\s*$", l)
        if m:
            n_sw[0] += 1
            emit("_sw%d = %s" % (n_sw[0], _expr(m.group(1))))
            emit("if False:")
            indent += 1
            emit("pass")
            switches.append([indent, False, n_sw[0]])
            continue
        m = re.match(r"^case\s+(.+?)\s*:\s*$", l)
        if m and switches:
            if switches[-1][1]:
                dedent()
            indent -= 1
            emit("elif _sw%d == %s:" % (switches[-1][2], _expr(m.group(1))))
            indent += 2
            switches[-1][1] = True
            switches[-1][0] = indent - 1
            emit("pass")
            continue
        if re.match(r"^break\s*;?$", l) and switches:
            continue
        m = re.match(r"^else\s+if\s*\((.*)\)\s*(\{)?\s*$", l)
        if m:
            emit("elif %s:" % _expr(m.group(1)))
            indent += 1
            if not m.group(2):
                once.append(indent)
            continue
        m = re.match(r"^else\s*(\{)?\s*(.*)$", l)
        if m and l.startswith("else"):
            emit("else:")
            indent += 1
            rest = m.group(2).strip()
            if not m.group(1):
                if rest:
                    for s in _stmt(rest):
                        emit(s)
                    dedent()
                    close_once(nxt_else)
                else:
                    once.append(indent)
            continue
        m = re.match(r"^for\s*\(\s*(?:var\s+)?(\w+)\s*=\s*(.+?);\s*\1\s*<\s*([^|]+?);\s*\1\+\+\s*\)\s*(\{)?\s*$", l)
        if m:
            emit("%s = %s" % (m.group(1), _expr(m.group(2))))
            emit("while %s < %s:" % (m.group(1), _expr(m.group(3))))
            indent += 1
            loops[indent] = m.group(1) + " += 1"
            if not m.group(4):
                once.append(indent)
            continue
        m = re.match(r"^for\s*\(\s*var\s+(\w+)\s*=\s*(.+?);\s*\1\s*<\s*(.+?)\s*\|\|\s*\1\s*<\s*(.+?);\s*\1\+\+\s*\)\s*(\{)?\s*$", l)
        if m:  # for(var c=0; c<a.length || c<b.length; c++)
            emit("%s = %s" % (m.group(1), _expr(m.group(2))))
            emit("while %s < %s or %s < %s:" % (m.group(1), _expr(m.group(3)), m.group(1), _expr(m.group(4))))
            indent += 1
            loops[indent] = m.group(1) + " += 1"
            if not m.group(5):
                once.append(indent)
            continue
        m = re.match(r"^for\s*\(\s*var\s+(\w+)\s+in\s+(.+?)\)\s*(\{)?\s*$", l)
        if m:
            emit("for %s in range(len(%s)):" % (m.group(1), _expr(m.group(2))))
            indent += 1
            if not m.group(3):
                once.append(indent)
            continue
        m = re.match(r"^(if|while)\s*\((.*)\)\s*(\{)?\s*$", l)
        if m:
            emit("%s %s:" % (m.group(1), _expr(m.group(2))))
            indent += 1
            if not m.group(3):
                once.append(indent)
            continue
        m = re.match(r"^(if|while)\s*\((.*?)\)\s+(\S.*)$", l)  # header and statement on one line
        if m and not m.group(3).startswith("{"):
            emit("%s %s:" % (m.group(1), _expr(m.group(2))))
            indent += 1
            for s in _stmt(m.group(3)):
                emit(s)
            dedent()
            close_once(nxt_else)
            continue
        for s in _stmt(l):
            emit(s)
        close_once(nxt_else)
    while indent > 1:
        dedent()
    src = "\n".join(out) + "\n"
    compile(src, "<js2py>", "exec")
    return src
