"""Parse tree -> live Unit graph.  Restatement of ``src/construct/*`` and
``src/unDusp.js``.  Arithmetic operators build units through
``quick`` (constant folding); ``for`` / ``then`` / ``!`` / ``~!`` need run-time
graph edits and are out of scope -- they raise.  ``at`` is
supported for times the reference fires before the first tick (t < 256 samples).
"""
from . import components as _components
from . import patches as _patches
from . import config, parse, quick
from .graph import DuspError, UnitOrPatch, Outlet, Inlet, _is_number


def _namespace():  # src/patchesAndComponents.js:1-8 (patches overwrite components)
    ns = {k: getattr(_components, k) for k in _components.__all__}
    ns.update({k: getattr(_patches, k) for k in _patches.__all__})
    return ns


# src/construct/shorthandConstructors.js:3-46
def _osc(wave):
    def make(frequency=None):
        osc = _components.Osc(frequency)
        if wave:
            osc.waveform = wave
        return osc
    return make


SHORTHAND_CONSTRUCTORS = {
    "O": _osc(None),
    "Z": _osc("saw"),
    "Sq": _osc("square"),
    "A": lambda time=None: _components.Shape("attack", time).trigger(),
    "D": lambda time=None: _components.Shape("decay", time).trigger(),
    "t": lambda: _components.Timer(),
    "LP": lambda freq=None: _components.Filter(None, freq),
    "HP": lambda freq=None: _components.Filter(None, freq, "HP"),
    "AP": lambda delaytime=None, feedback=None: _components.AllPass(delaytime, feedback),
    "random": lambda: config.host_random.random(),
}


def _index_get(index, key):
    return index.get(key) if isinstance(index, dict) else None  # property read on a Number -> undefined


def _index_set(index, key, val):
    if isinstance(index, dict):
        index[key] = val  # assignment to a Number primitive is silently dropped


def constructNumber(o):  # constructNumber.js:1-10
    if isinstance(o, str):
        o = parse.getNumber(o)
    if not o or o.get("type") != "number":
        return None
    return o["n"]


def constructObject(o, index=None):  # constructObject.js:3-52
    if index is None or (isinstance(index, int) and index == 0):
        index = {}
    if isinstance(o, str):
        o = parse.getObject(o)
    if not o or o["type"] != "object":
        return None
    ctor = _namespace().get(o["constructor"])
    if ctor is None:
        raise DuspError("Unknown object constructor: " + o["constructor"])
    # `o.arguments.map(constructExpression)`: map passes (element, position, array),
    # so each unnamed argument is constructed with index = its POSITION (:11)
    args = [constructExpression(a, i, o["arguments"]) for i, a in enumerate(o["arguments"])]
    obj = ctor(*args)
    if o.get("id"):
        obj.label = o["id"]
    existing = _index_get(index, obj.label)
    if existing is not None:
        if existing is not obj:
            raise DuspError("Duplicate objects for id:")
    else:
        _index_set(index, obj.label, obj)
    for arg in o["attributes"]:
        prop = arg["property"]
        upper = prop.upper()
        target = getattr(obj, upper, None)
        if isinstance(target, Inlet):
            prop = upper
        if arg["type"] == "attribute":
            setattr(obj, prop, constructExpression(arg["value"], index))
        else:
            raise DuspError("unknown argument type: ")
    flags = getattr(obj, "dusp_flags", None)
    if flags:
        for f in o["flags"]:
            fn = flags.get(f["flag"])
            if fn:
                fn(obj)
    return obj


def constructObjectReference(o, index=None):  # constructObjectReference.js:1-9
    if isinstance(o, str):
        o = parse.getObjectReference(o)
    got = _index_get(index, o["id"])
    if got is not None:
        return got
    raise DuspError("Error: Referencing an object which has not been declared: #" + o["id"])


def constructShorthand(o, index=None):  # constructShorthand.js:1-17
    if isinstance(o, str):
        o = parse.getShorthand(o)
    args = [constructNumber(a) for a in o["arguments"]]
    ctor = SHORTHAND_CONSTRUCTORS.get(o["constructorAlias"])
    if ctor:
        return ctor(*args)
    ctor = getattr(_components, o["constructorAlias"], None) if o["constructorAlias"] in _components.__all__ else None
    if ctor:
        return ctor(*args)
    raise DuspError("Unknown shorthand: " + o["constructorAlias"])


def constructOperation(o, index=None, destinations=None):  # constructOperation.js:1-87
    if not o.get("a") or not o.get("b") or not o.get("operator"):
        raise DuspError("could not construct operation")
    a = constructExpression(o["a"], index)
    b = constructExpression(o["b"], index)
    op = o["operator"]
    if op == "*":
        return quick.multiply(a, b)
    if op == "/":
        return quick.divide(a, b)
    if op == "+":
        return quick.add(a, b)
    if op == "-":
        return quick.subtract(a, b)
    if op == ",":
        return quick.concat(a, b)
    if op == "@":
        return _components.Pan(a, b)
    if op == "^":
        return quick.pow(a, b)
    if op == "->":
        if isinstance(b, UnitOrPatch):
            b.defaultInlet.set(a)
            return b
        raise DuspError("unknown use of -> operator")
    if op == "|<":
        return quick.clipBelow(b, a)
    if op == ">|":
        return quick.clipAbove(a, b)
    if op == "at":  # :63-69
        if not hasattr(a, "stop") or not hasattr(a, "trigger"):
            raise DuspError("invalid use of 'at' operator")
        a.stop()
        a.scheduleTrigger(b)
        return a
    if op == "!":  # regular retrigger :71-76
        if not hasattr(a, "stop") or not hasattr(a, "trigger"):
            raise DuspError("invalid use of '!' operator")
        a.trigger()
        _components.Retriggerer(a, b)
        return a
    if op == "~!":  # SporadicRetriggerer :78-82
        if not hasattr(a, "stop") or not hasattr(a, "trigger"):
            raise DuspError("invalide use of '!~' operator")
        _components.SporadicRetriggerer(a, b)
        return a
    if op == "for":  # :36-43
        if _is_number(a):
            a = _components.Repeater(a)
        if hasattr(a, "scheduleFinish"):
            a.scheduleFinish(b)
        else:
            raise DuspError("invalid use of 'for' operator. First operand has no scheduleFinish function")
        return a
    if op == "then":  # :45-60 -- a run-time graph edit: when `a` finishes, the inlet that read it is handed `b`
        out = None
        if not destinations:
            out = _components.Repeater()
            out.IN = a
            destinations = [lambda x: setattr(out, "IN", x)]

        def on_finish(dst=destinations):
            for f in dst:
                f(b)
        if isinstance(a, UnitOrPatch):  # (a Number silently drops the property in JS)
            a.onFinish = on_finish
        return out if out is not None else a
    raise DuspError("Unknown operator: " + op)


def constructObjectProperty(o, index=None):  # constructObjectProperty.js:1-4
    obj = constructExpression(o["object"], index)
    return getattr(obj, o["property"])


def constructExpression(o, index=None, destinations=None):  # constructExpression.js:1-37
    if isinstance(o, str):
        src = o
        o = parse.getExpression(o, index if isinstance(index, int) else 0)
        if o is None:
            raise DuspError("Can't construct expression: " + src)
    t = o["type"]
    if t == "object":
        return constructObject(o, index)
    if t == "number":
        return constructNumber(o)
    if t == "id":
        return constructObjectReference(o, index)
    if t == "operation":
        return constructOperation(o, index, destinations)
    if t == "objectProperty":
        return constructObjectProperty(o, index)
    if t == "shorthand":
        return constructShorthand(o, index)
    if t == "unnamedArgument":
        return constructExpression(o["value"], index)
    if t == "string":
        return o["string"]
    if t == "json":
        return o["o"]
    raise DuspError("Unknown expression type: " + str(t))


def unDusp(o):  # unDusp.js:4-16
    if o is None:
        return None
    if isinstance(o, str):
        return constructExpression(o)
    if isinstance(o, (int, float)):
        return o
    if isinstance(o, (UnitOrPatch, Outlet)):
        return o
    return None
