"""Dusp text -> parse tree.  Restatement of the reference's recursive-descent
scanner ``src/parseDSP/``: same token rules, same operator
table order (= binding order), same quirks (a number token is a run of
``[0-9.-]`` so ``200-100`` is ONE token parsed as 200; words are
``[a-zA-Z_]`` only; object arguments must be whitespace separated).

Nodes are dicts with the reference's field names (``type``, ``length``, ...).
"""
import json
import math
import re

# src/parseDSP/config.js:2-19 -- list order determines binding order
OPERATORS = ["->", "at", "^", "*", "/", "@", "+", "-", "~!", "!", ",", "->", ">|", "|<", "for", "then"]
BASE_SHORTHANDS = ["O", "Z", "Sq", "A", "D", "t", "random", "LP", "AP", "HP"]  # config.js:26

_NUMBER_CH = re.compile(r"[0-9.\-]")
_WORD_CH = re.compile(r"[a-zA-Z_]")
_WORD_DIGIT_CH = re.compile(r"[a-zA-Z0-9_]")
_PARSEFLOAT = re.compile(r"^[+-]?(\d+\.?\d*(?:[eE][+-]?\d+)?|\.\d+(?:[eE][+-]?\d+)?)")


def _shorthand_names():
    from . import components, patches
    return BASE_SHORTHANDS + list(components.__all__) + list(patches.__all__)  # config.js:28-30


def _ch(s, i):
    return s[i] if 0 <= i < len(s) else None


def parse_float(tok):
    """JavaScript parseFloat: longest numeric prefix, else NaN."""
    m = _PARSEFLOAT.match(tok.strip())
    return float(m.group(0)) if m else math.nan


def skipWhitespace(s, i0=0):  # skipWhitespace.js:2-11
    i = i0
    while i < len(s):
        if s[i].isspace():
            i += 1
            continue
        return i
    return i


def countWhitespace(s, i0=0):  # countWhitespace.js:2-10 (undefined -> None at end of string)
    i = i0
    while i < len(s):
        if s[i].isspace():
            i += 1
            continue
        return i - i0
    return None


def getWord(s, i0=0):  # getWord.js:3-20
    i = i0
    while i < len(s):
        if not _WORD_CH.match(s[i]):
            return None if i == i0 else s[i0:i]
        i += 1
    return s[i0:i]  # open-ended word (the reference warns); "" is falsy like in JS


def getWordWithDigits(s, i0=0):  # getWordWithDigits.js:3-15
    i = i0
    while i <= len(s):
        c = _ch(s, i)
        if c is None or not _WORD_DIGIT_CH.match(c):
            return None if i == i0 else s[i0:i]
        i += 1
    return None


def getNumber(s, i0=0):  # getNumber.js:3-17
    i = i0
    while i <= len(s):
        c = _ch(s, i)
        if c is None or not _NUMBER_CH.match(c):
            if i == i0:
                return None
            return {"type": "number", "n": parse_float(s[i0:i]), "length": i - i0}
        i += 1
    return None


def getString(s, i0=0):  # getString.js:1-23
    c = _ch(s, i0)
    if c not in ('"', "'"):
        return None
    i1 = i0
    while True:
        i1 = s.find(c, i1 + 1)  # reference restarts from i0+1 and would spin on an escaped quote
        if i1 == -1:
            return None
        if s[i1 - 1] != "\\":
            break
    return {"type": "string", "string": s[i0 + 1:i1], "length": i1 - i0 + 1}


def getObjectReference(s, i0=0):  # getObjectReference.js:3-18
    if _ch(s, i0) != "#":
        return None
    ref = getWordWithDigits(s, i0 + 1)
    if ref:
        return {"id": ref, "length": len(ref) + 1, "type": "id"}
    return None


def getSpecific(search, s, i0=0):  # getSpecific.js:1-8
    return search if s.startswith(search, i0) else None


def getOperator(s, i0=0):  # getOperator.js:1-14 (longest match wins)
    winner = ""
    for op in OPERATORS:
        got = getSpecific(op, s, i0)
        if got and len(got) > len(winner):
            winner = got
    return winner or None


def getOperatorOperand(s, i0=0):  # getOperatorOperand.js:1-23
    operator = getOperator(s, i0)
    if operator is None:
        return None
    bindingOrder = OPERATORS.index(operator)
    i2 = skipWhitespace(s, i0 + len(operator))
    b = getSimpleExpression(s, i2)
    if not b:
        return None
    return {"type": "operation", "operator": operator, "b": b, "bindingOrder": bindingOrder,
            "length": i2 - i0 + b["length"]}


def getDotProperty(s, i0):  # getDotProperty.js:1-16
    i1 = skipWhitespace(s, i0)
    if _ch(s, i1) != ".":
        return None
    i2 = skipWhitespace(s, i1 + 1)
    prop = getWord(s, i2)
    if not prop:
        return None
    return {"type": "property", "property": prop, "length": i2 - i0 + len(prop)}


def getShorthand(s, i0=0):  # getShorthand.js:1-33
    constr = getWord(s, i0)
    if not constr or constr not in _shorthand_names():
        return None
    i = i0 + len(constr)
    args = []
    n = getNumber(s, i)
    if n:
        args.append(n)
        i += n["length"]
        while _ch(s, i) == ",":
            i += 1
            n = getNumber(s, i)
            if not n:
                return None
            args.append(n)
            i += n["length"]
    return {"type": "shorthand", "constructorAlias": constr, "arguments": args, "length": i - i0}


def getAttribute(s, i0=0):  # getAttribute.js:4-29
    prop = getWord(s, i0)
    if not prop:
        return None
    i1 = i0 + len(prop) + (countWhitespace(s, i0 + len(prop)) or 0)
    if _ch(s, i1) not in ("=", ":"):
        return None
    i2 = i1 + 1 + (countWhitespace(s, i1 + 1) or 0)
    value = getExpression(s, i2)
    if not value:
        return None
    return {"type": "attribute", "property": prop, "value": value, "length": i2 - i0 + value["length"]}


def getArgument(s, i0):  # getArgument.js:1-26
    ref = getObjectReference(s, i0)
    if ref:
        return ref
    attr = getAttribute(s, i0)
    if attr:
        return attr
    arg = getExpression(s, i0)
    if arg:
        return {"type": "unnamedArgument", "value": arg, "length": arg["length"]}
    flag = getWord(s, i0)
    if flag:
        return {"type": "flag", "flag": flag, "length": len(flag)}
    return None


def getObject(s, i0=0):  # getObject.js:1-65
    if _ch(s, i0) != "[":
        return None
    i1 = skipWhitespace(s, i0 + 1)
    constructor = getWord(s, i1)
    if not constructor:
        return None
    obj = {"type": "object", "constructor": constructor, "arguments": [], "flags": [], "attributes": []}
    iN = i1 + len(constructor)
    while True:
        if _ch(s, iN) == "]":
            obj["length"] = iN - i0 + 1
            return obj
        if countWhitespace(s, iN):
            iN = skipWhitespace(s, iN)
            if _ch(s, iN) == "]":
                obj["length"] = iN - i0 + 1
                return obj
            arg = getArgument(s, iN)
            if not arg:
                return None
            t = arg["type"]
            if t == "id":
                obj["id"] = arg["id"]
            elif t == "attribute":
                obj["attributes"].append(arg)
            elif t == "unnamedArgument":
                obj["arguments"].append(arg)
            elif t == "flag":
                obj["flags"].append(arg)
            else:
                return None
            iN += arg["length"]
        else:
            return None
        if not iN < len(s):
            break
    return None


def getObjectOrObjectProperty(s, i0=0):  # getObjectOrObjectProperty.js:1-21
    obj = getObject(s, i0) or getObjectReference(s, i0) or getShorthand(s, i0)
    if not obj:
        return None
    prop = getDotProperty(s, i0 + obj["length"])
    if prop:
        return {"type": "objectProperty", "property": prop["property"], "object": obj,
                "length": obj["length"] + prop["length"]}
    return obj


def getJSON(s, i0=0):  # getJSON/index.js:1-27 (delegated to a standard JSON scanner)
    try:
        o, end = json.JSONDecoder().raw_decode(s, i0)
    except ValueError:
        return None
    return {"type": "json", "o": o, "length": end - i0}


def getSimpleExpression(s, startIndex=0):  # getExpression.js:59-100
    c = _ch(s, startIndex)
    if c == "{":
        return getJSON(s, startIndex)
    if c == "(":
        i = skipWhitespace(s, startIndex + 1)
        expr = getExpression(s, i)
        if not expr:
            return None
        i = skipWhitespace(s, i + expr["length"])
        if _ch(s, i) != ")":
            return None
        expr["length"] = i + 1 - startIndex
        expr["bracketed"] = True
        return expr
    return (getObjectReference(s, startIndex) or getNumber(s, startIndex)
            or getObjectOrObjectProperty(s, startIndex) or getShorthand(s, startIndex)
            or getString(s, startIndex))


def getExpression(s, i0=0):  # getExpression.js:1-57
    i0 = i0 or 0
    expr0 = getSimpleExpression(s, i0)
    if expr0 is None:
        return None
    iN = i0 + expr0["length"]
    oList = [expr0]
    while True:
        op = getOperatorOperand(s, skipWhitespace(s, iN))
        if op:
            oList.append(op)
            iN = skipWhitespace(s, iN) + op["length"]
        else:
            break
    length = iN - i0
    for o in oList:
        o.pop("length", None)
    while len(oList) > 1:  # precedence fold, :37-53
        for i in range(1, len(oList)):
            if i == len(oList) - 1 or oList[i]["bindingOrder"] < oList[i + 1]["bindingOrder"]:
                if i > 1:
                    oList[i]["a"] = oList[i - 1]["b"]
                    oList[i - 1]["b"] = oList[i]
                else:
                    oList[i]["a"] = oList[i - 1]
                    oList[i - 1] = oList[i]
                del oList[i]
                break
    oList[0]["length"] = length
    return oList[0]


def getExpressions(s, i0=0):  # getExpressions.js:1-25
    exprs = []
    i = i0
    while i < len(s):
        expr = getExpression(s, i)
        if expr:
            exprs.append(expr)
        elif i == i0:
            return None
        else:
            return {"expressions": exprs, "length": i - i0}
        i = skipWhitespace(s, i + expr["length"])
    if i != i0:
        return {"expressions": exprs, "length": i - i0}
    return None
