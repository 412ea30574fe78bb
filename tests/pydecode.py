"""A second, independent opinion on the JSON -> Go-value layer: what encoding/json + reflect.DeepEqual do with a
Kubernetes object, written in plain Python.  Test infrastructure.

Only the TYPE TABLE is shared with the product (it is read out of csrc/host/k8s.cpp: which field is a string, a
pointer, a slice, a map, a struct value).  Decoding and equality are restated here from Go's rules and share no code
with json.hpp / gval.hpp / k8s.cpp / the oracles:

  decode(value, type) -> a canonical Python value in which Go's ZERO is None and everything else is explicit
    leaf (s, i, b, f)   "" / 0 / false / null / absent -> None, else the value
    [T   slice          null / absent -> None (nil);  [] -> [] (non-nil, empty);  elements decoded in order
    *T   pointer        null / absent -> None (nil);  else ("ptr", decode(T))  (a pointer to a zero value is NOT nil)
    {s   map            null / absent -> None (nil);  {} -> {} (non-nil, empty)
    Name struct VALUE   dict of the non-zero fields; all fields zero -> None; unknown JSON keys are dropped (Go ignores them)
    q                   a Quantity is a JSON string; its parsing is not restated here (tests/cases.py QUANTITY_CASES pin it):
                        equal strings are equal quantities, and the tests only ever compare strings that parse apart
    ios                 IntOrString: a JSON number is {Type Int, IntVal}, a JSON string {Type String, StrVal}
    any                 interface{}: kept as parsed, null -> None
  reflect.DeepEqual(a, b)  ==  decode(a) == decode(b)   (nil != empty, pointers by pointee, slices in order, maps unordered)
"""
import json
import os
import re

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _schema():
    text = open(os.path.join(ROOT, "odh-kubeflow_b200", "csrc", "host", "k8s.cpp")).read()
    body = re.search(r'kSchema = R"\((.*?)\)";', text, re.S).group(1)
    types = {}
    for line in body.strip().splitlines():
        name, fields = line.split(":", 1)
        types[name.strip()] = dict(f.split(":", 1) for f in fields.split())
    return types


TYPES = _schema()


def _generic(v):
    """interface{} / already-structured values: null is nil, everything else as parsed (dict keys unordered)."""
    if isinstance(v, dict):
        return {k: _generic(x) for k, x in v.items()}
    if isinstance(v, list):
        return [_generic(x) for x in v]
    return v


def decode(v, t):
    if t in ("s", "i", "b", "f"):
        return None if v in (None, "", 0, False) else v
    if t == "q":
        return None if v is None else ("quantity", str(v))
    if t == "ios":
        if v is None or v == 0:
            return None                       # IntOrString{} == IntOrString{Type: Int, IntVal: 0}
        return ("int", v) if isinstance(v, int) else ("string", v)
    if t == "any":
        return _generic(v)
    if t[0] == "[":
        return None if v is None else [decode(x, t[1:]) for x in v]
    if t[0] == "*":
        return None if v is None else ("ptr", decode(v, t[1:]))
    if t == "{s":
        return None if v is None else {k: ("" if x is None else x) for k, x in v.items()}
    if t == "{q":
        return None if v is None else {k: decode(x, "q") for k, x in v.items()}
    fields = TYPES[t]
    if v is None:
        return None
    out = {}
    for k, x in v.items():
        if k in fields:                       # a key the Go struct does not have is ignored by encoding/json
            d = decode(x, fields[k])
            if d is not None:
                out[k] = d
    return out or None


def deep_equal(a_json, b_json, type_name):
    """reflect.DeepEqual of the two documents decoded into Go type `type_name`."""
    a = json.loads(a_json) if isinstance(a_json, str) else a_json
    b = json.loads(b_json) if isinstance(b_json, str) else b_json
    return decode(a, type_name) == decode(b, type_name)
