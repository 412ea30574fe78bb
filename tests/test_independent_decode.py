"""The decode layer held against an INDEPENDENT restatement (tests/pydecode.py: Go's encoding/json + reflect.DeepEqual
rules in plain Python, sharing only the type table with the product).

The semantic oracle walks trees built by the product's decoder, so oracle and packer would follow a decoder mistake
together.  Here every pair goes in as JSON TEXT on both routes: the product decodes it (json.hpp -> k8s.cpp -> GVal),
the oracle and the packer judge it; the Python route decodes the same text by its own rules.  The verdicts must agree
on real fleet objects with their drift classes, and on copies whose JSON was mutated exactly where Go is subtle
(null / [] / {} / zero values / unknown keys)."""
import copy
import json

import numpy as np
import pytest

import oracle as O
import pydecode

host, N = O.host, O.N


def _product_verdict(a_text, b_text, type_name="PodSpec"):
    a, b = host.Obj.from_json(type_name, a_text), host.Obj.from_json(type_name, b_text)
    eq = O.deep_equal(a, b)
    assert (a.digest() == b.digest()) == eq          # the packer's digest follows the oracle
    return eq


def _as_api_json(v):
    """The debug rendering shows Quantity and IntOrString as Go structs; the API (and both decoders) take a string or
    a number there.  Quantities become a string that is the same for equal structs and parses apart for different
    ones (the cached canonical string if there is one, else value and scale spelled as an exponent)."""
    if isinstance(v, dict):
        if v and set(v) <= {"i", "d", "s", "Format"}:
            if v.get("s"):
                return v["s"]
            i = v.get("i") or {}
            return "%de%d" % (i.get("value", 0), i.get("scale", 0))
        if v and set(v) <= {"Type", "IntVal", "StrVal"}:
            return v.get("StrVal", "") if v.get("Type") else v.get("IntVal", 0)
        return {k: _as_api_json(x) for k, x in v.items()}
    if isinstance(v, list):
        return [_as_api_json(x) for x in v]
    return v


def _pod(obj):
    return _as_api_json(obj.to_python()["spec"]["template"]["spec"])


@pytest.mark.parametrize("kind,seed,n", [(host.FLEET_KUBEFLOW, host.SEED_BASE + 2, 700), (host.FLEET_ODH, host.SEED_BASE + 4, 500)])
def test_fleet_podspecs_agree_with_the_python_decoder(kind, seed, n):
    fleet = host.Fleet(kind, seed, n, 40)
    differ = 0
    for i in range(n):
        a, b = json.dumps(_pod(fleet.get(i, "sts_from"))), json.dumps(_pod(fleet.get(i, "sts_to")))
        want = pydecode.deep_equal(a, b, "PodSpec")
        assert _product_verdict(a, b) == want, (i, fleet.drift_class(i))
        # and the verdict the fleet's own (never serialised) objects get
        l1 = O.copy_statefulset_fields(fleet.get(i, "sts_from"), fleet.get(i, "sts_to"), 1)
        assert bool(l1 & N.STS_PODSPEC) == (not want), (i, fleet.drift_class(i))
        differ += not want
    assert 0.15 * n < differ < 0.6 * n                # the drift classes that touch the PodSpec are all there


# JSON-level edits at the places where Go is subtle.  Each takes a PodSpec dict and returns an edited deep copy.
def _c0(p):
    return p["containers"][0]


MUTATIONS = {
    "command_empty": lambda p: _c0(p).__setitem__("command", []),
    "command_null": lambda p: _c0(p).__setitem__("command", None),
    "args_empty": lambda p: _c0(p).__setitem__("args", []),
    "env_null": lambda p: _c0(p).__setitem__("env", None),
    "env_drop": lambda p: _c0(p).pop("env", None),
    "envfrom_empty": lambda p: _c0(p).__setitem__("envFrom", []),
    "seccontext_empty": lambda p: _c0(p).__setitem__("securityContext", {}),
    "seccontext_null": lambda p: _c0(p).__setitem__("securityContext", None),
    "resources_empty": lambda p: _c0(p).__setitem__("resources", {}),
    "limits_empty": lambda p: _c0(p).setdefault("resources", {}).__setitem__("limits", {}),
    "workingdir_blank": lambda p: _c0(p).__setitem__("workingDir", ""),
    "tty_false": lambda p: _c0(p).__setitem__("tty", False),
    "stdin_true": lambda p: _c0(p).__setitem__("stdin", True),
    "nodeselector_empty": lambda p: p.__setitem__("nodeSelector", {}),
    "nodeselector_null": lambda p: p.__setitem__("nodeSelector", None),
    "tolerations_empty": lambda p: p.__setitem__("tolerations", []),
    "hostnetwork_false": lambda p: p.__setitem__("hostNetwork", False),
    "grace_zero": lambda p: p.__setitem__("terminationGracePeriodSeconds", 0),
    "grace_null": lambda p: p.__setitem__("terminationGracePeriodSeconds", None),
    "podseccontext_empty": lambda p: p.__setitem__("securityContext", {}),
    "automount_false": lambda p: p.__setitem__("automountServiceAccountToken", False),
    "volumes_null": lambda p: p.__setitem__("volumes", None),
    "imagepullsecrets_empty": lambda p: p.__setitem__("imagePullSecrets", []),
    "mount_readonly_false": lambda p: [m.__setitem__("readOnly", False) for m in (_c0(p).get("volumeMounts") or [])[:1]],
    "env_value_blank": lambda p: [e.__setitem__("value", "") for e in (_c0(p).get("env") or [])[:1]],
    "env_reorder": lambda p: _c0(p).__setitem__("env", list(reversed(_c0(p).get("env") or []))) if _c0(p).get("env") else None,
    "port_hostport_zero": lambda p: [x.__setitem__("hostPort", 0) for x in (_c0(p).get("ports") or [])[:1]],
}


@pytest.mark.parametrize("name", sorted(MUTATIONS))
def test_mutated_podspecs_agree_with_the_python_decoder(name):
    """The edit is applied to one side only, over 60 different fleet PodSpecs (ODH-shaped: two containers, six volumes):
    whatever Go would say about `edited` against `original`, the product must say the same — equal for a zero value
    written out or a null, different for an empty slice, map or pointer target that was nil."""
    fleet = host.Fleet(host.FLEET_ODH, host.SEED_BASE + 61, 60, 6)
    verdicts = set()
    for i in range(60):
        base = _pod(fleet.get(i, "sts_from"))
        edited = copy.deepcopy(base)
        MUTATIONS[name](edited)
        a, b = json.dumps(base), json.dumps(edited)
        want = pydecode.deep_equal(a, b, "PodSpec")
        assert _product_verdict(a, b) == want, (name, i)
        verdicts.add(want)
    assert verdicts                                     # (both verdicts occur for some edits: an env list may be absent)


def test_service_and_metadata_types_too():
    fleet = host.Fleet(host.FLEET_KUBEFLOW, host.SEED_BASE + 62, 300, 10)
    for i in range(300):
        a, b = _as_api_json(fleet.get(i, "svc_from").to_python()), _as_api_json(fleet.get(i, "svc_to").to_python())
        for path, t in ((("spec",), "ServiceSpec"), (("metadata",), "ObjectMeta")):
            x, y = a, b
            for k in path:
                x, y = x.get(k), y.get(k)
            ja, jb = json.dumps(x), json.dumps(y)
            assert _product_verdict(ja, jb, t) == pydecode.deep_equal(ja, jb, t), (i, t)


def test_unknown_keys_are_the_one_known_difference():
    base = {"containers": [{"name": "nb", "image": "img"}]}
    extra = {"containers": [{"name": "nb", "image": "img", "notAField": {"x": 1}}]}
    assert pydecode.deep_equal(base, extra, "PodSpec") is True           # Go ignores the key
    assert _product_verdict(json.dumps(base), json.dumps(extra)) is False  # kept here: DESIGN.md section 4
