"""Hand-built known-answer cases for the quirks of SURVEY.md 8(a)/8(c).

Expected values are DERIVED from the cited reference lines (the reference has no
test for reconcilehelper, nor for C2/C3/C8), so they pin the oracle and the
kernels to each other and to a careful reading of the Go code — not to a Go run.
COMMON = components/common/reconcilehelper/util.go,
NBCTL  = components/notebook-controller/controllers/notebook_controller.go,
CULL   = components/notebook-controller/controllers/culling_controller.go.
"""
import copy

NOW = (1767225600, 123456789)  # 2026-01-01T00:00:00.123456789Z

BASE_NOTEBOOK = {
    "metadata": {"name": "nb-a", "namespace": "team-x", "labels": {"app": "nb", "team": "x"},
                 "annotations": {"owner": "alice"}},
    "spec": {"template": {"spec": {
        "serviceAccountName": "default-editor",
        "containers": [{
            "name": "nb-a", "image": "img:1",
            "resources": {"requests": {"cpu": "500m", "memory": "1Gi"}},
            "volumeMounts": [{"name": "ws", "mountPath": "/home/jovyan"}]}],
        "volumes": [{"name": "ws", "persistentVolumeClaim": {"claimName": "ws-nb-a"}}]}}},
}


def _edit(obj, fn):
    o = copy.deepcopy(obj)
    fn(o)
    return o


def _c0(sts):
    return sts["spec"]["template"]["spec"]["containers"][0]


# Each case: (name, edit(desired_dict) -> live_dict, expected reason bits)
# reason bits: L=1 A=2 R=4 P=8 D1'(template labels copied)=16
STS_CASES = [
    ("identical", lambda d: None, 0),
    ("desired-only label is NOT detected (loop ranges over live keys, COMMON:109)",
     lambda d: d["metadata"]["labels"].pop("team"), 0),
    ("live-only label", lambda d: d["metadata"]["labels"].update({"extra": "1"}), 1),
    ("live label value changed", lambda d: d["metadata"]["labels"].update({"app": "other"}), 1),
    ("live (k,'') absent from desired reads as equal (missing key == '', COMMON:110)",
     lambda d: d["metadata"]["labels"].update({"blank": ""}), 0),
    ("live annotation: desired STS has none (NBCTL:439-443)",
     lambda d: d["metadata"].update({"annotations": {"rev": "3"}}), 2),
    ("live annotation with empty value is equal", lambda d: d["metadata"].update({"annotations": {"rev": ""}}), 0),
    ("replicas differ", lambda d: d["spec"].update({"replicas": 0}), 4),
    ("replicas differ and template labels differ: labels copied too (NBCTL:190-194)",
     lambda d: (d["spec"].update({"replicas": 0}), d["spec"]["template"]["metadata"]["labels"].update({"x": "y"})), 4 | 16),
    ("template labels differ but replicas equal: never reconciled", lambda d: d["spec"]["template"]["metadata"]["labels"].update({"x": "y"}), 0),
    ("env nil vs empty slice (reflect.DeepEqual, COMMON:128)", None, 8),  # filled below
    ("command [] vs absent", lambda d: _c0(d).update({"command": []}), 8),
    ("quantity 500m vs 0.5: same number, different scale", lambda d: _c0(d)["resources"]["requests"].update({"cpu": "0.5"}), 8),
    ("quantity 1Gi vs 1024Mi: only the cached string differs", lambda d: _c0(d)["resources"]["requests"].update({"memory": "1024Mi"}), 8),
    ("quantity 1Gi vs 1Gi", lambda d: _c0(d)["resources"]["requests"].update({"memory": "1Gi"}), 0),
    ("API-server defaults on the live PodSpec", lambda d: d["spec"]["template"]["spec"].update(
        {"restartPolicy": "Always", "dnsPolicy": "ClusterFirst", "terminationGracePeriodSeconds": 30}), 8),
    ("securityContext {} vs desired fsGroup", lambda d: d["spec"]["template"]["spec"].update({"securityContext": {}}), 8),
    ("extra volume", lambda d: d["spec"]["template"]["spec"]["volumes"].append({"name": "dshm", "emptyDir": {"medium": "Memory"}}), 8),
    ("third container beyond the slotted two", lambda d: d["spec"]["template"]["spec"]["containers"].extend(
        [{"name": "a"}, {"name": "b"}]), 8),
    ("field outside the schema table", lambda d: d["spec"]["template"]["spec"].update({"futureField": {"a": 1}}), 8),
    ("everything at once", lambda d: (d["metadata"]["labels"].update({"extra": "1"}), d["metadata"].update({"annotations": {"r": "1"}}),
                                      d["spec"].update({"replicas": 0}), _c0(d).update({"image": "img:2"})), 1 | 2 | 4 | 8),
]


def _vols(d):
    return d["spec"]["template"]["spec"]["volumes"]


def _req(d):
    return _c0(d)["resources"]["requests"]


# Cases that edit BOTH sides: (name, edit of desired, edit of live, expected reason).  They sit on the two places where
# the packer spends fewer slots on the usual value than on the general one (csrc/host/pack.cpp: a small Quantity in one
# slot, a volume whose source holds nothing but its reference in two) - the short and the long spelling of a tile
# must never compare equal to each other, and each must stay injective.
STS_BOTH_CASES = [
    ("pvc readOnly on live only", None, lambda d: _vols(d)[0]["persistentVolumeClaim"].update({"readOnly": True}), 8),
    ("pvc readOnly on both", lambda d: _vols(d)[0]["persistentVolumeClaim"].update({"readOnly": True}),
     lambda d: _vols(d)[0]["persistentVolumeClaim"].update({"readOnly": True}), 0),
    ("pvc readOnly false is the zero value", None, lambda d: _vols(d)[0]["persistentVolumeClaim"].update({"readOnly": False}), 0),
    ("volume without a source vs emptyDir {}", lambda d: _vols(d).append({"name": "x"}), lambda d: _vols(d).append({"name": "x", "emptyDir": {}}), 8),
    ("emptyDir {} vs emptyDir {medium: ''}", lambda d: _vols(d).append({"name": "x", "emptyDir": {}}),
     lambda d: _vols(d).append({"name": "x", "emptyDir": {"medium": ""}}), 0),
    ("emptyDir {} vs emptyDir {sizeLimit}", lambda d: _vols(d).append({"name": "x", "emptyDir": {}}),
     lambda d: _vols(d).append({"name": "x", "emptyDir": {"sizeLimit": "1Gi"}}), 8),
    ("same claim name under another source kind", lambda d: _vols(d).append({"name": "x", "secret": {"secretName": "n"}}),
     lambda d: _vols(d).append({"name": "x", "configMap": {"name": "n"}}), 8),
    ("two sources in one volume vs one", lambda d: _vols(d).append({"name": "x", "secret": {"secretName": "n"}}),
     lambda d: _vols(d).append({"name": "x", "secret": {"secretName": "n"}, "emptyDir": {}}), 8),
    ("configMap optional pointer false vs absent (a *bool)", lambda d: _vols(d).append({"name": "x", "configMap": {"name": "n"}}),
     lambda d: _vols(d).append({"name": "x", "configMap": {"name": "n", "optional": False}}), 8),
    ("volume: same source under another name", lambda d: _vols(d).append({"name": "x", "secret": {"secretName": "n"}}),
     lambda d: _vols(d).append({"name": "y", "secret": {"secretName": "n"}}), 8),
    ("volume: same name, other secret", lambda d: _vols(d).append({"name": "x", "secret": {"secretName": "n"}}),
     lambda d: _vols(d).append({"name": "x", "secret": {"secretName": "m"}}), 8),
    ("volume: plain source on both", lambda d: _vols(d).append({"name": "x", "secret": {"secretName": "n"}}),
     lambda d: _vols(d).append({"name": "x", "secret": {"secretName": "n"}}), 0),
    ("volume: a source the packer has no short form for, on both", lambda d: _vols(d).append({"name": "x", "hostPath": {"path": "/p"}}),
     lambda d: _vols(d).append({"name": "x", "hostPath": {"path": "/p"}}), 0),
    ("volume: hostPath path differs", lambda d: _vols(d).append({"name": "x", "hostPath": {"path": "/p"}}),
     lambda d: _vols(d).append({"name": "x", "hostPath": {"path": "/q"}}), 8),
    ("mount: same name, other path", None, lambda d: _c0(d)["volumeMounts"][0].update({"mountPath": "/elsewhere"}), 8),
    ("mount: other name, same path", None, lambda d: _c0(d)["volumeMounts"][0].update({"name": "other"}), 8),
    ("mount: readOnly set", None, lambda d: _c0(d)["volumeMounts"][0].update({"readOnly": True}), 8),
    ("mounts: names swapped between two mounts",
     lambda d: _c0(d).update({"volumeMounts": [{"name": "a", "mountPath": "/x"}, {"name": "b", "mountPath": "/y"}]}),
     lambda d: _c0(d).update({"volumeMounts": [{"name": "b", "mountPath": "/x"}, {"name": "a", "mountPath": "/y"}]}), 8),
    ("mounts: same two mounts, other order",
     lambda d: _c0(d).update({"volumeMounts": [{"name": "a", "mountPath": "/x"}, {"name": "b", "mountPath": "/y"}]}),
     lambda d: _c0(d).update({"volumeMounts": [{"name": "b", "mountPath": "/y"}, {"name": "a", "mountPath": "/x"}]}), 8),
    ("mount: an empty element against none", lambda d: _c0(d)["volumeMounts"].append({}), None, 8),
    ("container port: other name", None, lambda d: _c0(d)["ports"][0].update({"name": "other"}), 8),
    ("container port: other number", None, lambda d: _c0(d)["ports"][0].update({"containerPort": 9999}), 8),
    ("container port: hostPort set", None, lambda d: _c0(d)["ports"][0].update({"hostPort": 80}), 8),
    ("container ports: a second one on both", lambda d: _c0(d)["ports"].append({"containerPort": 1, "name": "z"}),
     lambda d: _c0(d)["ports"].append({"containerPort": 1, "name": "z"}), 0),
    ("memory 511Gi on both (one-slot spelling)", lambda d: _req(d).update({"memory": "511Gi"}), lambda d: _req(d).update({"memory": "511Gi"}), 0),
    ("memory 512Gi on both (2^39: the long spelling)", lambda d: _req(d).update({"memory": "512Gi"}), lambda d: _req(d).update({"memory": "512Gi"}), 0),
    ("memory 512Gi vs 511Gi", lambda d: _req(d).update({"memory": "512Gi"}), lambda d: _req(d).update({"memory": "511Gi"}), 8),
    ("memory 512Gi vs 549755813888 (same bytes, DecimalSI)", lambda d: _req(d).update({"memory": "512Gi"}),
     lambda d: _req(d).update({"memory": "549755813888"}), 8),
    ("memory 1Gi vs 1073741824 (same bytes, DecimalSI)", None, lambda d: _req(d).update({"memory": "1073741824"}), 8),
    ("cpu 1 vs 1000m", lambda d: _req(d).update({"cpu": "1"}), lambda d: _req(d).update({"cpu": "1000m"}), 8),
    ("cpu -1 vs 1", lambda d: _req(d).update({"cpu": "-1"}), lambda d: _req(d).update({"cpu": "1"}), 8),
    ("cpu -1 on both", lambda d: _req(d).update({"cpu": "-1"}), lambda d: _req(d).update({"cpu": "-1"}), 0),
    ("cpu 1e3 vs 1k (DecimalExponent vs DecimalSI)", lambda d: _req(d).update({"cpu": "1e3"}), lambda d: _req(d).update({"cpu": "1k"}), 8),
    ("cpu 8E on both (near the int64 edge)", lambda d: _req(d).update({"cpu": "8E"}), lambda d: _req(d).update({"cpu": "8E"}), 0),
    ("cpu 8E vs 8P", lambda d: _req(d).update({"cpu": "8E"}), lambda d: _req(d).update({"cpu": "8P"}), 8),
    ("cpu 100n vs 100u (scales -9 and -6)", lambda d: _req(d).update({"cpu": "100n"}), lambda d: _req(d).update({"cpu": "100u"}), 8),
    ("a quantity present with value 0 vs absent", lambda d: _req(d).update({"cpu": "0"}), lambda d: _req(d).pop("cpu"), 8),
    ("value 0 on both", lambda d: _req(d).update({"cpu": "0"}), lambda d: _req(d).update({"cpu": "0"}), 0),
]


def _env_nil_vs_empty(d):
    # desired has env [NB_PREFIX] (appended by setPrefixEnvVar); make live's second container carry env: []
    d["spec"]["template"]["spec"]["containers"].append({"name": "side", "env": []})


def sts_case_pairs(generate):
    """generate(notebook_dict) -> desired StatefulSet dict.  Yields (name, from, to, expected)."""
    desired = generate(BASE_NOTEBOOK)
    for name, edit, want in STS_CASES:
        if name.startswith("env nil vs empty"):
            d2 = _edit(desired, lambda d: d["spec"]["template"]["spec"]["containers"].append({"name": "side"}))
            yield name, d2, _edit(desired, _env_nil_vs_empty), want
            continue
        yield name, desired, _edit(desired, edit), want
    for name, edit_from, edit_to, want in STS_BOTH_CASES:
        yield name, _edit(desired, edit_from or (lambda d: None)), _edit(desired, edit_to or (lambda d: None)), want


# Service: L=1 A=2 S=4 Pt=8
SVC_CASES = [
    ("identical", lambda d: None, 0),
    ("clusterIP is deliberately not compared (COMMON:182)", lambda d: d["spec"].update({"clusterIP": "10.0.0.7", "type": "NodePort"}), 0),
    ("targetPort Int(8888) vs Str('8888') (IntOrString compared on Type,IntVal,StrVal)",
     lambda d: d["spec"]["ports"][0].update({"targetPort": "8888"}), 8),
    ("nodePort set on live", lambda d: d["spec"]["ports"][0].update({"nodePort": 30080}), 8),
    ("appProtocol pointer set", lambda d: d["spec"]["ports"][0].update({"appProtocol": "http"}), 8),
    ("second port", lambda d: d["spec"]["ports"].append({"name": "x", "port": 81}), 8),
    ("ports [] vs one port", lambda d: d["spec"].update({"ports": []}), 8),
    ("selector value changed", lambda d: d["spec"]["selector"].update({"statefulset": "other"}), 4),
    ("selector nil vs non-nil", lambda d: d["spec"].pop("selector"), 4),
    ("selector {} vs one entry", lambda d: d["spec"].update({"selector": {}}), 4),
    ("live label", lambda d: d["metadata"].update({"labels": {"a": "b"}}), 1),
    ("live label with empty value", lambda d: d["metadata"].update({"labels": {"a": ""}}), 0),
    ("live annotation", lambda d: d["metadata"].update({"annotations": {"a": "b"}}), 2),
]

# Services whose port lists are longer than the one port the tile spells out: ports[1:] are folded into one digest
# (csrc/host/pack.cpp: pack_service_ports), which must still make equal lists equal and everything else different.
# (name, desired ports, live ports, expected reason)
_P = lambda name, port, **kw: dict({"name": name, "port": port, "protocol": "TCP", "targetPort": port}, **kw)
SVC_PORT_LISTS = [
    ("two equal ports", [_P("a", 80), _P("b", 81)], [_P("a", 80), _P("b", 81)], 0),
    ("three equal ports", [_P("a", 80), _P("b", 81), _P("c", 82)], [_P("a", 80), _P("b", 81), _P("c", 82)], 0),
    ("second port: other name", [_P("a", 80), _P("b", 81)], [_P("a", 80), _P("x", 81)], 8),
    ("second port: other number", [_P("a", 80), _P("b", 81)], [_P("a", 80), _P("b", 82)], 8),
    ("second port: nodePort set", [_P("a", 80), _P("b", 81)], [_P("a", 80), _P("b", 81, nodePort=30081)], 8),
    ("second port: targetPort as a string", [_P("a", 80), _P("b", 81)], [_P("a", 80), _P("b", 81, targetPort="81")], 8),
    ("second port: appProtocol", [_P("a", 80), _P("b", 81)], [_P("a", 80), _P("b", 81, appProtocol="http")], 8),
    ("third port differs", [_P("a", 80), _P("b", 81), _P("c", 82)], [_P("a", 80), _P("b", 81), _P("c", 83)], 8),
    ("third port missing", [_P("a", 80), _P("b", 81), _P("c", 82)], [_P("a", 80), _P("b", 81)], 8),
    ("same ports, other order", [_P("a", 80), _P("b", 81)], [_P("b", 81), _P("a", 80)], 8),
    ("first port differs, rest equal", [_P("a", 80), _P("b", 81)], [_P("a", 90), _P("b", 81)], 8),
    ("one port against the same port twice", [_P("a", 80)], [_P("a", 80), _P("a", 80)], 8),
    ("second port equal to the first on both sides", [_P("a", 80), _P("a", 80)], [_P("a", 80), _P("a", 80)], 0),
    ("nil against empty", None, [], 8),
    ("nil against nil", None, None, 0),
    ("empty against empty", [], [], 0),
]


# VirtualService: 1 differs (true), 2 to.spec missing (true), 4 from.spec missing (false)
VS_CASES = [
    ("identical", lambda f, t: None, 0),
    ("number int64(80) vs float64(80): dynamic types differ",
     lambda f, t: t["spec"]["http"][0]["route"][0]["destination"]["port"].update({"number": 80.0}), 1),
    ("to.spec absent: copied, true (COMMON:209-212)", lambda f, t: t.pop("spec"), 2),
    ("to.spec null: found but not a map => err => copied, true", lambda f, t: t.update({"spec": None}), 2),
    ("to.spec is a string", lambda f, t: t.update({"spec": "x"}), 2),
    ("from.spec absent: false (COMMON:200-203)", lambda f, t: f.pop("spec"), 4),
    ("from.spec not a map: false (COMMON:204-206)", lambda f, t: f.update({"spec": [1]}), 4),
    ("labels/annotations are not compared", lambda f, t: t["metadata"].update({"labels": {"a": "b"}}), 0),
    ("header added", lambda f, t: t["spec"]["http"][0]["headers"]["request"]["set"].update({"X-A": "1"}), 1),
    ("headers.request.set {} vs absent", lambda f, t: t["spec"]["http"][0]["headers"]["request"].pop("set"), 1),
    ("hosts [] vs ['*']", lambda f, t: t["spec"].update({"hosts": []}), 1),
    ("hosts [''] vs ['*']", lambda f, t: t["spec"].update({"hosts": [""]}), 1),
    ("extra key with null value", lambda f, t: t["spec"].update({"tls": None}), 1),
    ("key order is irrelevant", lambda f, t: t.update({"spec": dict(reversed(list(t["spec"].items())))}), 0),
    ("rewrite uri changed", lambda f, t: t["spec"]["http"][0]["rewrite"].update({"uri": "/"}), 1),
    ("number as string", lambda f, t: t["spec"]["http"][0]["route"][0]["destination"]["port"].update({"number": "80"}), 1),
    ("second host", lambda f, t: t["spec"]["hosts"].append("b"), 1),
]

LA = "notebooks.kubeflow.org/last-activity"
CHK = "notebooks.kubeflow.org/last_activity_check_timestamp"
STOP = "kubeflow-resource-stopped"


def _nb(ann, kernels=(), terminals=()):
    return {"annotations": ann, "kernels": [dict(k) for k in kernels], "terminals": [dict(t) for t in terminals]}


def _k(t, state="idle"):
    return {"last_activity": t, "execution_state": state}


# (name, notebook, cull_idle_min, period_min, mode, expected: dict of src/argmax/period/cull/new_la/flags)
CULL_CASES = [
    ("two kernels with the same maximum: strict After keeps the first (CULL:344)",
     _nb({LA: "2025-12-31T00:00:00Z", CHK: "2025-12-31T23:00:00Z"},
         [_k("2025-12-31T10:00:00.5Z"), _k("2025-12-31T10:00:00.5Z")]), 1440, 1, 0,
     {"src": 2, "argmax": 0, "new_la": "2025-12-31T10:00:00Z"}),
    ("kernel max :36.99 vs annotation :36 — Format drops the fraction, equal => update (CULL:348,364)",
     _nb({LA: "2025-12-31T15:37:36Z", CHK: "2025-12-31T23:00:00Z"}, [_k("2025-12-31T15:37:36.990063Z")]), 1440, 1, 0,
     {"src": 2, "argmax": 0, "new_la": "2025-12-31T15:37:36Z"}),
    ("annotation :36.5 vs kernel :36.9 — truncated 36 < 36.5 => no update",
     _nb({LA: "2025-12-31T15:37:36.5Z", CHK: "2025-12-31T23:00:00Z"}, [_k("2025-12-31T15:37:36.9Z")]), 1440, 1, 0,
     {"src": 0}),
    ("a busy kernel => annotation := now at second precision (CULL:378-387)",
     _nb({LA: "2025-12-31T00:00:00Z", CHK: "2025-12-31T23:00:00Z"}, [_k("2025-12-31T10:00:00Z"), _k("bad", "busy")]),
     1440, 1, 0, {"src": 1, "new_la": "2026-01-01T00:00:00Z", "cull": False}),
    ("'starting' is not idle either", _nb({LA: "2025-12-31T00:00:00Z"}, [_k("2025-12-31T10:00:00Z", "starting")]),
     1440, 1, 0, {"src": 1}),
    ("one unparseable kernel time => '' => no update (CULL:332-341)",
     _nb({LA: "2025-12-31T00:00:00Z"}, [_k("2025-12-31T10:00:00Z"), _k("nope")]), 1440, 1, 0, {"src": 0}),
    ("unparseable annotation => compare fails => no update (CULL:353-357)",
     _nb({LA: "garbage"}, [_k("2025-12-31T10:00:00Z")]), 1440, 1, 0, {"src": 0, "cull": False}),
    ("terminals run after kernels on the updated annotation (CULL:178-179)",
     _nb({LA: "2025-12-31T00:00:00Z"}, [_k("2025-12-31T10:00:00Z")], [{"last_activity": "2025-12-31T11:00:00+01:00"}]),
     1440, 1, 0, {"src": 3, "argmax": 0, "new_la": "2025-12-31T11:00:00+01:00"}),
    ("terminal older than the kernel-updated annotation => kernel value stays",
     _nb({LA: "2025-12-31T00:00:00Z"}, [_k("2025-12-31T10:00:00Z")], [{"last_activity": "2025-12-31T09:00:00Z"}]),
     1440, 1, 0, {"src": 2, "argmax": 0, "new_la": "2025-12-31T10:00:00Z"}),
    ("zone is kept by Format: winner in -05:00",
     _nb({LA: "2025-12-31T00:00:00Z"}, [_k("2025-12-31T01:00:00Z"), _k("2025-12-31T06:00:00.25-05:00")]), 1440, 1, 0,
     {"src": 2, "argmax": 1, "new_la": "2025-12-31T06:00:00-05:00"}),
    ("now == last + idle exactly: After is strict => not idle (CULL:236)",
     _nb({LA: "2025-12-31T23:55:00.123456789Z"}), 5, 1, 0, {"cull": False}),
    ("one nanosecond later => idle", _nb({LA: "2025-12-31T23:55:00.123456788Z"}), 5, 1, 0, {"cull": True}),
    ("stop key present with empty value => not idle (CULL:223-226)",
     _nb({LA: "2020-01-01T00:00:00Z", STOP: ""}), 5, 1, 0, {"cull": False}),
    ("annotations nil => not idle (CULL:222)", _nb(None), 5, 1, 0, {"cull": False, "period": False}),
    ("check timestamp absent => period not passed (CULL:208-211)", _nb({LA: "2020-01-01T00:00:00Z"}), 5, 1, 0,
     {"period": False, "cull": True}),
    ("unparseable check timestamp => zero Time => period passed (CULL:212)",
     _nb({LA: "2025-12-31T23:59:00Z", CHK: "junk"}), 1440, 1, 0, {"period": True}),
    ("check + period == now exactly: Before is strict => not passed",
     _nb({CHK: "2025-12-31T23:59:00.123456789Z"}), 1440, 1, 0, {"period": False}),
    ("check + period one ns before now => passed", _nb({CHK: "2025-12-31T23:59:00.123456788Z"}), 1440, 1, 0,
     {"period": True}),
    # ---- the Reconcile flow (mode 1)
    ("flow: stop set at entry => annotations removed, nothing else (CULL:104-117)",
     _nb({LA: "2020-01-01T00:00:00Z", CHK: "2020-01-01T00:00:00Z", STOP: "x"}, [_k("2025-12-31T10:00:00Z", "busy")]),
     5, 1, 1, {"action": 16}),
    ("flow: annotations missing => initialised to now, period not passed (CULL:141-159)",
     _nb({}, [_k("2025-12-31T10:00:00Z")]), 5, 1, 1, {"action": 32}),
    ("flow: period not passed => predicates not evaluated",
     _nb({LA: "2020-01-01T00:00:00Z", CHK: "2025-12-31T23:59:30Z"}, [_k("2025-12-31T10:00:00Z", "busy")]), 5, 1, 1,
     {"action": 0}),
    ("flow: period passed, kernel activity within the idle window => updated, not culled",
     _nb({LA: "2025-12-30T00:00:00Z", CHK: "2025-12-31T23:58:00Z"}, [_k("2025-12-31T01:00:00Z")]), 1440, 1, 1,
     {"src": 2, "period": True, "cull": False}),
    ("flow: period passed, stale => cull bit",
     _nb({LA: "2025-12-29T00:00:00Z", CHK: "2025-12-31T23:58:00Z"}, [_k("2025-12-29T01:00:00Z")]), 1440, 1, 1,
     {"src": 2, "period": True, "cull": True, "new_la": "2025-12-29T01:00:00Z"}),
    ("flow: period 0 and annotations missing => init then everything in one pass",
     _nb({}, [_k("2025-12-31T10:00:00Z")]), 0, 0, 1, {"action": 32 | 4 | 8}),
]

# strings for the RFC3339 parser: (text, expected (sec, nsec, offset) or None).  The first
# block is what the reference's tests feed through time.Parse (culling_controller_test.go:184-248).
RFC3339_CASES = [
    ("should-fail", None),
    ("", None),
    ("2021-08-30T15:37:36.990063Z", (1630337856, 990063000, 0)),
    ("1900-08-30T15:37:36.990063Z", (-2188110144, 990063000, 0)),
    ("2026-01-01T00:00:00Z", (1767225600, 0, 0)),
    ("2026-01-01T00:00:00+00:00", (1767225600, 0, 0)),
    ("2026-01-01T01:00:00+01:00", (1767225600, 0, 3600)),
    ("2025-12-31T19:00:00-05:00", (1767225600, 0, -18000)),
    ("2026-01-01T00:00:00.123456789Z", (1767225600, 123456789, 0)),
    ("2026-01-01T00:00:00.1234567891234Z", (1767225600, 123456789, 0)),
    ("2026-01-01T00:00:00,5Z", (1767225600, 500000000, 0)),
    ("2026-01-01T0:00:00Z", (1767225600, 0, 0)),
    ("2026-01-01t00:00:00Z", None),
    ("2026-01-01T00:00:00z", None),
    ("2026-01-01 00:00:00Z", None),
    ("2026-01-01T00:00:00", None),
    ("2026-01-01T00:00:60Z", None),
    ("2026-01-01T24:00:00Z", None),
    ("2026-02-29T00:00:00Z", None),
    ("2024-02-29T00:00:00Z", (1709164800, 0, 0)),
    ("2026-13-01T00:00:00Z", None),
    ("2026-00-10T00:00:00Z", None),
    ("2026-01-00T00:00:00Z", None),
    ("2026-1-01T00:00:00Z", None),
    ("2026-01-01T00:00:00.Z", None),
    ("2026-01-01T00:00:00+0100", None),
    ("2026-01-01T00:00:00+01", None),
    ("2026-01-02T00:00:00+24:00", (1767225600, 0, 86400)),
    ("2026-01-01T00:00:00+25:00", None),
    ("2026-01-01T00:00:00Zjunk", None),
    ("0001-01-01T00:00:00Z", (-62135596800, 0, 0)),
    ("9999-12-31T23:59:59Z", (253402300799, 0, 0)),
]

# resource.Quantity: text -> (value, scale, format, s) for the int64 fast path
QUANTITY_CASES = [
    ("0", (0, 0, "", "0")),
    ("1", (1, 0, "DecimalSI", "1")),
    ("500m", (500, -3, "DecimalSI", "500m")),
    ("0.5", (5, -1, "DecimalSI", "")),
    ("0.2", (2, -1, "DecimalSI", "")),
    ("1000m", (1000, -3, "DecimalSI", "")),
    ("100m", (100, -3, "DecimalSI", "100m")),
    ("400m", (400, -3, "DecimalSI", "400m")),
    ("64Mi", (67108864, 0, "BinarySI", "")),
    ("1Gi", (1073741824, 0, "BinarySI", "1Gi")),
    ("1024Mi", (1073741824, 0, "BinarySI", "")),
    ("256Mi", (268435456, 0, "BinarySI", "")),
    ("128Mi", (134217728, 0, "BinarySI", "")),
    ("2", (2, 0, "DecimalSI", "2")),
    ("1k", (1, 3, "DecimalSI", "1k")),
    ("1e3", (1, 3, "DecimalExponent", "1e3")),
]
