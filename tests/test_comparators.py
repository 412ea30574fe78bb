"""the ODH comparator family — CompareNotebookHTTPRoutes,
CompareNotebookNetworkPolicies, CompareNotebookReferenceGrants (labels + spec through
reflect.DeepEqual), the RoleBinding subjects test and the kube-rbac-proxy ConfigMap loop.
Golden objects are transcribed from the reference's own ODH tests
(odh-notebook-controller/controllers/notebook_controller_test.go:69-116,132-150)."""
import copy
import json

import numpy as np
import pytest

import oracle as O

host, N = O.host, O.N

# notebook_controller_test.go:69-116: the HTTPRoute the API server hands back for "test-notebook"
EXPECTED_HTTPROUTE = {
    "metadata": {"name": "nb-default-test-notebook", "namespace": "redhat-ods-applications",
                 "labels": {"notebook-name": "test-notebook", "notebook-namespace": "default"}},
    "spec": {
        "parentRefs": [{"group": "gateway.networking.k8s.io", "kind": "Gateway", "name": "data-science-gateway",
                        "namespace": "openshift-ingress"}],
        "rules": [{"matches": [{"path": {"type": "PathPrefix", "value": "/notebook/default/test-notebook"}}],
                   "backendRefs": [{"group": "", "kind": "Service", "name": "test-notebook", "namespace": "default",
                                    "port": 8888, "weight": 1}]}]},
}
# what NewNotebookHTTPRoute builds (notebook_route.go:50-130): no group/kind/weight defaults
DESIRED_HTTPROUTE = {
    "metadata": EXPECTED_HTTPROUTE["metadata"],
    "spec": {"parentRefs": [{"name": "data-science-gateway", "namespace": "openshift-ingress"}],
             "rules": [{"matches": [{"path": {"type": "PathPrefix", "value": "/notebook/default/test-notebook"}}],
                        "backendRefs": [{"name": "test-notebook", "namespace": "default", "port": 8888}]}]},
}
NETPOL = {  # NewNotebookNetworkPolicy, notebook_network.go:140-175
    "metadata": {"name": "nb-ctrl-np", "namespace": "default"},
    "spec": {"podSelector": {"matchLabels": {"notebook-name": "nb"}},
             "ingress": [{"ports": [{"protocol": "TCP", "port": 8888}],
                          "from": [{"namespaceSelector": {"matchLabels": {"kubernetes.io/metadata.name": "odh"}}}]}],
             "policyTypes": ["Ingress"]},
}
REFGRANT = {"metadata": {"name": "notebook-httproute-access", "namespace": "default", "labels": {"a": "b"}},
            "spec": {"from": [{"group": "gateway.networking.k8s.io", "kind": "HTTPRoute", "namespace": "odh"}],
                     "to": [{"group": "", "kind": "Service"}]}}
ROLEBINDING = {"metadata": {"name": "rb"}, "roleRef": {"kind": "ClusterRole", "name": "r"},
               "subjects": [{"kind": "ServiceAccount", "name": "nb", "namespace": "default"}]}
CONFIGMAP = {"metadata": {"name": "cm", "labels": {"opendatahub.io/managed-by": "workbenches"}}, "data": {"config-file.yaml": "x: 1"}}


def _e(obj, fn):
    o = copy.deepcopy(obj)
    fn(o)
    return o


def _patched_route():
    # notebook_controller_test.go:134: merge patch replaces the rules list
    return _e(EXPECTED_HTTPROUTE, lambda o: o["spec"].update({"rules": [{"backendRefs": [{"name": "foo", "port": 8888}]}]}))


# (type, mode, labels_path, spec_path, desired, found, expected reason: 1 labels differ, 2 spec differs)
CASES = [
    ("HTTPRoute == itself (the test's BeMatchingK8sResource)", "HTTPRoute", 0, "metadata.labels", "spec", EXPECTED_HTTPROUTE, EXPECTED_HTTPROUTE, 0),
    ("HTTPRoute after the manual patch of the test", "HTTPRoute", 0, "metadata.labels", "spec", EXPECTED_HTTPROUTE, _patched_route(), 2),
    ("HTTPRoute built by the controller vs the defaulted live object", "HTTPRoute", 0, "metadata.labels", "spec", DESIRED_HTTPROUTE, EXPECTED_HTTPROUTE, 2),
    ("HTTPRoute label changed", "HTTPRoute", 0, "metadata.labels", "spec", EXPECTED_HTTPROUTE, _e(EXPECTED_HTTPROUTE, lambda o: o["metadata"]["labels"].update({"notebook-name": "x"})), 1),
    ("HTTPRoute labels nil vs {} (DeepEqual)", "HTTPRoute", 0, "metadata.labels", "spec", _e(EXPECTED_HTTPROUTE, lambda o: o["metadata"].pop("labels")), _e(EXPECTED_HTTPROUTE, lambda o: o["metadata"].update({"labels": {}})), 1),
    ("HTTPRoute name/namespace are not compared", "HTTPRoute", 0, "metadata.labels", "spec", EXPECTED_HTTPROUTE, _e(EXPECTED_HTTPROUTE, lambda o: o["metadata"].update({"name": "other", "resourceVersion": "7"})), 0),
    ("HTTPRoute hostnames [] vs absent", "HTTPRoute", 0, "metadata.labels", "spec", EXPECTED_HTTPROUTE, _e(EXPECTED_HTTPROUTE, lambda o: o["spec"].update({"hostnames": []})), 2),
    ("HTTPRoute weight pointer 0 vs nil", "HTTPRoute", 0, "metadata.labels", "spec", DESIRED_HTTPROUTE, _e(DESIRED_HTTPROUTE, lambda o: o["spec"]["rules"][0]["backendRefs"][0].update({"weight": 0})), 2),
    ("NetworkPolicy == itself", "NetworkPolicy", 0, "metadata.labels", "spec", NETPOL, NETPOL, 0),
    ("NetworkPolicy port changed", "NetworkPolicy", 0, "metadata.labels", "spec", NETPOL, _e(NETPOL, lambda o: o["spec"]["ingress"][0]["ports"][0].update({"port": 8443})), 2),
    ("NetworkPolicy port Int vs Str", "NetworkPolicy", 0, "metadata.labels", "spec", NETPOL, _e(NETPOL, lambda o: o["spec"]["ingress"][0]["ports"][0].update({"port": "8888"})), 2),
    ("NetworkPolicy policyTypes defaulted with Egress", "NetworkPolicy", 0, "metadata.labels", "spec", NETPOL, _e(NETPOL, lambda o: o["spec"]["policyTypes"].append("Egress")), 2),
    ("NetworkPolicy empty podSelector {} == absent (zero struct)", "NetworkPolicy", 0, "metadata.labels", "spec", _e(NETPOL, lambda o: o["spec"].update({"podSelector": {}})), _e(NETPOL, lambda o: o["spec"].pop("podSelector")), 0),
    ("ReferenceGrant == itself", "ReferenceGrant", 0, "metadata.labels", "spec", REFGRANT, REFGRANT, 0),
    ("ReferenceGrant second from", "ReferenceGrant", 0, "metadata.labels", "spec", REFGRANT, _e(REFGRANT, lambda o: o["spec"]["from"].append({"group": "g", "kind": "k", "namespace": "n"})), 2),
    ("ReferenceGrant label value", "ReferenceGrant", 0, "metadata.labels", "spec", REFGRANT, _e(REFGRANT, lambda o: o["metadata"]["labels"].update({"a": "c"})), 1),
    ("RoleBinding subjects equal", "RoleBinding", 0, "metadata.none", "subjects", ROLEBINDING, _e(ROLEBINDING, lambda o: o["metadata"].update({"labels": {"x": "y"}})), 0),
    ("RoleBinding subject namespace differs", "RoleBinding", 0, "metadata.none", "subjects", ROLEBINDING, _e(ROLEBINDING, lambda o: o["subjects"][0].update({"namespace": "other"})), 2),
    ("RoleBinding subjects nil vs []", "RoleBinding", 0, "metadata.none", "subjects", _e(ROLEBINDING, lambda o: o.pop("subjects")), _e(ROLEBINDING, lambda o: o.update({"subjects": []})), 2),
    ("ConfigMap equal", "ConfigMap", 1, "metadata.labels", "data", CONFIGMAP, CONFIGMAP, 0),
    ("ConfigMap data value differs", "ConfigMap", 1, "metadata.labels", "data", CONFIGMAP, _e(CONFIGMAP, lambda o: o["data"].update({"config-file.yaml": "x: 2"})), 2),
    ("ConfigMap extra found key: lengths differ", "ConfigMap", 1, "metadata.labels", "data", CONFIGMAP, _e(CONFIGMAP, lambda o: o["data"].update({"k": "v"})), 2),
    ("ConfigMap label differs", "ConfigMap", 1, "metadata.labels", "data", CONFIGMAP, _e(CONFIGMAP, lambda o: o["metadata"]["labels"].update({"opendatahub.io/managed-by": "x"})), 1),
    ("ConfigMap nil data == {} data (len 0 both)", "ConfigMap", 1, "metadata.labels", "data", _e(CONFIGMAP, lambda o: o.pop("data")), _e(CONFIGMAP, lambda o: o.update({"data": {}})), 0),
    ("ConfigMap desired '' vs found other key: missing reads as '' (quirk)", "ConfigMap", 1, "metadata.labels", "data",
     _e(CONFIGMAP, lambda o: o.update({"data": {"a": ""}})), _e(CONFIGMAP, lambda o: o.update({"data": {"b": "x"}})), 0),
]


def _objs(case):
    _, typ, mode, lp, sp, desired, found, want = case
    return host.Obj.from_json(typ, desired), host.Obj.from_json(typ, found), mode, lp, sp, want


@pytest.mark.parametrize("case", CASES, ids=[c[0][:60] for c in CASES])
def test_comparator_cases_oracles(case):
    d, f, mode, lp, sp, want = _objs(case)
    assert O.compare_generic(d, f, lp, sp, mode) == want
    b = host.GenBatch(mode, lp, sp).add(d, f)
    assert int(O.packed_diff_generic(b)[0][0]) == want


def _fleet_batch(n, kind=host.FLEET_ODH):
    fleet = host.Fleet(kind, host.SEED_BASE + 11, n, max(n // 20, 1))
    objs = fleet.objects(0, n)
    b = host.GenBatch(N.GEN_DEEPEQUAL, "metadata.labels", "spec.template.spec")
    for o in objs:
        b.add(o["sts_from"], o["sts_to"])
    return objs, b


def test_flattened_trees_equal_iff_deep_equal_on_a_fleet():
    """The flattening rule on real PodSpec trees: vectors equal <=> reflect.DeepEqual (L1)."""
    objs, b = _fleet_batch(1200)
    reason, _ = O.packed_diff_generic(b)
    for i, o in enumerate(objs):
        assert O.compare_generic(o["sts_from"], o["sts_to"], "metadata.labels", "spec.template.spec", 0) == reason[i], i
    # cross pairs (different notebooks) must all differ
    b2 = host.GenBatch(N.GEN_DEEPEQUAL, "metadata.labels", "spec.template.spec")
    for i in range(0, 1200, 7):
        b2.add(objs[i]["sts_from"], objs[(i * 13 + 1) % 1200]["sts_to"])
    assert (O.packed_diff_generic(b2)[0] & 2).all()
    assert 0.2 < (reason & 2).astype(bool).mean() < 0.8


@pytest.mark.gpu
@pytest.mark.parametrize("case", CASES, ids=[c[0][:60] for c in CASES])
def test_comparator_cases_gpu(ctx, case):
    d, f, mode, lp, sp, want = _objs(case)
    reason, bm = ctx.diff_generic(host.GenBatch(mode, lp, sp).add(d, f))
    assert int(reason[0]) == want and int(bm[0]) == int(want != 0)


@pytest.mark.gpu
def test_comparator_fleet_gpu(ctx):
    objs, b = _fleet_batch(20000)
    got = ctx.diff_generic(b)
    want = O.packed_diff_generic(b)
    assert np.array_equal(got[0], want[0]) and np.array_equal(got[1], want[1])
    for i in range(0, 20000, 301):
        assert O.compare_generic(objs[i]["sts_from"], objs[i]["sts_to"], "metadata.labels", "spec.template.spec", 0) == got[0][i]
    # the flattened comparison agrees with the slotted StatefulSet kernel on the PodSpec verdict
    fleet = host.Fleet(host.FLEET_ODH, host.SEED_BASE + 11, 20000, 1000)
    sts_reason = ctx.diff_sts(fleet.pack(svc=False, vs=False)[0])[0]
    assert np.array_equal((got[0] & 2) != 0, (sts_reason & 8) != 0)
    # ConfigMap flavour at scale: labels maps of the same fleet
    cm = host.GenBatch(N.GEN_CONFIGMAP, "metadata.labels", "metadata.annotations")
    for o in objs[:5000]:
        cm.add(o["sts_from"], o["sts_to"])
    g, w = ctx.diff_generic(cm), O.packed_diff_generic(cm)
    assert np.array_equal(g[0], w[0]) and np.array_equal(g[1], w[1])


# ---- the desired side from the restated ODH builders (nbh_odh_build) ---------------------------------------
TEST_NB = {"metadata": {"name": "test-notebook", "namespace": "default"},
           "spec": {"template": {"spec": {"containers": [{"name": "test-notebook", "image": "i"}]}}}}
# notebook_controller_test.go:800-838: the NetworkPolicy the ODH test expects for "test-notebook" (controller pod in
# redhat-ods-applications, suite_test.go)
EXPECTED_NETPOL = {"metadata": {"name": "test-notebook-ctrl-np", "namespace": "default"},
                   "spec": {"podSelector": {"matchLabels": {"notebook-name": "test-notebook"}},
                            "ingress": [{"ports": [{"protocol": "TCP", "port": 8888}],
                                         "from": [{"namespaceSelector": {"matchLabels": {"kubernetes.io/metadata.name": "redhat-ods-applications"}}}]}],
                            "policyTypes": ["Ingress"]}}
EXPECTED_PROXY_NETPOL = {"metadata": {"name": "test-notebook-kube-rbac-proxy-np", "namespace": "default"},
                         "spec": {"podSelector": {"matchLabels": {"notebook-name": "test-notebook"}},
                                  "ingress": [{"ports": [{"protocol": "TCP", "port": 8443}]}], "policyTypes": ["Ingress"]}}


def test_odh_builders_give_the_objects_the_reference_tests_expect():
    nb = host.Notebook(TEST_NB)
    route = host.odh_build(nb, "HTTPRoute")
    # the controller's own object (no defaults yet) equals the transcription above; against the object the API server
    # hands back (group / kind / weight defaulted) the reference's comparator already says "different"
    assert O.compare_generic(route, host.Obj.from_json("HTTPRoute", DESIRED_HTTPROUTE), "metadata.labels", "spec", 0) == 0
    assert O.compare_generic(route, host.Obj.from_json("HTTPRoute", EXPECTED_HTTPROUTE), "metadata.labels", "spec", 0) == 2
    assert route.to_python()["metadata"] == EXPECTED_HTTPROUTE["metadata"]
    for what, expected in (("NetworkPolicy", EXPECTED_NETPOL), ("KubeRbacProxyNetworkPolicy", EXPECTED_PROXY_NETPOL)):
        got = host.odh_build(nb, what)
        assert O.compare_generic(got, host.Obj.from_json("NetworkPolicy", expected), "metadata.labels", "spec", 0) == 0, what
        assert got.to_python()["metadata"] == expected["metadata"]
    rg = host.odh_build(nb, "ReferenceGrant")
    want = {"metadata": {"name": "notebook-httproute-access", "namespace": "default",
                         "labels": {"app.kubernetes.io/managed-by": "odh-notebook-controller", "opendatahub.io/component": "notebook-controller"}},
            "spec": {"from": [{"group": "gateway.networking.k8s.io", "kind": "HTTPRoute", "namespace": "redhat-ods-applications"}],
                     "to": [{"group": "", "kind": "Service"}]}}
    assert O.compare_generic(rg, host.Obj.from_json("ReferenceGrant", want), "metadata.labels", "spec", 0) == 0
    rb = host.odh_build(nb, "RoleBinding", rolebinding_name="rb", role_kind="ClusterRole", role_name="r")
    assert O.compare_generic(rb, host.Obj.from_json("RoleBinding", _e(ROLEBINDING, lambda o: o["subjects"][0].update({"name": "test-notebook"}))),
                             "metadata.none", "subjects", 0) == 0
    # a name that does not fit a DNS label switches to generateName (notebook_route.go:68-79)
    long_nb = host.Notebook(_e(TEST_NB, lambda o: o["metadata"].update({"name": "n" * 40, "namespace": "s" * 30})))
    meta = host.odh_build(long_nb, "HTTPRoute").to_python()["metadata"]
    assert "name" not in meta and meta["generateName"] == "nb-" + "s" * 10 + "-" + "n" * 10 + "-"
    # gateway overrides (NOTEBOOK_GATEWAY_NAME / NOTEBOOK_GATEWAY_NAMESPACE)
    pr = host.odh_build(nb, "HTTPRoute", gateway_name="gw", gateway_namespace="gwns").to_python()["spec"]["parentRefs"][0]
    assert pr == {"name": "gw", "namespace": "gwns"}
    with pytest.raises(host.HostError):
        host.odh_build(nb, "Ingress")


def _api_defaults(kind, obj):
    """What the API server adds to a freshly created object of the kind (the defaults the reference's expected objects show)."""
    o = copy.deepcopy(obj)
    if kind == "HTTPRoute":
        for p in o["spec"]["parentRefs"]:
            p.update({"group": "gateway.networking.k8s.io", "kind": "Gateway"})
        for r in o["spec"]["rules"]:
            for b in r["backendRefs"]:
                b.update({"group": "", "kind": "Service", "weight": 1})
    return o


def _odh_pairs(n):
    """(kind, desired, live) for n ODH fleet notebooks: the builders' objects against live copies that are identical,
    defaulted by the API server, or edited by hand (the cases the ODH reconcilers exist for)."""
    fleet = host.Fleet(host.FLEET_ODH, host.SEED_BASE + 71, n, max(n // 25, 1))
    rng = np.random.default_rng(71)
    out = []
    for i in range(n):
        nb = fleet.get(i, "notebook")
        kind = ("HTTPRoute", "NetworkPolicy", "KubeRbacProxyNetworkPolicy", "ReferenceGrant")[i % 4]
        typ = "NetworkPolicy" if "NetworkPolicy" in kind else kind
        desired = host.odh_build(nb, kind)
        live = _as_api(desired.to_python())
        u = rng.random()
        if u < 0.35:
            pass
        elif u < 0.6:
            live = _api_defaults(kind, live)
        elif u < 0.75:
            live["metadata"].setdefault("labels", {})["edited-by"] = "someone"
        elif u < 0.9:
            if "policyTypes" in live["spec"]:
                live["spec"]["policyTypes"] = ["Egress"]                  # the manual patch of notebook_controller_test.go:902
            elif "rules" in live["spec"]:
                live["spec"]["rules"] = [{"backendRefs": [{"name": "foo", "port": 8888}]}]   # ... and of :134
            else:
                live["spec"]["to"].append({"group": "", "kind": "Secret"})
        else:
            live["metadata"]["resourceVersion"] = str(int(rng.integers(1, 10 ** 6)))          # not compared
        out.append((typ, desired, host.Obj.from_json(typ, live)))
    return out


def _as_api(v):
    if isinstance(v, dict):
        if v and set(v) <= {"Type", "IntVal", "StrVal"}:
            return v.get("StrVal", "") if v.get("Type") else v.get("IntVal", 0)
        return {k: _as_api(x) for k, x in v.items()}
    if isinstance(v, list):
        return [_as_api(x) for x in v]
    return v


def test_odh_builder_pairs_packed_equals_semantic():
    pairs = _odh_pairs(1200)
    b = host.GenBatch(N.GEN_DEEPEQUAL, "metadata.labels", "spec")
    for _, d, f in pairs:
        b.add(d, f)
    reason = O.packed_diff_generic(b)[0]
    for i, (_, d, f) in enumerate(pairs):
        assert O.compare_generic(d, f, "metadata.labels", "spec", 0) == reason[i], i
    # every outcome occurs: equal, labels only, spec only
    assert {0, 1, 2} <= set(int(x) for x in reason)


@pytest.mark.gpu
def test_odh_builder_pairs_gpu(ctx):
    pairs = _odh_pairs(6000)
    b = host.GenBatch(N.GEN_DEEPEQUAL, "metadata.labels", "spec")
    for _, d, f in pairs:
        b.add(d, f)
    got, want = ctx.diff_generic(b), O.packed_diff_generic(b)
    assert np.array_equal(got[0], want[0]) and np.array_equal(got[1], want[1])
    for i in range(0, 6000, 7):
        assert O.compare_generic(pairs[i][1], pairs[i][2], "metadata.labels", "spec", 0) == got[0][i]
