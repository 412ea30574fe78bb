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
