"""CPU tests: the oracle against the reference's own known-answer cases
(tests/golden/reference_cases.json) and against the derived cases of cases.py;
the host packer + packed oracle (L2) against the semantic oracle (L1)."""
import importlib
import json
import os
import time

import numpy as np
import pytest

import cases
import oracle as O

host = O.host
GOLDEN = json.load(open(os.path.join(os.path.dirname(__file__), "golden", "reference_cases.json")))


def _resolve_now(meta, now_sec):
    """'now-6m' -> RFC3339 string of now shifted by -6 minutes, as the Go tests build inputs."""
    ann = meta.get("annotations")
    if not ann:
        return meta
    out = {}
    for k, v in ann.items():
        if isinstance(v, str) and v.startswith("now"):
            minutes = int(v[3:-1])
            v = O.format_rfc3339(now_sec + minutes * 60)
        out[k] = v
    return {"annotations": out}


# ------------------------------------------------ reference's own test vectors
@pytest.mark.parametrize("case", GOLDEN["stop_annotation_is_set"]["cases"], ids=lambda c: c["name"])
def test_ref_stop_annotation_is_set(case):
    now = int(time.time())
    assert O.stop_annotation_is_set(_resolve_now(case["meta"], now)) == case["result"]


@pytest.mark.parametrize("case", GOLDEN["all_kernels_are_idle"]["cases"], ids=lambda c: c["name"])
def test_ref_all_kernels_are_idle(case):
    assert O.all_kernels_are_idle(case["kernels"]) == case["result"]


@pytest.mark.parametrize("case", GOLDEN["notebook_is_idle"]["cases"], ids=lambda c: c["name"])
def test_ref_notebook_is_idle(case):
    t = time.time_ns()
    now = (t // 10**9, t % 10**9)
    meta = _resolve_now(case["meta"], now[0])
    assert O.notebook_is_idle(meta, now, case["cull_idle_time"]) == case["result"]
    # and through the packed oracle (host parse + L2 loops)
    b = host.CullBatch().add({"annotations": meta.get("annotations")})
    action = O.packed_cull(b, now, case["cull_idle_time"], 1, 0)[0]
    assert bool(action[0] & 8) == case["result"]


def test_ref_bdd_statefulset_labels():
    g = GOLDEN["bdd_statefulset_labels"]
    sts = host.generate_statefulset(host.Notebook(g["notebook"])).to_python()
    assert sts["metadata"]["labels"] == g["expected_statefulset_labels"]
    assert sts["spec"]["replicas"] == 1
    assert sts["spec"]["template"]["spec"]["containers"][0]["workingDir"] == "/home/jovyan"


# ------------------------------------------------------------- derived cases
def _desired_sts_dict(nb_dict):
    """desired StatefulSet as a JSON dict that decodes back to the same object."""
    return _typed_json(host.generate_statefulset(host.Notebook(nb_dict)))


def _typed_json(obj):
    """Obj -> JSON dict in Kubernetes wire form (quantities / IntOrString re-encoded)."""
    def fix(v):
        if isinstance(v, dict):
            if set(v) <= {"i", "d", "s", "Format"} and ("i" in v or "s" in v or "Format" in v or "d" in v) and _is_quantity(v):
                return _quantity_text(v)
            if set(v) <= {"Type", "IntVal", "StrVal"} and v:
                return v.get("StrVal", "") if v.get("Type") else v.get("IntVal", 0)
            return {k: fix(x) for k, x in v.items()}
        if isinstance(v, list):
            return [fix(x) for x in v]
        return v
    return fix(obj.to_python())


def _is_quantity(v):
    return isinstance(v.get("i", {}), dict) and set(v.get("i", {})) <= {"value", "scale"}


def _quantity_text(v):
    if v.get("s"):
        return v["s"]
    if "d" in v:
        raise ValueError("an inf.Dec quantity has no text this helper can write: build the JSON from the original string")
    val, scale = v.get("i", {}).get("value", 0), v.get("i", {}).get("scale", 0)
    fmt = v.get("Format", "")
    if fmt == "BinarySI":
        for suf, sh in (("Gi", 30), ("Mi", 20), ("Ki", 10)):
            if val % (1 << sh) == 0:
                return "%d%s" % (val >> sh, suf)
        return str(val)
    if scale == -3:
        return "%dm" % val
    if scale == -1:
        return "%d.%d" % (val // 10, val % 10)
    if scale == 0:
        return str(val)
    return "%de%d" % (val, scale)


def test_typed_json_round_trip():
    d = _desired_sts_dict(cases.BASE_NOTEBOOK)
    assert host.StatefulSet(d).digest() == host.generate_statefulset(host.Notebook(cases.BASE_NOTEBOOK)).digest()


STS_PAIRS = list(cases.sts_case_pairs(_desired_sts_dict))


@pytest.mark.parametrize("name,frm,to,want", STS_PAIRS, ids=[p[0][:60] for p in STS_PAIRS])
def test_derived_sts(name, frm, to, want):
    f, t = host.StatefulSet(frm), host.StatefulSet(to)
    batch = host.StsBatch().add(f, t)
    assert int(O.packed_diff_sts(batch)[0][0]) == want           # L2 on the packed pair
    assert O.copy_statefulset_fields(f, t, 1) == want            # L1 on the objects
    # Go side effect: afterwards `to` carries the desired owned fields => a second pass is clean
    assert O.copy_statefulset_fields(f, t, 1) == 0


def test_sts_nil_replicas_is_flagged():
    d = _desired_sts_dict(cases.BASE_NOTEBOOK)
    live = json.loads(json.dumps(d))
    del live["spec"]["replicas"]
    f, t = host.StatefulSet(d), host.StatefulSet(live)
    assert O.copy_statefulset_fields(f, t, 1) & 0x80
    assert int(O.packed_diff_sts(host.StsBatch().add(f, host.StatefulSet(live)))[0][0]) & 0x80


def _svc_pairs():
    desired = _typed_json(host.generate_service(host.Notebook(cases.BASE_NOTEBOOK)))
    for name, edit, want in cases.SVC_CASES:
        yield name, desired, cases._edit(desired, edit), want


SVC_PAIRS = list(_svc_pairs())


@pytest.mark.parametrize("name,frm,to,want", SVC_PAIRS, ids=[p[0][:60] for p in SVC_PAIRS])
def test_derived_svc(name, frm, to, want):
    f, t = host.Service(frm), host.Service(to)
    assert int(O.packed_diff_svc(host.SvcBatch().add(f, t))[0][0]) == want
    assert O.copy_service_fields(f, t) == want
    assert O.copy_service_fields(f, t) == 0


def _svc_port_pairs():
    desired = host.generate_service(host.Notebook(cases.BASE_NOTEBOOK)).to_python()
    for name, fp, tp, want in cases.SVC_PORT_LISTS:
        f, t = json.loads(json.dumps(desired)), json.loads(json.dumps(desired))
        for side, ports in ((f, fp), (t, tp)):
            if ports is None:
                side["spec"].pop("ports")
            else:
                side["spec"]["ports"] = json.loads(json.dumps(ports))
        yield name, f, t, want


SVC_PORT_PAIRS = list(_svc_port_pairs())


@pytest.mark.parametrize("name,frm,to,want", SVC_PORT_PAIRS, ids=[p[0] for p in SVC_PORT_PAIRS])
def test_service_port_lists_longer_than_the_tile_spells_out(name, frm, to, want):
    f, t = host.Service(frm), host.Service(to)
    assert O.copy_service_fields(f, host.Service(to)) == want          # the object-level port decides
    assert int(O.packed_diff_svc(host.SvcBatch().add(f, t))[0][0]) == want


def _vs_pairs():
    desired = host.generate_virtualservice(host.Notebook(cases.BASE_NOTEBOOK)).to_python()
    for name, edit, want in cases.VS_CASES:
        f, t = json.loads(json.dumps(desired)), json.loads(json.dumps(desired))
        edit(f, t)
        yield name, f, t, want


VS_PAIRS = list(_vs_pairs())


@pytest.mark.parametrize("name,frm,to,want", VS_PAIRS, ids=[p[0][:60] for p in VS_PAIRS])
def test_derived_vs(name, frm, to, want):
    f, t = host.Unstructured(frm), host.Unstructured(to)
    assert int(O.packed_diff_vs(host.VsBatch().add(f, t))[0][0]) == want
    assert O.copy_virtualservice(f, t) == want
    if want in (1, 2):
        assert O.copy_virtualservice(f, t) == 0  # spec was copied


def check_cull_expectation(action, argmax, new_la, want):
    if "action" in want:
        assert action == want["action"]
    if "src" in want:
        assert action & 3 == want["src"]
    if "argmax" in want:
        assert argmax == want["argmax"]
    if "period" in want:
        assert bool(action & 4) == want["period"]
    if "cull" in want:
        assert bool(action & 8) == want["cull"]
    if "new_la" in want:
        assert new_la == want["new_la"]


@pytest.mark.parametrize("name,nb,idle,period,mode,want", cases.CULL_CASES, ids=[c[0][:60] for c in cases.CULL_CASES])
def test_derived_cull(name, nb, idle, period, mode, want):
    a1, am1, la1 = O.cull(nb, cases.NOW, idle, period, mode)
    check_cull_expectation(a1, am1, la1, want)
    b = host.CullBatch().add(nb)
    a2, am2, _, _, _ = O.packed_cull(b, cases.NOW, idle, period, mode)
    assert (int(a2[0]), int(am2[0])) == (a1, am1)
    assert b.new_last_activity(0, a2[0], am2[0], cases.NOW[0]) == la1


@pytest.mark.parametrize("text,want", cases.RFC3339_CASES, ids=[c[0] or "empty" for c in cases.RFC3339_CASES])
def test_rfc3339_parsers_agree(text, want):
    assert O.parse_rfc3339(text) == want          # oracle's layout-driven parser
    assert host.parse_rfc3339(text) == want       # product's packer
    if want:
        assert O.format_rfc3339(want[0], want[2]) == host.format_rfc3339(want[0], want[2])
        again = O.parse_rfc3339(O.format_rfc3339(want[0], want[2]))
        assert again == (want[0], 0, want[2])     # Format drops the fraction (CULL:348)


@pytest.mark.parametrize("text,want", cases.QUANTITY_CASES, ids=[c[0] for c in cases.QUANTITY_CASES])
def test_quantity_struct(text, want):
    q = host.Obj.from_json("ResourceRequirements", {"limits": {"x": text}}).to_python()["limits"]["x"]
    got = (q.get("i", {}).get("value", 0), q.get("i", {}).get("scale", 0), q.get("Format", ""), q.get("s", ""))
    assert got == want


# ------------------------------------------- packer + L2 against L1 on fleets
@pytest.mark.parametrize("kind,seed,n", [(host.FLEET_KUBEFLOW, host.SEED_BASE + 2, 1500), (host.FLEET_ODH, host.SEED_BASE + 4, 700)])
def test_fleet_packed_equals_semantic(kind, seed, n):
    fleet = host.Fleet(kind, seed, n, 64)
    sts, svc, vs = fleet.pack()
    r_sts, bm_sts, _ = O.packed_diff_sts(sts, threads=2)
    r_svc, _, _ = O.packed_diff_svc(svc)
    r_vs, bm_vs, _ = O.packed_diff_vs(vs)
    for i in range(n):
        assert O.copy_statefulset_fields(fleet.get(i, "sts_from"), fleet.get(i, "sts_to"), 1) == r_sts[i], i
        assert O.copy_service_fields(fleet.get(i, "svc_from"), fleet.get(i, "svc_to")) == r_svc[i], i
        assert O.copy_virtualservice(fleet.get(i, "vs_from"), fleet.get(i, "vs_to")) == r_vs[i], i
    bits = pkg_bits(bm_sts, n)
    assert np.array_equal(bits, (r_sts & 0x0F) != 0)
    assert np.array_equal(pkg_bits(bm_vs, n), (r_vs & 3) != 0)
    # every drift class shows up with the reason it was built to trigger
    want = {0: {0}, 1: {8}, 2: {1}, 3: {0}, 4: {0}, 5: {4, 20}, 6: {2}, 7: {8}, 8: {8}, 9: {8}}
    for i in range(n):
        assert int(r_sts[i]) in want[fleet.drift_class(i)], (i, fleet.drift_class(i), r_sts[i])


def pkg_bits(bitmap, n):
    return O.pkg.engine.bitmap_to_bools(bitmap, n)


def test_pack_equal_iff_deep_equal():
    """The packing rule: tiles equal <=> reflect.DeepEqual (L1) on every generated pair,
    including pairs of DIFFERENT notebooks."""
    fleet = host.Fleet(host.FLEET_ODH, 99, 400, 8)
    sts, _, _ = fleet.pack(svc=False, vs=False)
    pods = sts.arrays()["pod_to"].reshape(-1, 192)
    frm = sts.arrays()["pod_from"].reshape(-1, 192)
    objs_to = [fleet.get(i, "sts_to") for i in range(fleet.n)]
    objs_from = [fleet.get(i, "sts_from") for i in range(fleet.n)]

    def podspec(o):
        return o.get("spec.template.spec")

    for i in range(0, fleet.n, 7):
        a, b = podspec(objs_from[i]), podspec(objs_to[i])
        assert O.deep_equal(a, b) == bool(np.array_equal(frm[i], pods[i])), i
        j = (i * 31 + 5) % fleet.n
        c = podspec(objs_to[j])
        assert O.deep_equal(b, c) == bool(np.array_equal(pods[i], pods[j])), (i, j)


def test_cull_fleet_packed_equals_semantic():
    cf = host.CullFleet(host.SEED_BASE + 3, 6000)
    b = cf.pack()
    now = (cf.now_sec, cf.now_nsec)
    for mode in (0, 1):
        a2, am2, cb, pb, _ = O.packed_cull(b, now, 1440, 1, mode, threads=3)
        for i in range(cf.n):
            a1, am1, la1 = O.cull(cf.json(i), now, 1440, 1, mode)
            assert (a1, am1) == (int(a2[i]), int(am2[i])), (mode, i, cf.json(i))
            assert la1 == b.new_last_activity(i, a2[i], am2[i], now[0]), (mode, i)
        assert np.array_equal(pkg_bits(cb, cf.n), (a2 & 8) != 0)
        assert np.array_equal(pkg_bits(pb, cf.n), (a2 & 4) != 0)
    # the distribution exercises every branch
    a2 = O.packed_cull(b, now, 1440, 1, 0)[0]
    assert set((a2 & 3).tolist()) == {0, 1, 2, 3}


def test_builders_follow_reference_rules():
    nb = json.loads(json.dumps(cases.BASE_NOTEBOOK))
    nb["metadata"]["annotations"].update({"kubectl.kubernetes.io/x": "1", "notebooks.kubeflow.org/last-activity": "t",
                                          "kubeflow-resource-stopped": ""})
    sts = host.generate_statefulset(host.Notebook(nb)).to_python()
    assert sts["spec"]["replicas"] == 0                                   # key presence, empty value (NBCTL:435)
    t = sts["spec"]["template"]
    assert t["metadata"]["annotations"] == {"owner": "alice", "kubeflow-resource-stopped": ""}  # NBCTL:486-491
    assert t["metadata"]["labels"] == {"app": "nb", "team": "x", "statefulset": "nb-a", "notebook-name": "nb-a",
                                       "opendatahub.io/workbenches": "true"}
    c0 = t["spec"]["containers"][0]
    assert c0["ports"] == [{"containerPort": 8888, "name": "notebook-port", "protocol": "TCP"}]
    assert c0["env"] == [{"name": "NB_PREFIX", "value": "/notebook/team-x/nb-a"}]
    assert t["spec"]["securityContext"] == {"fsGroup": 100}
    no_fs = host.generate_statefulset(host.Notebook(nb), {"ADD_FSGROUP": "false"}).to_python()
    assert "securityContext" not in no_fs["spec"]["template"]["spec"]
    nb2 = json.loads(json.dumps(nb))
    nb2["spec"]["template"]["spec"]["containers"][0]["env"] = [{"name": "NB_PREFIX", "value": "/mine"}]
    nb2["spec"]["template"]["spec"]["containers"][0]["ports"] = [{"containerPort": 9999}]
    sts2 = host.generate_statefulset(host.Notebook(nb2)).to_python()
    assert sts2["spec"]["template"]["spec"]["containers"][0]["env"] == [{"name": "NB_PREFIX", "value": "/mine"}]  # NBCTL:420-425
    svc = host.generate_service(host.Notebook(nb2)).to_python()
    assert svc["spec"]["ports"] == [{"name": "http-notebook", "port": 80, "protocol": "TCP", "targetPort": {"IntVal": 9999}}]
    assert svc["spec"]["selector"] == {"statefulset": "nb-a"}
    nb["metadata"]["annotations"]["notebooks.kubeflow.org/http-headers-request-set"] = '{"X-A": "1"}'
    nb["metadata"]["annotations"]["notebooks.kubeflow.org/http-rewrite-uri"] = "/r/"
    vs = host.generate_virtualservice(host.Notebook(nb), {"CLUSTER_DOMAIN": "c.local"}).to_python()
    h = vs["spec"]["http"][0]
    assert vs["metadata"]["name"] == "notebook-team-x-nb-a"
    assert h["headers"]["request"]["set"] == {"X-A": "1"} and h["rewrite"]["uri"] == "/r/"
    assert h["match"][0]["uri"]["prefix"] == "/notebook/team-x/nb-a/"
    assert h["route"][0]["destination"] == {"host": "nb-a.team-x.svc.c.local", "port": {"number": 80}}
    assert vs["spec"]["hosts"] == ["*"] and vs["spec"]["gateways"] == ["kubeflow/kubeflow-gateway"]
    nb["metadata"]["annotations"]["notebooks.kubeflow.org/http-headers-request-set"] = "{not json"
    vs = host.generate_virtualservice(host.Notebook(nb)).to_python()
    assert vs["spec"]["http"][0]["headers"]["request"]["set"] == {}       # NBCTL:608-611
    with pytest.raises(host.HostError):
        host.generate_statefulset(host.Notebook({"metadata": {"name": "x"}, "spec": {"template": {"spec": {}}}}))


def test_odh_webhook_shape():
    nb = host.Notebook(cases.BASE_NOTEBOOK)
    host.odh_mutate(nb, 1 | 2 | 4 | 8 | 16)
    d = nb.to_python()
    pod = d["spec"]["template"]["spec"]
    assert d["metadata"]["annotations"]["kubeflow-resource-stopped"] == "odh-notebook-controller-lock"  # ODHWH:105-114
    assert [c["name"] for c in pod["containers"]] == ["nb-a", "kube-rbac-proxy"]
    assert {v["name"] for v in pod["volumes"]} == {"ws", "trusted-ca", "runtime-images", "elyra-dsp-details",
                                                   "kube-rbac-proxy-config", "kube-rbac-proxy-tls-certificates"}
    assert pod["serviceAccountName"] == "nb-a"                            # ODHWH:324
    proxy = pod["containers"][1]
    assert len(proxy["args"]) == 11 and proxy["resources"]["limits"]["cpu"]["s"] == "100m"
    env_names = [e["name"] for e in pod["containers"][0]["env"]]
    assert sorted(env_names) == sorted(["PIP_CERT", "REQUESTS_CA_BUNDLE", "SSL_CERT_FILE", "PIPELINES_SSL_SA_CERTS",
                                        "KF_PIPELINES_SSL_SA_CERTS", "GIT_SSL_CAINFO"])
    assert host.generate_statefulset(nb).to_python()["spec"]["replicas"] == 0


def test_ref_kube_rbac_proxy_injection():
    """notebook_webhook_test.go:339-417: what InjectKubeRbacProxy must leave in the Notebook — the shape
    every config #4 object starts from (builders.cpp odh_webhook_mutate, flag 16)."""
    g = GOLDEN["kube_rbac_proxy_injection"]
    nb = host.Notebook(g["notebook"])
    host.odh_mutate(nb, 16)
    pod = nb.to_python()["spec"]["template"]["spec"]
    e = g["expect"]
    proxies = [c for c in pod["containers"] if c["name"] == e["container_name"]]
    assert len(proxies) == 1 and len(pod["containers"]) == 2
    c = proxies[0]
    assert c["imagePullPolicy"] == e["imagePullPolicy"]
    for side in ("requests", "limits"):
        assert {k: _quantity_text(v) for k, v in c["resources"][side].items()} == e[side]      # .String() of each Quantity
    assert [{"name": p["name"], "containerPort": p["containerPort"]} for p in c["ports"]] == e["ports"]
    assert sorted(m["name"] for m in c["volumeMounts"]) == sorted(e["volumeMounts"])
    assert sorted(v["name"] for v in pod["volumes"]) == sorted(e["volumes"])
    assert pod["serviceAccountName"] == e["serviceAccountName"]
    assert all(p in c for p in e["probes"])


def test_ref_kube_rbac_proxy_custom_resources():
    g = GOLDEN["kube_rbac_proxy_custom_resources"]
    nb = host.Notebook(g["notebook"])
    host.odh_mutate(nb, 16)
    pod = nb.to_python()["spec"]["template"]["spec"]
    e = g["expect"]
    assert [c["name"] for c in pod["containers"]] == e["containers"]
    c = pod["containers"][-1]
    for side in ("requests", "limits"):
        assert {k: _quantity_text(v) for k, v in c["resources"][side].items()} == e[side]
    assert sorted(v["name"] for v in pod["volumes"]) == sorted(e["volumes"])


def test_ref_kube_rbac_proxy_validation_failure():
    g = GOLDEN["kube_rbac_proxy_validation_failure"]
    nb = host.Notebook(g["notebook"])
    before = nb.digest()
    with pytest.raises(host.HostError, match=g["expect"]["error"]):
        host.odh_mutate(nb, 16)
    assert nb.digest() == before                                           # left exactly as it was
    pod = nb.to_python()["spec"]["template"]["spec"]
    assert [c["name"] for c in pod["containers"]] == g["expect"]["containers"]
    assert pod.get("serviceAccountName", "") == g["expect"]["serviceAccountName"] and "volumes" not in pod


@pytest.mark.parametrize("ann,ok", [({"cpu-request": " 50m "}, True), ({"cpu-request": "-1"}, False), ({"memory-limit": "sixty"}, False),
                                    ({"memory-request": "65Mi"}, False), ({"memory-request": "64Mi"}, True),
                                    ({"cpu-request": "0.1"}, True), ({"cpu-request": "0.11"}, False), ({"cpu-limit": ""}, True)])
def test_auth_sidecar_resource_rules(ann, ok):
    """parseAndValidateAuthSidecarResources (notebook_webhook.go:124-172): trimmed, empty = absent, no negatives,
    request <= limit by VALUE (0.1 == 100m), defaults 100m / 64Mi fill what is missing."""
    nb = host.Notebook({"metadata": {"name": "n", "namespace": "d", "annotations":
                                      {"notebooks.opendatahub.io/auth-sidecar-" + k: v for k, v in ann.items()}},
                        "spec": {"template": {"spec": {"containers": [{"name": "n", "image": "i"}]}}}})
    if ok:
        host.odh_mutate(nb, 16)
        assert len(nb.to_python()["spec"]["template"]["spec"]["containers"]) == 2
    else:
        with pytest.raises(host.HostError):
            host.odh_mutate(nb, 16)


@pytest.mark.parametrize("kind", [host.FLEET_KUBEFLOW, host.FLEET_ODH])
def test_copy_half_of_copy_fields_matches_the_semantic_oracle(kind):
    """Copy*Fields = compare (kernel) + copy (csrc/host/apply.cpp).  Given the reason bits, the product's copy must
    leave `to` exactly as the object-level restatement of the Go functions leaves it (digest equality <=> DeepEqual)."""
    lib = O.pkg._native.nbh()
    fleet = host.Fleet(kind, host.SEED_BASE + 77, 600, 16)
    for i in range(600):
        for which, orc, apply in (("sts", lambda f, t: O.copy_statefulset_fields(f, t, 1), lambda f, t, r: lib.nbh_apply_sts_fields(f._h, t._h, r, 1)),
                                  ("svc", O.copy_service_fields, lambda f, t, r: lib.nbh_apply_svc_fields(f._h, t._h)),
                                  ("vs", O.copy_virtualservice, lambda f, t, r: lib.nbh_apply_vs_fields(f._h, t._h, r))):
            frm, want_to, got_to = fleet.get(i, which + "_from"), fleet.get(i, which + "_to"), fleet.get(i, which + "_to")
            reason = orc(frm, want_to)                        # mutates want_to like the Go code
            assert apply(frm, got_to, reason) == 0
            assert got_to.digest() == want_to.digest(), (which, i, fleet.drift_class(i), reason)


# ---- Feast config mount and cluster-proxy env injection (config #4 input shape, G4) -------------------------
def _feast_state(nb_obj):
    pod = nb_obj.to_python()["spec"]["template"]["spec"]
    vols = [v for v in pod.get("volumes") or [] if v["name"] == "odh-feast-config"]
    mounts = {c["name"]: [{k: m.get(k, False if k == "readOnly" else None) for k in ("name", "mountPath", "readOnly")}
                          for m in c.get("volumeMounts") or [] if m["name"] == "odh-feast-config"]
              for c in pod["containers"]}
    return pod, vols, mounts


def test_ref_feast_label_gate():
    """notebook_feast_config_test.go:40-105 (isFeastEnabled): only the exact label value "true" mounts."""
    g = GOLDEN["feast_config"]
    for case in g["is_enabled"]:
        body = {"metadata": {"name": "test-notebook", "namespace": "default"},
                "spec": {"template": {"spec": {"containers": [{"name": "test-notebook", "image": "i"}]}}}}
        if case["labels"] is not None:
            body["metadata"]["labels"] = case["labels"]
        nb = host.Notebook(body)
        host.odh_mutate(nb, host.ODH_FEAST)
        _, vols, _ = _feast_state(nb)
        assert bool(vols) == case["want"], case["at"]


@pytest.mark.parametrize("k", range(4))
def test_ref_feast_mount(k):
    """notebook_feast_config_test.go:108-301 (mountFeastConfig), reached the way the webhook reaches it: label "true"
    and the config map named <notebook>-feast-config (NewFeastConfig, notebook_feast_config.go:149-158)."""
    case = GOLDEN["feast_config"]["mount"][k]
    body = json.loads(json.dumps(case["notebook"]))
    body["metadata"]["labels"] = {"opendatahub.io/feast-integration": "true"}
    nb = host.Notebook(body)
    host.odh_mutate(nb, host.ODH_FEAST)
    pod, vols, mounts = _feast_state(nb)
    e = case["expect"]
    assert [v["name"] for v in pod.get("volumes") or []] == e["volumes"], case["at"]
    assert len(vols) == 1 and vols[0]["configMap"]["name"] == e["configMap"]
    assert ("items" not in vols[0]["configMap"]) == e["items_nil"]           # all keys of the ConfigMap, mounted dynamically
    assert mounts == e["mounts"], case["at"]
    assert len(pod["containers"]) == len(case["notebook"]["spec"]["template"]["spec"]["containers"])


@pytest.mark.parametrize("k", range(2))
def test_ref_feast_unmount(k):
    """notebook_feast_config_test.go:304-397 (unmountFeastConfig), reached the way the webhook reaches it: the label
    is not "true" and the volume is still there (notebook_webhook.go:440-444)."""
    case = GOLDEN["feast_config"]["unmount"][k]
    nb = host.Notebook(case["notebook"])
    host.odh_mutate(nb, host.ODH_FEAST)
    pod, vols, mounts = _feast_state(nb)
    assert vols == [] and mounts == case["expect"]["mounts"], case["at"]
    if k == 0:
        # append(s[:i], s[i+1:]...) leaves EMPTY, non-nil slices behind: reflect.DeepEqual tells them from nil ones
        assert pod["volumes"] == [] and pod["containers"][0]["volumeMounts"] == []
    else:
        assert pod.get("volumes") is None                                      # untouched: still nil


def test_proxy_env_injection_quirks():
    """InjectProxyConfigEnvVars (notebook_webhook.go:566-616; no spec of its own in the reference, so these
    expectations are derived from the cited lines): the three variables are appended to the IMAGE container only;
    one that is already present keeps its value, because the update at :583-585 assigns to the range variable."""
    body = {"metadata": {"name": "wb", "namespace": "ns"},
            "spec": {"template": {"spec": {"containers": [
                {"name": "sidecar", "image": "s"},
                {"name": "wb", "image": "i", "env": [{"name": "A", "value": "1"}, {"name": "HTTP_PROXY", "value": "http://mine:1"}]}]}}}}
    nb = host.Notebook(body)
    host.odh_mutate(nb, 0, proxy=("http://p:3128", "http://p:3129", ".svc"))
    cs = nb.to_python()["spec"]["template"]["spec"]["containers"]
    assert "env" not in cs[0] or not cs[0]["env"]
    env = [(e["name"], e["value"]) for e in cs[1]["env"]]
    assert env[:2] == [("A", "1"), ("HTTP_PROXY", "http://mine:1")]             # kept, not updated
    assert sorted(env[2:]) == [("HTTPS_PROXY", "http://p:3129"), ("NO_PROXY", ".svc")]
    # no container carries the notebook's name: nothing happens (the loop at :570 never matches)
    nb2 = host.Notebook({"metadata": {"name": "other", "namespace": "ns"},
                         "spec": {"template": {"spec": {"containers": [{"name": "wb", "image": "i"}]}}}})
    before = nb2.digest()
    host.odh_mutate(nb2, 0, proxy=("a", "b", "c"))
    assert nb2.digest() == before


def test_odh_fleet_carries_feast_and_proxy_shapes():
    """Config #4's fleet: a labelled fraction with the sixth volume, a fraction whose label went back to "false"
    (volume removed again), a fraction admitted under a cluster-wide proxy; packed and object-level oracles agree on
    all of them (the drift on top includes live pods that predate either mutation)."""
    fleet = host.Fleet(host.FLEET_ODH, host.SEED_BASE + 4, 2500, 50)
    feast = off = proxy = 0
    for i in range(0, 2500, 5):
        nb = fleet.get(i, "notebook").to_python()
        lab = (nb["metadata"].get("labels") or {}).get("opendatahub.io/feast-integration")
        pod = nb["spec"]["template"]["spec"]
        has = any(v["name"] == "odh-feast-config" for v in pod.get("volumes") or [])
        assert has == (lab == "true"), i
        feast += has
        off += lab == "false"
        proxy += any(e["name"] == "NO_PROXY" for e in pod["containers"][0].get("env") or [])
    assert 30 < feast < 100 and 5 < off < 35 and 80 < proxy < 170
    sts, svc, vs = fleet.pack()
    l1 = O.l1_diff_all(fleet)
    assert np.array_equal(O.packed_diff_sts(sts)[0], l1["sts"])
    assert np.array_equal(O.packed_diff_vs(vs)[0], l1["vs"])


def test_culler_env_parsing_follows_strconv_atoi(monkeypatch):
    """initGlobalVars (CULL:525-558): Atoi accepts a sign and digits only; a bad CULL_IDLE_TIME falls back to 1440,
    a bad IDLENESS_CHECK_PERIOD is an error; empty means default (GetEnvDefault, CULL:505-511)."""
    culler = importlib.import_module("odh-kubeflow_b200.culler")
    for text, want in (("", 1440), ("30", 30), ("+5", 5), ("-3", -3), (" 5", 1440), ("5 ", 1440), ("1_000", 1440),
                       ("abc", 1440), ("1.5", 1440), ("9223372036854775808", 1440), ("007", 7)):
        monkeypatch.setenv("CULL_IDLE_TIME", text)
        monkeypatch.delenv("IDLENESS_CHECK_PERIOD", raising=False)
        culler.initGlobalVars()
        assert culler.CULL_IDLE_TIME == want and culler.IDLENESS_CHECK_PERIOD == 1, text
    monkeypatch.delenv("CULL_IDLE_TIME", raising=False)
    for text in (" 2", "2m", "1_0", "one"):
        monkeypatch.setenv("IDLENESS_CHECK_PERIOD", text)
        with pytest.raises(ValueError):
            culler.initGlobalVars()
    monkeypatch.setenv("IDLENESS_CHECK_PERIOD", "15")
    culler.initGlobalVars()
    assert (culler.CULL_IDLE_TIME, culler.IDLENESS_CHECK_PERIOD) == (1440, 15)
    monkeypatch.delenv("IDLENESS_CHECK_PERIOD")
    culler.initGlobalVars()


_SIDECAR = GOLDEN["auth_sidecar_resources_table"]


@pytest.mark.parametrize("case", _SIDECAR["cases"], ids=[c["name"] for c in _SIDECAR["cases"]])
def test_ref_auth_sidecar_resources_table(case):
    """auth_proxy_resources_test.go:140-418, all 17 rows: the Quantity spellings the reference itself accepts, refuses,
    trims, compares and prints — reached through InjectKubeRbacProxy, whose proxy container carries the four values."""
    meta = {"name": "test-notebook", "namespace": "default"}
    if case["annotations"] is not None:
        meta["annotations"] = {_SIDECAR["annotation_prefix"] + k: v for k, v in case["annotations"].items()}
    nb = host.Notebook({"metadata": meta, "spec": {"template": {"spec": {"containers": [{"name": "test-notebook", "image": "test-image"}]}}}})
    if case["error"]:
        before = nb.digest()
        with pytest.raises(host.HostError):
            host.odh_mutate(nb, 16)
        assert nb.digest() == before                       # nothing injected (notebook_webhook.go:177-183)
        return
    host.odh_mutate(nb, 16)
    proxy = [c for c in nb.to_python()["spec"]["template"]["spec"]["containers"] if c["name"] == "kube-rbac-proxy"][0]
    got = {"cpu_request": _quantity_text(proxy["resources"]["requests"]["cpu"]), "memory_request": _quantity_text(proxy["resources"]["requests"]["memory"]),
           "cpu_limit": _quantity_text(proxy["resources"]["limits"]["cpu"]), "memory_limit": _quantity_text(proxy["resources"]["limits"]["memory"])}
    assert got == case["expect"]


def test_ref_kube_rbac_proxy_replaces_an_existing_container():
    g = GOLDEN["kube_rbac_proxy_update_existing"]
    nb = host.Notebook(g["notebook"])
    host.odh_mutate(nb, 16)
    pod = nb.to_python()["spec"]["template"]["spec"]
    e = g["expect"]
    assert len(pod["containers"]) == e["containers"]
    c = [x for x in pod["containers"] if x["name"] == "kube-rbac-proxy"][0]
    assert c["image"] != e["old_image"] and c["imagePullPolicy"] == e["imagePullPolicy"]
    for side in ("requests", "limits"):
        assert {k: _quantity_text(v) for k, v in c["resources"][side].items()} == e[side]
    assert e["args_absent"] not in c["args"] and len(c["args"]) > e["args_more_than"]
    assert [{"name": p["name"], "containerPort": p["containerPort"]} for p in c["ports"]] == e["ports"]
    names = {v["name"] for v in pod["volumes"]}
    assert set(e["volumes_include"]) <= names
