"""The Notebook shapes the reference ships and tests with (tests/golden/shape_fixtures.json, made by
tests/golden/make_shape_fixtures.py from the reference tree) pushed through the whole path: builders
(NBCTL:433-658) -> drift -> L1 semantic oracle == L2 packed oracle == kernels.  Expected values of the
builder rules are derived by hand from the cited reference lines."""
import json
import os

import numpy as np
import pytest

import oracle as O

pkg = O.pkg
host, N = pkg.host, pkg._native

HERE = os.path.dirname(os.path.abspath(__file__))
FIX = json.load(open(os.path.join(HERE, "golden", "shape_fixtures.json")))
SHAPES = sorted(FIX["notebooks"])


def nb_of(key):
    return json.loads(json.dumps(FIX["notebooks"][key]["object"]))


@pytest.mark.parametrize("key", SHAPES)
def test_builders_on_reference_shapes(key):
    nb = nb_of(key)
    name, ns = nb["metadata"]["name"], nb["metadata"]["namespace"]
    c_in = nb["spec"]["template"]["spec"]["containers"][0]
    sts = host.generate_statefulset(host.Notebook(nb)).to_python()
    assert sts["metadata"]["name"] == name and sts["metadata"]["namespace"] == ns
    assert sts["spec"]["replicas"] == 1                                               # NBCTL:434-437: no stop annotation
    assert sts["metadata"].get("labels", {}) == nb["metadata"].get("labels", {})      # NBCTL:474-477
    assert sts["spec"]["selector"]["matchLabels"] == {"statefulset": name}            # NBCTL:456-460
    want_tmpl = {"statefulset": name, "notebook-name": name, "opendatahub.io/workbenches": "true"}
    want_tmpl.update(nb["metadata"].get("labels", {}))                                # NBCTL:462-466,480-483
    assert sts["spec"]["template"]["metadata"]["labels"] == want_tmpl
    kept = {k: v for k, v in nb["metadata"].get("annotations", {}).items() if "kubectl" not in k and "notebook" not in k}
    assert sts["spec"]["template"]["metadata"].get("annotations", {}) == kept         # NBCTL:486-491
    pod = sts["spec"]["template"]["spec"]
    c0 = pod["containers"][0]
    assert c0["workingDir"] == (c_in.get("workingDir") or "/home/jovyan")             # NBCTL:495-497
    if "ports" in c_in:                                                               # NBCTL:498-506: only when nil
        assert [p["containerPort"] for p in c0["ports"]] == [p["containerPort"] for p in c_in["ports"]]
    else:
        assert c0["ports"] == [{"containerPort": 8888, "name": "notebook-port", "protocol": "TCP"}]
    assert c0["env"][-1] == {"name": "NB_PREFIX", "value": "/notebook/%s/%s" % (ns, name)}   # NBCTL:417-431,508
    assert [e["name"] for e in c0["env"][:-1]] == [e["name"] for e in c_in.get("env", [])]
    assert pod["securityContext"] == {"fsGroup": 100}                                 # NBCTL:514-521
    assert pod.get("serviceAccountName", "") == nb["spec"]["template"]["spec"].get("serviceAccountName", "")
    assert len(pod.get("volumes", [])) == len(nb["spec"]["template"]["spec"].get("volumes", []))

    svc = host.generate_service(host.Notebook(nb)).to_python()                        # NBCTL:525-552
    port = c_in["ports"][0]["containerPort"] if c_in.get("ports") else 8888
    assert svc["spec"]["ports"] == [{"name": "http-notebook", "port": 80, "protocol": "TCP", "targetPort": {"IntVal": port}}]
    assert svc["spec"]["selector"] == {"statefulset": name} and svc["spec"]["type"] == "ClusterIP"

    vs = host.generate_virtualservice(host.Notebook(nb)).to_python()                  # NBCTL:558-658
    assert vs["metadata"]["name"] == "notebook-%s-%s" % (ns, name)
    h = vs["spec"]["http"][0]
    assert h["match"][0]["uri"]["prefix"] == "/notebook/%s/%s/" % (ns, name)
    assert h["rewrite"]["uri"] == "/notebook/%s/%s/" % (ns, name)
    assert h["route"][0]["destination"] == {"host": "%s.%s.svc.cluster.local" % (name, ns), "port": {"number": 80}}


def _drifted_pairs(key):
    """(kind, desired, live, expected reason) over one reference shape: the quirks of COMMON:107-219."""
    nb = nb_of(key)
    name = nb["metadata"]["name"]
    mk = {"sts": lambda: host.generate_statefulset(host.Notebook(nb)), "svc": lambda: host.generate_service(host.Notebook(nb)),
          "vs": lambda: host.generate_virtualservice(host.Notebook(nb))}
    out = [("sts", mk["sts"](), mk["sts"](), 0), ("svc", mk["svc"](), mk["svc"](), 0), ("vs", mk["vs"](), mk["vs"](), 0)]
    live = mk["sts"](); live.set_str("spec.template.spec.containers.0.image", "someone/changed:it")
    out.append(("sts", mk["sts"](), live, N.STS_PODSPEC))
    live = mk["sts"](); live.erase("spec.template.spec.securityContext")                  # API server view without fsGroup
    out.append(("sts", mk["sts"](), live, N.STS_PODSPEC))
    live = mk["sts"](); live.set_str("spec.template.spec.dnsPolicy", "ClusterFirst")      # a defaulted field (SURVEY 8a vi)
    out.append(("sts", mk["sts"](), live, N.STS_PODSPEC))
    stopped = json.loads(json.dumps(nb)); stopped["metadata"].setdefault("annotations", {})["kubeflow-resource-stopped"] = ""
    out.append(("sts", host.generate_statefulset(host.Notebook(stopped)), mk["sts"](), N.STS_REPLICAS))  # 0 vs 1, same template labels
    live = mk["svc"](); live.set_str("spec.selector.statefulset", name + "-old")
    out.append(("svc", mk["svc"](), live, N.SVC_SELECTOR))
    live = mk["svc"](); live.set_str("spec.clusterIP", "10.0.0.7")                        # deliberately not compared (COMMON:182)
    out.append(("svc", mk["svc"](), live, 0))
    live = mk["vs"](); live.set_str("spec.http.0.rewrite.uri", "/elsewhere/")
    out.append(("vs", mk["vs"](), live, N.VS_SPEC_DIFFERS))
    live = mk["vs"](); live.erase("spec")
    out.append(("vs", mk["vs"](), live, N.VS_TO_MISSING))
    return out


def _batches(pairs):
    b = {"sts": host.StsBatch(), "svc": host.SvcBatch(), "vs": host.VsBatch()}
    want = {"sts": [], "svc": [], "vs": []}
    for kind, frm, to, reason in pairs:
        b[kind].add(frm, to)
        want[kind].append(reason)
    return b, want


@pytest.mark.parametrize("key", SHAPES)
def test_oracles_on_reference_shapes(key):
    pairs = _drifted_pairs(key)
    b, want = _batches(pairs)
    # L1: the object-level restatement of Copy*Fields (mutates a clone of `to`, like the Go code)
    l1 = {"sts": O.copy_statefulset_fields, "svc": O.copy_service_fields, "vs": O.copy_virtualservice}
    for kind, frm, to, reason in pairs:
        assert l1[kind](frm, to.clone()) == reason, (key, kind, reason)
    # L2: the packed loops over what the packer made of the same objects
    assert O.packed_diff_sts(b["sts"])[0].tolist() == want["sts"]
    assert O.packed_diff_sts_sparse(b["sts"].sparse())[0].tolist() == want["sts"]
    assert O.packed_diff_svc(b["svc"])[0].tolist() == want["svc"]
    assert O.packed_diff_vs(b["vs"])[0].tolist() == want["vs"]


def test_kernels_example_of_the_idleness_proposal():
    ks = FIX["kernels_example"]["kernels"]
    assert O.all_kernels_are_idle(ks)                                                  # CULL:315-327
    now = (1767225600, 0)
    nb = {"annotations": {"notebooks.kubeflow.org/last-activity": "2021-08-30T13:00:00Z",
                          "notebooks.kubeflow.org/last_activity_check_timestamp": "2021-08-30T13:00:00Z"},
          "kernels": ks, "terminals": []}
    action, argmax, text = O.cull(nb, now, 1440, 1, 1)
    assert action & N.ACT_SRC_MASK == 2 and argmax == 0
    assert text == "2021-08-30T14:02:56Z"                                              # CULL:348: Format drops the fraction
    assert action & N.ACT_CULL                                                         # years idle


@pytest.fixture(scope="module")
def ctx():
    c = pkg.Context()
    yield c
    c.close()


@pytest.mark.gpu
@pytest.mark.parametrize("key", SHAPES)
def test_kernels_on_reference_shapes(ctx, key):
    b, want = _batches(_drifted_pairs(key))
    assert ctx.diff_sts(b["sts"])[0].tolist() == want["sts"]
    assert ctx.diff_sts_sparse(b["sts"].sparse())[0].tolist() == want["sts"]
    assert ctx.diff_svc(b["svc"])[0].tolist() == want["svc"]
    assert ctx.diff_svc_gen(b["svc"].gen())[0].tolist() == want["svc"]
    assert ctx.diff_vs(b["vs"])[0].tolist() == want["vs"]
    assert ctx.diff_vs_gen(b["vs"].gen(), host.vs_gen_consts())[0].tolist() == want["vs"]


@pytest.mark.gpu
def test_kernels_example_through_the_text_path(ctx):
    ks = FIX["kernels_example"]["kernels"]
    nb = {"annotations": {"notebooks.kubeflow.org/last-activity": "2021-08-30T13:00:00Z",
                          "notebooks.kubeflow.org/last_activity_check_timestamp": "2021-08-30T13:00:00Z"},
          "kernels": ks, "terminals": []}
    tb = pkg.engine.CullTextBatch([nb])
    action, argmax = ctx.cull_text(tb, 1767225600, 0, 1440, 1, 1)[:2]
    assert int(action[0]) & N.ACT_SRC_MASK == 2 and int(argmax[0]) == 0 and int(action[0]) & N.ACT_CULL
