"""CopyDeploymentSetFields (components/common/reconcilehelper/util.go:136-163) — SURVEY 8a row D4.  The reference
defines it next to CopyStatefulSetFields and never calls it; it differs in ONE line: `from.Spec.Replicas !=
to.Spec.Replicas` (:152) compares the pointers, which two separately decoded objects never share, so the function
reports an update (and assigns the pointer) unless both are nil — even when the numbers are equal.  Expected values
below are derived from the cited lines (no reference test exists)."""
import json

import numpy as np
import pytest

import oracle as O

pkg = O.pkg
host, N = pkg.host, pkg._native

BASE = {"apiVersion": "apps/v1", "kind": "Deployment",
        "metadata": {"name": "d", "namespace": "ns", "labels": {"app": "d"}},
        "spec": {"replicas": 2, "selector": {"matchLabels": {"app": "d"}},
                 "template": {"metadata": {"labels": {"app": "d"}},
                              "spec": {"containers": [{"name": "c", "image": "nginx"}]}}}}


def _edit(fn):
    d = json.loads(json.dumps(BASE))
    fn(d)
    return d


CASES = [
    ("identical objects, replicas 2 and 2: the POINTERS differ (:152)", BASE, BASE, N.STS_REPLICAS),
    ("both replicas nil: the only way to be equal", _edit(lambda d: d["spec"].pop("replicas")), _edit(lambda d: d["spec"].pop("replicas")), 0),
    ("desired nil, live 3: no panic, update", _edit(lambda d: d["spec"].pop("replicas")), _edit(lambda d: d["spec"].update(replicas=3)), N.STS_REPLICAS),
    ("live nil", BASE, _edit(lambda d: d["spec"].pop("replicas")), N.STS_REPLICAS),
    ("image differs too", BASE, _edit(lambda d: d["spec"]["template"]["spec"]["containers"][0].update(image="nginx:2")), N.STS_REPLICAS | N.STS_PODSPEC),
    ("live-only label, both nil", _edit(lambda d: d["spec"].pop("replicas")),
     _edit(lambda d: (d["spec"].pop("replicas"), d["metadata"]["labels"].update(extra="x"))), N.STS_LABELS),
    ("live annotation, both nil", _edit(lambda d: d["spec"].pop("replicas")),
     _edit(lambda d: (d["spec"].pop("replicas"), d["metadata"].update(annotations={"a": "b"}))), N.STS_ANNOTATIONS),
    ("template labels differ: not looked at by this function", _edit(lambda d: d["spec"].pop("replicas")),
     _edit(lambda d: (d["spec"].pop("replicas"), d["spec"]["template"]["metadata"]["labels"].update(v="2"))), 0),
]


@pytest.mark.parametrize("name,frm,to,want", CASES, ids=[c[0][:50] for c in CASES])
def test_oracles_on_deployments(name, frm, to, want):
    f, t = host.Deployment(frm), host.Deployment(to)
    before_from = f.digest()
    assert O.packed_diff_deploy(host.StsBatch().add(f, t))[0].tolist() == [want]
    live = t.clone()
    assert O.copy_deployment_fields(f, live) == want
    assert f.digest() == before_from
    # after the call `to` carries from's labels, annotations, replicas (when assigned) and PodSpec
    for path in ("metadata.labels", "metadata.annotations", "spec.template.spec"):
        a, b = f.get(path), live.get(path)
        assert (a is None) == (b is None) and (a is None or a.digest() == b.digest()), path
    if want & N.STS_REPLICAS:
        a, b = f.get("spec.replicas"), live.get("spec.replicas")
        assert (a is None) == (b is None) and (a is None or a.digest() == b.digest())
    # the product's copy half gives the same object
    got = t.clone()
    assert N.nbh().nbh_apply_sts_fields(f._h, got._h, want, 0) == 0
    assert got.digest() == live.digest()


@pytest.fixture(scope="module")
def ctx():
    c = pkg.Context()
    yield c
    c.close()


@pytest.mark.gpu
def test_kernel_on_deployments(ctx):
    b = host.StsBatch()
    for _, frm, to, _ in CASES:
        b.add(host.Deployment(frm), host.Deployment(to))
    reason, bitmap = ctx.diff_deploy(b)
    assert reason.tolist() == [c[3] for c in CASES]
    assert pkg.engine.bitmap_to_bools(bitmap, len(CASES)).tolist() == [c[3] != 0 for c in CASES]
    # the same pairs as StatefulSets: equal numbers are equal, a nil pointer is the panic case
    r2, _ = ctx.diff_sts(b, allow_nil=True)
    assert r2[0] == 0 and (r2[1] & N.STS_NILREPLICAS) and (r2[3] & N.STS_NILREPLICAS)


@pytest.mark.gpu
def test_kernel_on_a_fleet_of_deployment_shaped_pairs(ctx):
    fleet = host.Fleet(host.FLEET_ODH, host.SEED_BASE + 61, 20_001, 300)
    sts, _, _ = fleet.pack(svc=False, vs=False)
    got, want = ctx.diff_deploy(sts), O.packed_diff_deploy(sts)
    assert np.array_equal(got[0], want[0]) and np.array_equal(got[1], want[1])
    assert (got[0] & N.STS_REPLICAS).all()                  # generateStatefulSet always sets replicas: pointers differ
