"""The packer spends ONE word on the usual form of a Quantity, a container port, a volume mount and a volume
(csrc/host/pack.cpp) and several on the general form.  Exhaustively over small pools that mix both forms: two PodSpecs
that differ in exactly one such element have equal tiles exactly when reflect.DeepEqual calls the elements equal
(object-level oracle), the semantic tiles exactly when equality.Semantic.DeepEqual does, and the sparse tiles agree
with the dense ones."""
import copy
import itertools

import numpy as np
import pytest

import cases
import oracle as O

host, N = O.host, O.N

VOLUMES = [
    {"name": "a"}, {"name": "b"}, {"name": ""},
    {"name": "a", "emptyDir": {}}, {"name": "a", "emptyDir": {"medium": ""}}, {"name": "a", "emptyDir": {"medium": "Memory"}},
    {"name": "a", "emptyDir": {"sizeLimit": "1Gi"}}, {"name": "a", "emptyDir": {"medium": "Memory", "sizeLimit": "1Gi"}},
    {"name": "a", "secret": {"secretName": "s"}}, {"name": "a", "secret": {"secretName": "t"}}, {"name": "b", "secret": {"secretName": "s"}},
    {"name": "a", "secret": {"secretName": "s", "optional": False}}, {"name": "a", "secret": {"secretName": "s", "defaultMode": 420}},
    {"name": "a", "secret": {}}, {"name": "a", "configMap": {"name": "s"}}, {"name": "a", "configMap": {}},
    {"name": "a", "persistentVolumeClaim": {"claimName": "s"}}, {"name": "a", "persistentVolumeClaim": {"claimName": "s", "readOnly": True}},
    {"name": "a", "persistentVolumeClaim": {"claimName": "s", "readOnly": False}},
    {"name": "a", "secret": {"secretName": "s"}, "emptyDir": {}}, {"name": "a", "secret": {"secretName": "s"}, "configMap": {"name": "s"}},
    {"name": "a", "hostPath": {"path": "/p"}}, {"name": "a", "hostPath": {"path": "/q"}}, {"name": "a", "hostPath": {"path": "/p"}, "secret": {"secretName": "s"}},
]
MOUNTS = [
    {}, {"name": "a"}, {"mountPath": "/x"}, {"name": "a", "mountPath": "/x"}, {"name": "a", "mountPath": "/y"}, {"name": "b", "mountPath": "/x"},
    {"name": "a", "mountPath": "/x", "readOnly": True}, {"name": "a", "mountPath": "/x", "readOnly": False},
    {"name": "a", "mountPath": "/x", "subPath": "s"}, {"name": "a", "mountPath": "/x", "subPath": ""},
    {"name": "/x", "mountPath": "a"},
]
PORTS = [
    {}, {"name": "p"}, {"containerPort": 80}, {"name": "p", "containerPort": 80}, {"name": "p", "containerPort": 81},
    {"name": "q", "containerPort": 80}, {"name": "p", "containerPort": 80, "protocol": "TCP"}, {"name": "p", "containerPort": 80, "protocol": "UDP"},
    {"name": "p", "containerPort": 80, "hostPort": 80}, {"name": "p", "containerPort": 80, "hostIP": ""},
]
QUANTITIES = ["0", "1", "1000m", "1.0", "-1", "100m", "0.1", "1Gi", "1024Mi", "1073741824", "511Gi", "512Gi", "549755813888",
              "1e3", "1k", "1000", "8E", "8P", "100n", "100u", "1.5Gi", "12345678901234567890", None]


def _pod(kind, value):
    nb = copy.deepcopy(cases.BASE_NOTEBOOK)
    spec = nb["spec"]["template"]["spec"]
    c0 = spec["containers"][0]
    if kind == "volume":
        spec["volumes"] = [value]
    elif kind == "mount":
        c0["volumeMounts"] = [value]
    elif kind == "port":
        c0["ports"] = [value]
    elif kind == "quantity":
        c0["resources"] = {"requests": {"cpu": value} if value is not None else {}}
    elif kind == "second mount":                         # the same, behind another element: slot positions shift nothing
        c0["volumeMounts"] = [{"name": "first", "mountPath": "/1"}, value]
    elif kind == "sidecar quantity":
        spec["containers"].append({"name": "side", "image": "s", "resources": {"limits": {"memory": value} if value is not None else {}}})
    return spec


@pytest.mark.parametrize("kind,pool", [("volume", VOLUMES), ("mount", MOUNTS), ("second mount", MOUNTS), ("port", PORTS),
                                       ("quantity", QUANTITIES), ("sidecar quantity", QUANTITIES)])
def test_tiles_equal_iff_the_elements_are(kind, pool):
    specs = [_pod(kind, v) for v in pool]
    pods = [host.Obj.from_json("PodSpec", sp) for sp in specs]          # what the object-level oracle walks
    base = host.generate_statefulset(host.Notebook(cases.BASE_NOTEBOOK)).to_python()
    for k in ("spec",):
        base[k] = {"replicas": 1, "template": {"metadata": {"labels": {"statefulset": "nb-a"}}}}
    sets = []
    for sp in specs:                                                     # what the packer packs: the same JSON, inside a StatefulSet
        d = copy.deepcopy(base)
        d["spec"]["template"]["spec"] = copy.deepcopy(sp)
        sets.append(host.StatefulSet(d))
    idx = list(itertools.product(range(len(pool)), repeat=2))
    sts, rb = host.StsBatch(), host.RestartBatch()
    for i, j in idx:
        sts.add(sets[i], sets[j])
        rb.add(sets[i], sets[j], sets[i])
    dense = O.packed_diff_sts(sts)[0]
    sparse = O.packed_diff_sts_sparse(sts.sparse())[0]
    sem = O.packed_restart_block(rb)[0]
    assert np.array_equal(dense, sparse)
    n_equal = 0
    for k, (i, j) in enumerate(idx):
        deq, seq = O.deep_equal(pods[i], pods[j]), O.semantic_equal(pods[i], pods[j])
        assert bool(dense[k] & N.STS_PODSPEC) == (not deq), (kind, pool[i], pool[j])
        assert (int(sem[k]) == 4) == seq, (kind, pool[i], pool[j])
        n_equal += deq
    assert len(pool) <= n_equal < len(idx)               # the diagonal at least, never everything
