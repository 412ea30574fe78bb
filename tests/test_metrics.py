"""SURVEY 8(f) rank 4 (first half): Metrics.scrape (pkg/metrics/metrics.go:82-98) — the
notebook_running gauge — as a segmented count on the GPU."""
import collections

import numpy as np
import pytest

import oracle as O

host, pkg = O.host, O.pkg
h64 = pkg.engine.hash64


def _pack(stss, ns_ids):
    """host-side packing of the scrape inputs from StatefulSet objects"""
    ns, nm, lb, fl = [], [], [], []
    for s in stss:
        d = s.to_python()
        labels = d.get("spec", {}).get("template", {}).get("metadata", {}).get("labels")
        present = labels is not None and "notebook-name" in labels
        ns.append(ns_ids.setdefault(d["metadata"].get("namespace", ""), len(ns_ids)))
        nm.append(h64(d["metadata"].get("name", "")))
        lb.append(h64(labels["notebook-name"]) if present else 0)
        fl.append(1 if present else 0)
    return np.array(ns, np.uint32), np.array(nm, np.uint64), np.array(lb, np.uint64), np.array(fl, np.uint8)


def _sts(name, ns, labels):
    t = {"metadata": {}} if labels is None else {"metadata": {"labels": labels}}
    return host.StatefulSet({"metadata": {"name": name, "namespace": ns}, "spec": {"replicas": 1, "template": t}})


KNOWN = [
    ("label equals the name", _sts("nb", "a", {"notebook-name": "nb"}), True),
    ("label differs", _sts("nb", "a", {"notebook-name": "other"}), False),
    ("label missing", _sts("nb", "a", {"statefulset": "nb"}), False),
    ("no template labels at all", _sts("nb", "a", None), False),
    ("generated name: nb-xyz with label of the notebook", _sts("nb-x7k2p", "a", {"notebook-name": "a-very-long-notebook-name"}), False),
    ("empty name and present empty label: ok && '' == ''", _sts("", "a", {"notebook-name": ""}), True),
]


@pytest.mark.parametrize("desc,sts,want", KNOWN, ids=[k[0][:50] for k in KNOWN])
def test_known_answers_oracle(desc, sts, want):
    assert O.counts_as_running(sts) == want


@pytest.mark.gpu
def test_known_answers_and_fleet_gpu(ctx):
    ids = {}
    arrs = _pack([k[1] for k in KNOWN], ids)
    counts = ctx.count_running(*arrs, n_namespaces=len(ids))
    assert counts.tolist() == [sum(1 for k in KNOWN if k[2])]
    # a fleet: live StatefulSets in list order, some with a drifted / missing label
    fleet = host.Fleet(host.FLEET_KUBEFLOW, host.SEED_BASE + 13, 4000, 90)
    objs = [o["sts_to"] for o in fleet.objects(0, 4000)]
    for j in range(0, 4000, 9):
        objs[j].set_str("spec.template.metadata.labels.notebook-name", "renamed")
    for j in range(0, 4000, 31):
        objs[j].erase("spec.template.metadata.labels.notebook-name")
    ids = {}
    arrs = _pack(objs, ids)
    counts = ctx.count_running(*arrs, n_namespaces=len(ids))
    want = collections.Counter()
    for s, ns in zip(objs, arrs[0]):
        if O.counts_as_running(s):
            want[int(ns)] += 1
    assert counts.tolist() == [want.get(k, 0) for k in range(len(ids))]
    assert counts.sum() < 4000 and counts.sum() > 3000


@pytest.mark.gpu
def test_count_at_scale_and_hot_namespace(ctx):
    rng = np.random.RandomState(3)
    m, n_ns = 2_000_000, 20_000
    ns = np.minimum((rng.zipf(1.3, m) - 1), n_ns - 1).astype(np.uint32)       # one very hot namespace
    ns[::1000] = n_ns + 5                                                     # out-of-range ids are ignored
    name = rng.randint(1, 1 << 62, m).astype(np.uint64)
    label = name.copy()
    label[rng.rand(m) < 0.2] = 12345
    flags = (rng.rand(m) < 0.9).astype(np.uint8)
    got = ctx.count_running(ns, name, label, flags, n_ns)
    hit = (flags == 1) & (label == name) & (ns < n_ns)
    want = np.bincount(ns[hit], minlength=n_ns).astype(np.uint32)
    assert np.array_equal(got, want)
    assert ctx.count_running(ns[:0], name[:0], label[:0], flags[:0], 7).tolist() == [0] * 7
