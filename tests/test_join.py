"""The owner-reference join (notebook_controller.go:157-170): which live StatefulSet
does each Notebook control?  CPU: semantic vs packed oracle; GPU: the hash-join kernels."""
import copy
import json

import numpy as np
import pytest

import oracle as O

host = O.host
NONE = 0xFFFFFFFF


def _nb(name, ns, uid):
    return host.Notebook({"metadata": {"name": name, "namespace": ns, "uid": uid},
                          "spec": {"template": {"spec": {"containers": [{"name": name}]}}}})


def _sts(name, ns, refs):
    return host.StatefulSet({"metadata": {"name": name, "namespace": ns, "ownerReferences": refs}, "spec": {"replicas": 1}})


def _ref(uid, controller=True, kind="Notebook"):
    r = {"apiVersion": "kubeflow.org/v1beta1", "kind": kind, "name": "x", "uid": uid}
    if controller is not None:
        r["controller"] = controller
    return r


# (description, notebook, StatefulSet list, expected index or -1) — derived from
# metav1.IsControlledBy/GetControllerOf and the loop at NBCTL:165-170
KNOWN = [
    ("single match", ("a", "ns1", "u1"), [("s0", "ns1", [_ref("u1")])], 0),
    ("no StatefulSet yet => create (NBCTL:172)", ("a", "ns1", "u1"), [], -1),
    ("first of two wins (break at NBCTL:168)", ("a", "ns1", "u1"), [("s0", "ns1", [_ref("u9")]), ("s1", "ns1", [_ref("u1")]), ("s2", "ns1", [_ref("u1")])], 1),
    ("owner reference without controller=true is not a controller", ("a", "ns1", "u1"), [("s0", "ns1", [_ref("u1", None)]), ("s1", "ns1", [_ref("u1", False)])], -1),
    ("only the FIRST controller reference counts (GetControllerOf)", ("a", "ns1", "u1"), [("s0", "ns1", [_ref("u2"), _ref("u1")])], -1),
    ("non-controller reference before the controller one is skipped", ("a", "ns1", "u1"), [("s0", "ns1", [_ref("u7", False), _ref("u1")])], 0),
    ("same UID in another namespace is outside the List (InNamespace)", ("a", "ns1", "u1"), [("s0", "ns2", [_ref("u1")])], -1),
    ("matching is by UID, not by name/kind", ("a", "ns1", "u1"), [("a", "ns1", [_ref("u2")]), ("zzz", "ns1", [_ref("u1", True, "Other")])], 1),
    ("no owner references at all", ("a", "ns1", "u1"), [("s0", "ns1", [])], -1),
]


def _keys(nb, stss):
    return np.array([host.join_key_owner(nb)], np.uint64), np.array([host.join_key_controlled(s) for s in stss], np.uint64)


@pytest.mark.parametrize("desc,nb,lst,want", KNOWN, ids=[k[0][:50] for k in KNOWN])
def test_known_answers_oracles(desc, nb, lst, want):
    n = _nb(*nb)
    stss = [_sts(*s) for s in lst]
    assert O.find_controlled(n, stss) == want
    nbk, stk = _keys(n, stss)
    got = int(O.packed_owner_join(nbk, stk)[0][0])
    assert got == (NONE if want < 0 else want)


def _fleet_lists(n, seed=host.SEED_BASE + 6):
    fleet = host.Fleet(host.FLEET_KUBEFLOW, seed, n, max(n // 40, 1))
    nb_keys, sts_keys, owner = fleet.join_lists(n)
    return fleet, nb_keys, sts_keys, owner


def test_packed_join_equals_semantic_on_a_fleet():
    n = 600
    fleet, nb_keys, sts_keys, owner = _fleet_lists(n)
    objs = fleet.objects(0, n)
    # materialise the StatefulSet list as objects
    stss = []
    for j, ow in enumerate(owner):
        if ow >= 0:
            stss.append(objs[ow]["sts_to"])
        else:
            o = objs[j % n]["sts_to"].get("")          # a private copy of some StatefulSet
            if ow == -1:
                o.erase("metadata.ownerReferences")
            else:
                uid = objs[j % n]["notebook"].to_python()["metadata"]["uid"]
                o.set_str("metadata.ownerReferences.0.uid", uid + "-deleted")
            stss.append(o)
    # keys from the objects equal the keys the generator emitted
    match, _ = O.packed_owner_join(nb_keys, sts_keys)
    for i in range(n):
        assert host.join_key_owner(objs[i]["notebook"]) == nb_keys[i]
        want = O.find_controlled(objs[i]["notebook"], stss)
        assert int(match[i]) == (NONE if want < 0 else want), i
    # the list really contains the interesting classes
    counts = np.bincount(owner[owner >= 0], minlength=n)
    assert (counts == 0).any() and (counts == 2).any() and (owner == -1).any() and (owner == -2).any()
    for j in np.nonzero(owner >= 0)[0][:100]:
        assert host.join_key_controlled(stss[j]) == sts_keys[j]


@pytest.mark.gpu
@pytest.mark.parametrize("desc,nb,lst,want", KNOWN, ids=[k[0][:50] for k in KNOWN])
def test_known_answers_gpu(ctx, desc, nb, lst, want):
    n = _nb(*nb)
    stss = [_sts(*s) for s in lst]
    nbk, stk = _keys(n, stss)
    assert int(ctx.owner_join(nbk, stk)[0]) == (NONE if want < 0 else want)


@pytest.mark.gpu
@pytest.mark.parametrize("n", [1, 33, 1000, 200_000])
def test_fleet_join_gpu(ctx, n):
    fleet, nb_keys, sts_keys, owner = _fleet_lists(n)
    got = ctx.owner_join(nb_keys, sts_keys)
    want, _ = O.packed_owner_join(nb_keys, sts_keys)
    assert np.array_equal(got, want)
    if n >= 1000:
        assert (got == NONE).mean() > 0.01
    # property: every match points at a list entry owned by that notebook, and it is the first one
    hit = got != NONE
    assert np.array_equal(owner[got[hit]], np.nonzero(hit)[0])
    first = np.full(n, -1, np.int64)
    for j in range(len(owner) - 1, -1, -1):
        if owner[j] >= 0:
            first[owner[j]] = j
    assert np.array_equal(np.where(hit, got.astype(np.int64), -1), first)


@pytest.mark.gpu
def test_join_device_api_and_edge_cases(ctx):
    import ctypes as C
    import torch
    N = O.N
    lib = N.nbg()
    fleet, nb_keys, sts_keys, owner = _fleet_lists(50_000, seed=77)
    want, _ = O.packed_owner_join(nb_keys, sts_keys)
    d_nb, d_st = torch.from_numpy(nb_keys.view(np.int64)).cuda(), torch.from_numpy(sts_keys.view(np.int64)).cuda()
    d_match = torch.empty(len(nb_keys), dtype=torch.int32, device="cuda")
    need = lib.nbg_owner_join_scratch_bytes(len(sts_keys))
    scratch = torch.full((need,), 0x5A, dtype=torch.uint8, device="cuda")     # dirty on purpose: the call clears it
    stream = torch.cuda.current_stream().cuda_stream
    for _ in range(2):                                                        # twice over the same dirty table
        rc = lib.nbg_owner_join_device(ctx.handle, 0, d_nb.data_ptr(), len(nb_keys), d_st.data_ptr(), len(sts_keys),
                                       d_match.data_ptr(), scratch.data_ptr(), need, stream)
        assert rc == 0
        torch.cuda.synchronize()
        assert np.array_equal(d_match.cpu().numpy().view(np.uint32), want)
    # too-small scratch and a null pointer are rejected
    assert lib.nbg_owner_join_device(ctx.handle, 0, d_nb.data_ptr(), len(nb_keys), d_st.data_ptr(), len(sts_keys),
                                     d_match.data_ptr(), scratch.data_ptr(), need - 1, stream) == N.NBG_E_INVAL
    assert lib.nbg_owner_join_device(ctx.handle, 0, d_nb.data_ptr(), len(nb_keys), d_st.data_ptr(), len(sts_keys),
                                     d_match.data_ptr(), None, need, stream) == N.NBG_E_INVAL
    # an empty StatefulSet list: everything unmatched
    assert (ctx.owner_join(nb_keys[:100], np.empty(0, np.uint64)) == NONE).all()
    assert ctx.owner_join(np.empty(0, np.uint64), sts_keys).size == 0
