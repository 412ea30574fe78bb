"""Size-aware namespace -> shard assignment (nbg_balance_shards) and the lean packing mode.

The partition unit of the reference is the namespace (NBCTL:159: one List per namespace), so a shard map may
move whole namespaces only.  nbg_shard_of is the stateless map; nbg_balance_shards is the table a deployment
computes from its namespace sizes when the fleet is as unequal as a Zipf one."""
import ctypes as C

import numpy as np
import pytest

import oracle as O

host, N = O.host, O.N


def _balance(h, sz, G):
    table = np.empty(len(h), np.uint32)
    load = np.zeros(G, np.uint64)
    h, sz = np.ascontiguousarray(h, np.uint64), np.ascontiguousarray(sz, np.uint64)
    rc = N.nbg().nbg_balance_shards(h.ctypes.data, sz.ctypes.data, len(h), G, table.ctypes.data, load.ctypes.data)
    assert rc == N.NBG_OK
    return table, load


def _lpt_python(h, sz, G):
    """The documented rule, restated: decreasing size (ties: increasing hash), lightest shard, lowest number."""
    order = sorted(range(len(h)), key=lambda k: (-int(sz[k]), int(h[k]), k))
    load = [0] * G
    table = np.empty(len(h), np.uint32)
    for k in order:
        s = min(range(G), key=lambda j: (load[j], j))
        table[k] = s
        load[s] += int(sz[k])
    return table, np.array(load, np.uint64)


@pytest.mark.parametrize("G", [1, 2, 3, 8])
def test_balance_is_the_documented_lpt_rule(G):
    rng = np.random.default_rng(G)
    h = rng.integers(1, 2 ** 63, size=500, dtype=np.uint64)
    sz = (rng.zipf(1.3, size=500) % 10_000).astype(np.uint64)          # heavy tail, zeros and ties included
    table, load = _balance(h, sz, G)
    want_t, want_l = _lpt_python(h, sz, G)
    assert np.array_equal(table, want_t) and np.array_equal(load, want_l)
    assert (table < G).all()
    assert load.sum() == sz.sum()
    assert np.array_equal(load, np.bincount(table, weights=sz.astype(np.float64), minlength=G).astype(np.uint64))
    # the LPT bound: no shard is more than one namespace above the mean
    assert load.max() <= sz.sum() / G + sz.max()


def test_balance_does_not_depend_on_the_order_of_the_arrays():
    rng = np.random.default_rng(3)
    h = rng.permutation(np.arange(1, 401, dtype=np.uint64))           # distinct hashes
    sz = rng.integers(0, 50, size=400).astype(np.uint64)              # many ties in size
    t1, l1 = _balance(h, sz, 4)
    perm = rng.permutation(400)
    t2, l2 = _balance(h[perm], sz[perm], 4)
    assert np.array_equal(t1[perm], t2) and np.array_equal(l1, l2)


def test_balance_bad_arguments_and_empty():
    lib = N.nbg()
    t = np.zeros(4, np.uint32)
    assert lib.nbg_balance_shards(None, None, 0, 4, None, None) == N.NBG_OK          # nothing to place
    assert lib.nbg_balance_shards(None, None, 4, 4, t.ctypes.data, None) == N.NBG_E_INVAL
    h = np.arange(1, 5, dtype=np.uint64)
    assert lib.nbg_balance_shards(h.ctypes.data, h.ctypes.data, 4, 0, t.ctypes.data, None) == N.NBG_E_INVAL


def test_zipf_fleet_hash_shards_against_balanced_shards():
    f = host.Fleet(host.FLEET_KUBEFLOW, host.SEED_BASE + 5, 200_000, 4000)
    h, sz = f.ns_table()
    assert sz.sum() == f.n and len(h) == 4000
    assert int(h[7]) == N.nbg().nbg_hash64(b"ns-000007", 9) == f.ns_hash(next(i for i in range(f.n) if f.get(i, "notebook").to_python()["metadata"]["namespace"] == "ns-000007"))
    for G in (2, 4, 8):
        t_hash, l_hash = f.shard_table(G, balanced=False)
        t_bal, l_bal = f.shard_table(G, balanced=True)
        assert l_hash.sum() == l_bal.sum() == f.n
        floor = max(1.0, float(sz.max()) * G / f.n)                   # the largest namespace cannot be split
        assert l_bal.max() / l_bal.mean() <= max(floor * 1.001, 1.001)
        assert l_bal.max() <= l_hash.max()
        # the index lists are a partition of the fleet, and the hash table gives what nbh_fleet_shard_indices gives
        parts = [f.shard_indices_by_table(t_bal, s) for s in range(G)]
        assert [len(p) for p in parts] == [int(x) for x in l_bal]
        allidx = np.concatenate(parts)
        assert len(np.unique(allidx)) == f.n
        assert np.array_equal(f.shard_indices_by_table(t_hash, G - 1), f.shard_indices(G, G - 1))


def test_lean_pack_keeps_exactly_the_sparse_form():
    f = host.Fleet(host.FLEET_ODH, host.SEED_BASE + 4, 3000, 60)
    idx = np.arange(5, 3000, 7, dtype=np.uint64)
    full, svc_f, vs_f = f.pack(idx=idx)
    lean, svc_l, vs_l = f.pack(idx=idx, lean=True)
    assert lean.lean and not full.lean
    a, b = full.sparse().arrays(), lean.sparse().arrays()
    assert set(a) == set(b) and all(np.array_equal(a[k], b[k]) for k in a)
    for x, y in ((svc_f, svc_l), (vs_f, vs_l)):                        # the small batches are untouched by the mode
        ax, ay = x.arrays(), y.arrays()
        assert all(np.array_equal(ax[k], ay[k]) for k in ax)
    with pytest.raises(host.HostError):
        lean.arrays()
    v = lean.view
    assert not v.pod_from and not v.pod_to                             # no dense tiles behind a lean batch
    # and the packed oracle gives the same verdicts from either
    assert np.array_equal(O.packed_diff_sts_sparse(lean.sparse())[0], O.packed_diff_sts(full)[0])
