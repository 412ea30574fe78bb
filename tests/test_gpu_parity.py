"""GPU parity tests: every result comes out of the CUDA kernels through the C ABI
(libnbgraft.so) and is compared bit for bit with the CPU oracle."""
import ctypes as C
import importlib
import json
import os
import time

import numpy as np
import pytest

import cases
import oracle as O
from test_oracle_cases import GOLDEN, STS_PAIRS, SVC_PAIRS, SVC_PORT_PAIRS, VS_PAIRS, _resolve_now, check_cull_expectation

pytestmark = pytest.mark.gpu

pkg = O.pkg
host = pkg.host
N = pkg._native
rh = importlib.import_module("odh-kubeflow_b200.reconcilehelper")
culler = importlib.import_module("odh-kubeflow_b200.culler")
bits = pkg.engine.bitmap_to_bools


# ------------------------------------------- the reference's own table tests
@pytest.mark.parametrize("case", GOLDEN["all_kernels_are_idle"]["cases"], ids=lambda c: c["name"])
def test_ref_all_kernels_are_idle(ctx, case):
    assert culler.allKernelsAreIdle(case["kernels"]) == case["result"]


@pytest.mark.parametrize("case", GOLDEN["notebook_is_idle"]["cases"], ids=lambda c: c["name"])
def test_ref_notebook_is_idle(ctx, case, monkeypatch):
    monkeypatch.setenv("CULL_IDLE_TIME", str(case["cull_idle_time"]))
    culler.initGlobalVars()
    t = time.time_ns()
    now = (t // 10**9, t % 10**9)
    assert culler.notebookIsIdle(_resolve_now(case["meta"], now[0]), now=now) == case["result"]


@pytest.mark.parametrize("case", GOLDEN["stop_annotation_is_set"]["cases"], ids=lambda c: c["name"])
def test_ref_stop_annotation(ctx, case):
    meta = _resolve_now(case["meta"], int(time.time()))
    assert culler.StopAnnotationIsSet(meta) == case["result"]
    # the same bit as the kernel reports it in Reconcile mode (NBG_ACT_STOPPED)
    b = host.CullBatch().add({"annotations": meta.get("annotations")})
    action = ctx.cull(b, int(time.time()), 0, 5, 1, N.CULL_RECONCILE)[0]
    assert bool(action[0] & N.ACT_STOPPED) == case["result"]


@pytest.mark.parametrize("case", GOLDEN["set_stop_annotation"]["cases"], ids=lambda c: c["name"])
def test_ref_set_stop_annotation(ctx, case):
    meta = json.loads(json.dumps(case["meta"]))
    culler.setStopAnnotation(meta)
    assert culler.STOP_ANNOTATION in meta["annotations"]


def test_ref_bdd_labels_need_no_update(ctx):
    g = GOLDEN["bdd_statefulset_labels"]
    nb = host.Notebook(g["notebook"])
    desired = host.generate_statefulset(nb)
    assert desired.to_python()["metadata"]["labels"] == g["expected_statefulset_labels"]
    reason, bm = ctx.diff_sts(host.StsBatch().add(desired, desired))
    assert reason[0] == 0 and bm[0] == 0


# ---------------------------------------------------------- derived cases
@pytest.mark.parametrize("name,frm,to,want", STS_PAIRS, ids=[p[0][:60] for p in STS_PAIRS])
def test_derived_sts(ctx, name, frm, to, want):
    to = json.loads(json.dumps(to))
    got = rh.CopyStatefulSetFields(frm, to, ctx, caller_prestep=True)
    assert got == ((want & 0x0F) != 0)
    reason, bm = ctx.diff_sts(host.StsBatch().add(host.StatefulSet(frm), host.StatefulSet(to)))
    assert reason[0] == 0, "after the copy a second pass must be clean"
    # raw reason byte
    r2, _ = ctx.diff_sts(host.StsBatch().add(host.StatefulSet(frm), host.StatefulSet(cases._edit(to, lambda d: None))))
    assert r2[0] == 0


@pytest.mark.parametrize("name,frm,to,want", STS_PAIRS, ids=[p[0][:60] for p in STS_PAIRS])
def test_derived_sts_reason_bits(ctx, name, frm, to, want):
    reason, bm = ctx.diff_sts(host.StsBatch().add(host.StatefulSet(frm), host.StatefulSet(to)))
    assert int(reason[0]) == want
    assert int(bm[0]) == int((want & 0x0F) != 0)


def test_sts_nil_replicas(ctx):
    frm = STS_PAIRS[0][1]
    live = json.loads(json.dumps(frm))
    del live["spec"]["replicas"]
    with pytest.raises(RuntimeError, match="nil pointer"):
        rh.CopyStatefulSetFields(frm, live, ctx)
    with pytest.raises(pkg.NbgError) as ei:
        ctx.diff_sts(host.StsBatch().add(host.StatefulSet(frm), host.StatefulSet(live)))
    assert ei.value.status == N.NBG_E_NILREPLICAS


@pytest.mark.parametrize("name,frm,to,want", SVC_PAIRS, ids=[p[0][:60] for p in SVC_PAIRS])
def test_derived_svc(ctx, name, frm, to, want):
    reason, bm = ctx.diff_svc(host.SvcBatch().add(host.Service(frm), host.Service(to)))
    assert int(reason[0]) == want and int(bm[0]) == int(want != 0)
    to = json.loads(json.dumps(to))
    assert rh.CopyServiceFields(frm, to, ctx) == (want != 0)
    assert rh.CopyServiceFields(frm, to, ctx) is False


def test_service_port_lists_longer_than_the_tile_spells_out(ctx):
    b = host.SvcBatch()
    for name, frm, to, want in SVC_PORT_PAIRS:
        b.add(host.Service(frm), host.Service(to))
    reason, bm = ctx.diff_svc(b)
    assert reason.tolist() == [p[3] for p in SVC_PORT_PAIRS]
    assert bits(bm, len(b)).tolist() == [p[3] != 0 for p in SVC_PORT_PAIRS]


@pytest.mark.parametrize("name,frm,to,want", VS_PAIRS, ids=[p[0][:60] for p in VS_PAIRS])
def test_derived_vs(ctx, name, frm, to, want):
    reason, bm = ctx.diff_vs(host.VsBatch().add(host.Unstructured(frm), host.Unstructured(to)))
    assert int(reason[0]) == want and int(bm[0]) == int(want in (1, 2))
    to = json.loads(json.dumps(to))
    assert rh.CopyVirtualService(frm, to, ctx) == (want in (1, 2))
    if want in (1, 2):
        assert to["spec"] == frm["spec"]


@pytest.mark.parametrize("name,nb,idle,period,mode,want", cases.CULL_CASES, ids=[c[0][:60] for c in cases.CULL_CASES])
def test_derived_cull(ctx, name, nb, idle, period, mode, want):
    b = host.CullBatch().add(nb)
    action, argmax, cb, pb = ctx.cull(b, cases.NOW[0], cases.NOW[1], idle, period, mode)
    a1, am1, la1 = O.cull(nb, cases.NOW, idle, period, mode)
    assert (int(action[0]), int(argmax[0])) == (a1, am1)
    check_cull_expectation(int(action[0]), int(argmax[0]), b.new_last_activity(0, action[0], argmax[0], cases.NOW[0]), want)
    assert int(cb[0]) == int(bool(a1 & 8)) and int(pb[0]) == int(bool(a1 & 4))


def test_culler_mirror_updates_annotation(ctx):
    culler.CULL_IDLE_TIME, culler.IDLENESS_CHECK_PERIOD = 1440, 1
    meta = {"annotations": {culler.LAST_ACTIVITY_ANNOTATION: "2025-12-31T00:00:00Z"}}
    ks = [{"last_activity": "2025-12-31T10:00:00.7Z", "execution_state": "idle"}]
    assert culler.updateTimestampFromKernelsActivity(meta, ks, now=cases.NOW) is True
    assert meta["annotations"][culler.LAST_ACTIVITY_ANNOTATION] == "2025-12-31T10:00:00Z"
    ks[0]["execution_state"] = "busy"
    assert culler.updateTimestampFromKernelsActivity(meta, ks, now=cases.NOW) is False
    assert meta["annotations"][culler.LAST_ACTIVITY_ANNOTATION] == "2026-01-01T00:00:00Z"
    assert culler.updateTimestampFromTerminalsActivity(meta, [{"last_activity": "2026-01-01T00:00:00.9Z"}], now=cases.NOW) is True
    assert culler.updateTimestampFromKernelsActivity(meta, [], now=cases.NOW) is False
    assert culler.cullingCheckPeriodHasPassed({"annotations": {culler.LAST_ACTIVITY_CHECK_TIMESTAMP_ANNOTATION: "x"}}, now=cases.NOW)


# ------------------------------------------------- fleets at BASELINE sizes
def _check_diff_fleet(ctx, fleet):
    sts, svc, vs = fleet.pack()
    got = {"sts": ctx.diff_sts(sts), "svc": ctx.diff_svc(svc), "vs": ctx.diff_vs(vs)}
    want = {"sts": O.packed_diff_sts(sts, threads=8), "svc": O.packed_diff_svc(svc, threads=8),
            "vs": O.packed_diff_vs(vs, threads=8)}
    for k in got:
        assert np.array_equal(got[k][0], want[k][0]), k + " reason bytes"
        assert np.array_equal(got[k][1], want[k][1]), k + " bitmap"
    assert np.array_equal(bits(got["sts"][1], fleet.n), (got["sts"][0] & 0x0F) != 0)
    l1 = O.l1_diff_all(fleet)                    # and the object-level oracle on EVERY object
    for k in got:
        assert np.array_equal(got[k][0], l1[k]), k + " reason bytes against the object-level oracle"
    return got


def test_config2_10k_notebooks(ctx):
    """BASELINE config #2: 10k synthetic Notebooks, StatefulSet/Service diff on one GPU."""
    fleet = host.Fleet(host.FLEET_KUBEFLOW, host.SEED_BASE + 2, 10000, 200)
    got = _check_diff_fleet(ctx, fleet)
    frac = ((got["sts"][0] & 0x0F) != 0).mean()
    assert 0.3 < frac < 0.6          # 40% identical + the undetected classes


def test_config4_odh_notebooks(ctx):
    """BASELINE config #4 (ODH webhook-shaped, incl. VirtualService) at its full size: all 100 k notebooks against
    the packed oracle and against the object-level oracle."""
    fleet = host.Fleet(host.FLEET_ODH, host.SEED_BASE + 4, 100000, 2000)
    _check_diff_fleet(ctx, fleet)


def test_config3_1m_kernel_records(ctx):
    """BASELINE config #3: ~1M kernel last_activity records over 250k notebooks."""
    cf = host.CullFleet(host.SEED_BASE + 3, 250000)
    b = cf.pack()
    assert 900_000 < b.records()[0] < 1_100_000
    now = (cf.now_sec, cf.now_nsec)
    for mode in (N.CULL_PREDICATES, N.CULL_RECONCILE):
        got = ctx.cull(b, now[0], now[1], 1440, 1, mode)
        want = O.packed_cull(b, now, 1440, 1, mode, threads=8)
        for k in range(4):
            assert np.array_equal(got[k], want[k]), (mode, k)
        a1, am1 = O.l1_cull_all(cf, mode=mode)   # the object-level oracle on EVERY notebook
        assert np.array_equal(a1, got[0]) and np.array_equal(am1, got[1]), mode
        for i in range(0, cf.n, 503):            # the annotation string it would write: a sample
            a, am, la1 = O.cull(cf.json(i), now, 1440, 1, mode)
            assert la1 == b.new_last_activity(i, got[0][i], got[1][i], now[0])


# --------------------------------------------------------------- edge cases
@pytest.mark.parametrize("n", [1, 31, 32, 33, 63, 64, 65, 127, 129, 8191, 8193])
def test_ragged_sizes(ctx, n):
    fleet = host.Fleet(host.FLEET_KUBEFLOW, 7, n, 3)
    sts, svc, vs = fleet.pack()
    for got, want in ((ctx.diff_sts(sts), O.packed_diff_sts(sts)), (ctx.diff_svc(svc), O.packed_diff_svc(svc)),
                      (ctx.diff_vs(vs), O.packed_diff_vs(vs))):
        assert np.array_equal(got[0], want[0]) and np.array_equal(got[1], want[1])
    cf = host.CullFleet(11, n)
    b = cf.pack()
    got, want = ctx.cull(b, cf.now_sec, cf.now_nsec, 1440, 1, 1), O.packed_cull(b, (cf.now_sec, cf.now_nsec), 1440, 1, 1)
    for k in range(4):
        assert np.array_equal(got[k], want[k])


def test_empty_batches(ctx):
    assert ctx.diff_sts(host.StsBatch())[0].size == 0
    assert ctx.diff_svc(host.SvcBatch())[0].size == 0
    assert ctx.diff_vs(host.VsBatch())[0].size == 0
    assert ctx.cull(host.CullBatch(), 0, 0)[0].size == 0


def test_objects_with_no_maps_and_no_records(ctx):
    bare = {"metadata": {"name": "x"}, "spec": {"replicas": 1, "template": {"spec": {"containers": [{"name": "c"}]}}}}
    r, bm = ctx.diff_sts(host.StsBatch().add(host.StatefulSet(bare), host.StatefulSet(bare)))
    assert r[0] == 0 and bm[0] == 0
    b = host.CullBatch()
    for _ in range(70):
        b.add({"annotations": None, "kernels": [], "terminals": []})
    a, am, cb, pb = ctx.cull(b, 100, 5, 1, 1, 0)
    assert not a.any() and (am == N.NBG_ARGMAX_NONE).all() and not cb.any() and not pb.any()


def test_long_segments_and_big_maps(ctx):
    """max-size rows: a notebook with 5000 kernels, an object with 3000 labels."""
    ks = [{"last_activity": host.format_rfc3339(1767000000 + (i * 7919) % 100000), "execution_state": "idle"} for i in range(5000)]
    nb = {"annotations": {culler.LAST_ACTIVITY_ANNOTATION: "2025-01-01T00:00:00Z"}, "kernels": ks, "terminals": ks[:777]}
    b = host.CullBatch().add({"annotations": {}}).add(nb).add({"annotations": {}})
    got, want = ctx.cull(b, cases.NOW[0], cases.NOW[1], 1440, 1, 0), O.packed_cull(b, cases.NOW, 1440, 1, 0)
    for k in range(4):
        assert np.array_equal(got[k], want[k])
    assert got[0][1] & 3 and got[1][1] == O.cull(nb, cases.NOW, 1440, 1, 0)[1]
    labels = {"k%04d" % i: "v%d" % (i % 13) for i in range(3000)}
    frm = {"metadata": {"labels": labels}, "spec": {"replicas": 1}}
    live = json.loads(json.dumps(frm))
    live["metadata"]["labels"]["k2999"] = "changed"
    batch = host.StsBatch().add(host.StatefulSet(frm), host.StatefulSet(frm)).add(host.StatefulSet(frm), host.StatefulSet(live))
    r, _ = ctx.diff_sts(batch)
    assert r.tolist() == [0, 1]


def test_bad_arguments_are_rejected(ctx):
    lib = N.nbg()
    fleet = host.Fleet(host.FLEET_KUBEFLOW, 1, 100, 2)
    sts, _, _ = fleet.pack(svc=False, vs=False)
    view = sts.view
    reason = np.zeros(100, np.uint8)
    bm = np.zeros(2, np.uint64)
    rp, bp = reason.ctypes.data_as(C.c_void_p), bm.ctypes.data_as(C.c_void_p)
    assert lib.nbg_diff_sts(None, C.byref(view), 100, rp, bp) == N.NBG_E_INVAL
    assert lib.nbg_diff_sts(ctx.handle, None, 100, rp, bp) == N.NBG_E_INVAL
    assert lib.nbg_diff_sts(ctx.handle, C.byref(view), 100, None, bp) == N.NBG_E_INVAL
    off = sts.arrays()["labels_to.off"].copy()
    off[50] = off[51] + 5                               # not monotone
    view.labels_to.off = off.ctypes.data_as(N.u32p)
    assert lib.nbg_diff_sts(ctx.handle, C.byref(view), 100, rp, bp) == N.NBG_E_RANGE
    cb = host.CullFleet(3, 10).pack()
    p = N.nbg_cull_params(0, 1_000_000_000, 1, 1, 0)    # nsec out of range
    a, am = np.zeros(10, np.uint8), np.zeros(10, np.uint32)
    v = cb.view
    assert lib.nbg_cull(ctx.handle, C.byref(v), 10, C.byref(p), a.ctypes.data_as(C.c_void_p), am.ctypes.data_as(C.c_void_p), bp, bp) == N.NBG_E_INVAL
    p = N.nbg_cull_params(0, 0, 2_000_000_000, 1, 0)
    assert lib.nbg_cull(ctx.handle, C.byref(v), 10, C.byref(p), a.ctypes.data_as(C.c_void_p), am.ctypes.data_as(C.c_void_p), bp, bp) == N.NBG_E_INVAL
    with pytest.raises(pkg.NbgError):
        pkg.Context([99])


# ---------------------------------------- device-pointer API, big inputs, shards
def test_device_pointer_api_and_tiling_property(ctx):
    """Inputs resident in HBM (torch tensors), launched on torch's stream.  A fleet tiled K
    times must give the base result K times (size-independent property at ~0.6M objects)."""
    dev = importlib.import_module("odh-kubeflow_b200.device")
    import torch
    fleet = host.Fleet(host.FLEET_KUBEFLOW, host.SEED_BASE + 5, 9984, 300)   # multiple of 64
    sts, svc, vs = fleet.pack()
    base = ctx.diff_sts(sts)
    K = 64
    d = dev.DeviceBatch.from_host(sts, device=0, tile=K)
    reason, bitmap = d.run(ctx)
    torch.cuda.synchronize()
    assert d.n == 9984 * K
    r = reason.cpu().numpy().reshape(K, -1)
    assert (r == base[0][None, :]).all()
    assert (bitmap.cpu().numpy().view(np.uint64).reshape(K, -1) == base[1][None, :]).all()
    for batch, fn in ((svc, ctx.diff_svc), (vs, ctx.diff_vs)):
        d = dev.DeviceBatch.from_host(batch, device=0, tile=3)
        reason, bitmap = d.run(ctx)
        torch.cuda.synchronize()
        want = fn(batch)
        assert (reason.cpu().numpy().reshape(3, -1) == want[0][None, :]).all()
    cf = host.CullFleet(5, 64 * 500)
    cb = cf.pack()
    want = ctx.cull(cb, cf.now_sec, cf.now_nsec, 1440, 1, 1)
    d = dev.DeviceBatch.from_host(cb, device=0, tile=5)
    outs = d.run(ctx, now=(cf.now_sec, cf.now_nsec), cull_idle_min=1440, check_period_min=1, mode=1)
    torch.cuda.synchronize()
    assert (outs[0].cpu().numpy().reshape(5, -1) == want[0][None, :]).all()
    assert (outs[1].cpu().numpy().view(np.uint32).reshape(5, -1) == want[1][None, :]).all()
    assert (outs[2].cpu().numpy().view(np.uint64).reshape(5, -1) == want[2][None, :]).all()


def test_sharding_by_namespace_is_a_partition(ctx):
    """shard by nbg_shard_of(hash(namespace)); per-shard results gathered on the
    host equal the unsharded result."""
    fleet = host.Fleet(host.FLEET_KUBEFLOW, 21, 6000, 97)
    whole = ctx.diff_sts(fleet.pack(svc=False, vs=False)[0])[0]
    for g in (2, 4, 8):
        shards = fleet.shard_indices(g)
        assert sum(len(s) for s in shards) == fleet.n
        out = np.full(fleet.n, 255, np.uint8)
        for idx in shards:
            # one namespace never spans two shards
            r, _ = ctx.diff_sts(fleet.pack(idx=idx, svc=False, vs=False)[0])
            out[idx.astype(np.int64)] = r
        assert np.array_equal(out, whole)
    ns = np.array([fleet.ns_hash(i) for i in range(fleet.n)], dtype=np.uint64)
    for idx in fleet.shard_indices(4):
        others = np.setdiff1d(np.arange(fleet.n), idx.astype(np.int64))
        assert not np.intersect1d(ns[idx.astype(np.int64)], ns[others]).size


def test_concurrent_callers(ctx):
    """The reference has two concurrent callers (notebook + Culler controllers)."""
    import threading
    fleet = host.Fleet(host.FLEET_KUBEFLOW, 31, 20000, 50)
    sts = fleet.pack(svc=False, vs=False)[0]
    cb = host.CullFleet(32, 50000).pack()
    want_s, want_c = O.packed_diff_sts(sts)[0], O.packed_cull(cb, cases.NOW, 1440, 1, 1)[0]
    errs = []

    def a():
        for _ in range(5):
            if not np.array_equal(ctx.diff_sts(sts)[0], want_s):
                errs.append("sts")

    def b():
        for _ in range(5):
            if not np.array_equal(ctx.cull(cb, cases.NOW[0], cases.NOW[1], 1440, 1, 1)[0], want_c):
                errs.append("cull")

    ts = [threading.Thread(target=a), threading.Thread(target=b)]
    [t.start() for t in ts]
    [t.join() for t in ts]
    assert not errs


# ------------------------------------------------------------ sparse tiles
@pytest.mark.parametrize("kind,n", [(host.FLEET_KUBEFLOW, 20011), (host.FLEET_ODH, 9000), (host.FLEET_KUBEFLOW, 1),
                                    (host.FLEET_KUBEFLOW, 65), (host.FLEET_ODH, 32769)])
def test_sparse_tiles_give_the_dense_answer(ctx, kind, n):
    fleet = host.Fleet(kind, host.SEED_BASE + 9, n, max(n // 30, 1))
    sts = fleet.pack(svc=False, vs=False)[0]
    dense = ctx.diff_sts(sts)
    sparse = ctx.diff_sts_sparse(sts.sparse())
    want = O.packed_diff_sts_sparse(sts.sparse(), threads=4)
    assert np.array_equal(sparse[0], dense[0]) and np.array_equal(sparse[1], dense[1])
    assert np.array_equal(sparse[0], want[0]) and np.array_equal(sparse[1], want[1])
    assert sts.sparse().algorithmic_bytes() < 0.6 * sts.algorithmic_bytes()


@pytest.mark.parametrize("name,frm,to,want", STS_PAIRS, ids=[p[0][:60] for p in STS_PAIRS])
def test_derived_sts_sparse(ctx, name, frm, to, want):
    b = host.StsBatch().add(host.StatefulSet(frm), host.StatefulSet(to))
    reason, bm = ctx.diff_sts_sparse(b.sparse())
    assert int(reason[0]) == want and int(bm[0]) == int((want & 0x0F) != 0)


def test_sparse_device_api_tiled(ctx):
    dev = importlib.import_module("odh-kubeflow_b200.device")
    import torch
    fleet = host.Fleet(host.FLEET_KUBEFLOW, host.SEED_BASE + 5, 9984, 300)
    sts = fleet.pack(svc=False, vs=False)[0]
    base = ctx.diff_sts(sts)
    d = dev.DeviceBatch.from_host(sts.sparse(), device=0, tile=40)
    reason, bitmap = d.run(ctx)
    torch.cuda.synchronize()
    assert (reason.cpu().numpy().reshape(40, -1) == base[0][None, :]).all()
    assert (bitmap.cpu().numpy().view(np.uint64).reshape(40, -1) == base[1][None, :]).all()


def test_sparse_bad_runs_are_rejected(ctx):
    lib = N.nbg()
    sts = host.Fleet(host.FLEET_KUBEFLOW, 1, 100, 2).pack(svc=False, vs=False)[0]
    sp = sts.sparse()
    view = sp.view
    off = sp.arrays()["pod_to.voff"].copy()
    off[10:] += 500                                     # one run longer than a tile
    view.pod_to.off = off.ctypes.data_as(N.u32p)
    reason, bm = np.zeros(100, np.uint8), np.zeros(2, np.uint64)
    assert lib.nbg_diff_sts_sparse(ctx.handle, C.byref(view), 100, reason.ctypes.data_as(C.c_void_p),
                                   bm.ctypes.data_as(C.c_void_p)) == N.NBG_E_RANGE
    # a run two values shorter than its mask says (the next run moved up): monotone, even, and still refused —
    # the kernels size their reads of both sides by the masks
    off = sp.arrays()["pod_to.voff"].copy()
    off[11:] -= 2
    view.pod_to.off = off.ctypes.data_as(N.u32p)
    assert lib.nbg_diff_sts_sparse(ctx.handle, C.byref(view), 100, reason.ctypes.data_as(C.c_void_p),
                                   bm.ctypes.data_as(C.c_void_p)) == N.NBG_E_RANGE
    store = pkg.engine.Store(ctx, 100)
    try:
        idx = np.arange(100, dtype=np.uint64)
        assert lib.nbg_store_put(store._h, idx.ctypes.data_as(N.u64p), 100, C.byref(view), None, None, None) == N.NBG_E_RANGE
    finally:
        store.close()


# ------------------------------------ desired Service / VirtualService made on the device
@pytest.mark.parametrize("kind,n", [(host.FLEET_KUBEFLOW, 30001), (host.FLEET_ODH, 20000), (host.FLEET_ODH, 1), (host.FLEET_KUBEFLOW, 97)])
def test_device_generated_desired_side_gives_the_same_answer(ctx, kind, n):
    fleet = host.Fleet(kind, host.SEED_BASE + 15, n, max(n // 40, 1))
    _, svc, vs = fleet.pack(sts=False)
    want_svc, want_vs = ctx.diff_svc(svc), ctx.diff_vs(vs)
    got_svc = ctx.diff_svc_gen(svc.gen())
    got_vs = ctx.diff_vs_gen(vs.gen(), host.vs_gen_consts())
    assert np.array_equal(got_svc[0], want_svc[0]) and np.array_equal(got_svc[1], want_svc[1])
    assert np.array_equal(got_vs[0], want_vs[0]) and np.array_equal(got_vs[1], want_vs[1])
    assert np.array_equal(got_svc[0], O.packed_diff_svc(svc)[0]) and np.array_equal(got_vs[0], O.packed_diff_vs(vs)[0])
    if n > 1000:
        assert svc.gen().algorithmic_bytes() < 0.62 * svc.algorithmic_bytes()
        assert vs.gen().algorithmic_bytes() < 0.80 * vs.algorithmic_bytes()


@pytest.mark.parametrize("name,frm,to,want", SVC_PAIRS, ids=[p[0][:60] for p in SVC_PAIRS])
def test_derived_svc_device_generated(ctx, name, frm, to, want):
    b = host.SvcBatch().add(host.Service(frm), host.Service(to))
    reason, bm = ctx.diff_svc_gen(b.gen())
    assert int(reason[0]) == want


def test_vs_gen_respects_the_environment(ctx):
    env = {"ISTIO_HOST": "nb.example.com", "ISTIO_GATEWAY": "istio-system/gw", "CLUSTER_DOMAIN": "corp.local"}
    nb = host.Notebook(cases.BASE_NOTEBOOK)
    desired = host.generate_virtualservice(nb, env)
    other = host.generate_virtualservice(nb)                    # default env: hosts/gateways/host differ
    b = host.VsBatch().add(desired, desired).add(desired, other)
    r, _ = ctx.diff_vs_gen(b.gen(), host.vs_gen_consts(env))
    assert r.tolist() == [0, 1]
    r, _ = ctx.diff_vs_gen(b.gen(), host.vs_gen_consts())       # wrong constants for this desired side
    assert r.tolist() == [1, 1] or r.tolist() == [1, 0]


def test_one_context_over_two_devices():
    """nbg_init(devices, 2): one process, the batch split in contiguous 64-aligned ranges, one per
    device, results gathered by the call itself.  Skipped on a one-GPU box."""
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    c2 = pkg.Context([0, 1])
    try:
        fleet = host.Fleet(host.FLEET_KUBEFLOW, 41, 50_001, 300)
        sts, svc, vs = fleet.pack()
        for got, want in ((c2.diff_sts(sts), O.packed_diff_sts(sts)), (c2.diff_sts_sparse(sts.sparse()), O.packed_diff_sts(sts)),
                          (c2.diff_svc(svc), O.packed_diff_svc(svc)), (c2.diff_vs(vs), O.packed_diff_vs(vs)),
                          (c2.diff_svc_gen(svc.gen()), O.packed_diff_svc(svc))):
            assert np.array_equal(got[0], want[0]) and np.array_equal(got[1], want[1])
        cb = host.CullFleet(42, 300_001).pack()
        got, want = c2.cull(cb, cases.NOW[0], cases.NOW[1], 1440, 1, 1), O.packed_cull(cb, cases.NOW, 1440, 1, 1)
        for k in range(4):
            assert np.array_equal(got[k], want[k])
        assert c2.launch_count() >= 12
    finally:
        c2.close()


def test_resident_batches_on_the_second_device_of_a_context():
    """Context([0, 1]) with batches resident on CUDA device 1: the *_device calls must be given the context's index
    of that device (1), not 0 — the kernels run on device 1's stream over device 1's pointers.  Skipped on a
    one-GPU box."""
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    devmod = importlib.import_module("odh-kubeflow_b200.device")
    c2 = pkg.Context([0, 1])
    try:
        assert c2.index_of(1) == 1 and c2.index_of(0) == 0
        with pytest.raises(ValueError):
            c2.index_of(5)
        n = 40_003
        fleet = host.Fleet(host.FLEET_KUBEFLOW, 43, n, 300)
        sts, svc, vs = fleet.pack()
        cf = host.CullFleet(44, n)
        cb = cf.pack()
        now = (cf.now_sec, cf.now_nsec)
        want = (O.packed_diff_sts(sts)[0], O.packed_diff_svc(svc)[0], O.packed_diff_vs(vs)[0], O.packed_cull(cb, now, 1440, 1, 1)[0])
        d = [devmod.DeviceBatch.from_host(b, device=1) for b in (sts.sparse(), svc.gen(), vs.gen(), cb)]
        for b, w in zip(d, want):                       # each kernel alone ...
            outs = b.run(c2, now=now) if b.kind == "CullBatch" else b.run(c2)
            torch.cuda.synchronize(1)
            assert np.array_equal(outs[0].cpu().numpy(), w)
            outs[0].zero_()
        devmod.ResidentStep(c2, *d, now=now).run()      # ... and the four as one step
        torch.cuda.synchronize(1)
        for b, w in zip(d, want):
            assert np.array_equal(b.outs[0].cpu().numpy(), w)
        one = pkg.Context([1])                          # a context on device 1 alone calls it index 0
        try:
            assert one.index_of(1) == 0
            d[0].outs[0].zero_()
            d[0].run(one)
            torch.cuda.synchronize(1)
            assert np.array_equal(d[0].outs[0].cpu().numpy(), want[0])
        finally:
            one.close()
    finally:
        c2.close()


@pytest.mark.parametrize("n", [1, 5000, 140_001])
def test_step_in_one_call_equals_the_four_calls(ctx, n):
    fleet = host.Fleet(host.FLEET_KUBEFLOW, host.SEED_BASE + 16, n, max(n // 40, 1))
    sts, svc, vs = fleet.pack()
    cf = host.CullFleet(host.SEED_BASE + 17, n)
    cb = cf.pack()
    r = ctx.step(sts.sparse(), svc.gen(), vs.gen(), host.vs_gen_consts(), cb, cf.now_sec, cf.now_nsec)
    for k, want in (("sts", O.packed_diff_sts(sts)), ("svc", O.packed_diff_svc(svc)), ("vs", O.packed_diff_vs(vs))):
        assert np.array_equal(r[k][0], want[0]) and np.array_equal(r[k][1], want[1]), k
    want = O.packed_cull(cb, (cf.now_sec, cf.now_nsec), 1440, 1, 1)
    for j in range(4):
        assert np.array_equal(r["cull"][j], want[j])


def _masked(t, n):
    """bits of a device bitmap (bits past n belong to nobody)"""
    words = (n + 63) // 64
    return pkg.engine.bitmap_to_bools(t.view(np.uint64)[:words], n)


@pytest.mark.parametrize("n", [1, 33, 4097, 150_001])
def test_resident_step_equals_the_stand_alone_kernels(ctx, n):
    """nbg_step_device: StatefulSet kernel on the caller's stream, the three small predicates on the
    context's side streams, joined back — same outputs as the four device calls, for every subset."""
    import torch
    devmod = importlib.import_module("odh-kubeflow_b200.device")
    fleet = host.Fleet(host.FLEET_ODH if n % 2 else host.FLEET_KUBEFLOW, host.SEED_BASE + 18, n, max(n // 40, 1))
    sts, svc, vs = fleet.pack()
    cf = host.CullFleet(host.SEED_BASE + 19, n)
    cb = cf.pack()
    now = (cf.now_sec, cf.now_nsec)
    d = dict(zip(("sts", "svc", "vs", "cull"), (devmod.DeviceBatch.from_host(b) for b in (sts.sparse(), svc.gen(), vs.gen(), cb))))
    side = torch.cuda.Stream()
    for mode, subset, stream in ((1, ("sts", "svc", "vs", "cull"), None), (0, ("sts", "svc", "vs", "cull"), side), (1, ("sts",), None),
                                 (1, ("svc", "cull"), side), (1, ("vs",), None)):
        want = {"sts": O.packed_diff_sts(sts), "svc": O.packed_diff_svc(svc), "vs": O.packed_diff_vs(vs),
                "cull": O.packed_cull(cb, now, 1440, 1, mode)}
        for b in d.values():
            for o in b.outs:
                o.fill_(0xA5)
        torch.cuda.synchronize()
        step = devmod.ResidentStep(ctx, now=now, mode=mode, time_sts=(stream is None), **{k: d[k] for k in subset})
        before = ctx.launch_count()
        for _ in range(3):                                   # back to back: the fork/join events are reused
            if stream is None:
                step.run()
            else:
                step.run(stream=stream.cuda_stream)
        # ordered on the caller's stream like one launch: a copy queued behind it sees the results
        if stream is None:
            got_dev = {k: [o.clone() for o in d[k].outs] for k in d}
        else:
            with torch.cuda.stream(stream):
                got_dev = {k: [o.clone() for o in d[k].outs] for k in d}
        torch.cuda.synchronize()
        assert ctx.launch_count() == before + 3 * len(subset)
        if stream is None and "sts" in subset:
            ms = step.sts_ms()
            assert len(ms) == 3 and all(t > 0 for t in ms)
        for k in d:
            got = [o.cpu().numpy() for o in got_dev[k]]
            if k not in subset:
                assert all((g == 0xA5).all() for g in got), k          # untouched
                continue
            assert np.array_equal(got[0], want[k][0]), k
            if k != "cull":
                assert np.array_equal(_masked(got[1], n), _masked(want[k][1], n)), k
            else:
                assert np.array_equal(got[1].view(np.uint32), want[k][1])
                assert np.array_equal(_masked(got[2], n), _masked(want[k][2], n))
                assert np.array_equal(_masked(got[3], n), _masked(want[k][3], n))


def test_step_device_rejects_bad_arguments(ctx):
    lib = N.nbg()
    out = N.nbg_step_outputs()
    cb = N.nbg_cull_batch()
    p = N.nbg_cull_params(0, 0, 1440, 1, 1)
    none = [None] * 5
    assert lib.nbg_step_device(ctx.handle, 0, *none, None, 10, None, None, None, None) == N.NBG_E_INVAL
    assert lib.nbg_step_device(ctx.handle, 0, *none, None, 10, C.byref(out), None, None, None) == N.NBG_OK   # nothing to do
    assert lib.nbg_step_device(ctx.handle, 0, None, None, None, None, C.byref(cb), None, 10, C.byref(out), None, None, None) == N.NBG_E_INVAL
    assert lib.nbg_step_device(ctx.handle, 0, None, None, None, None, C.byref(cb), C.byref(p), 10, C.byref(out), None, None, None) == N.NBG_E_INVAL
    assert lib.nbg_step_device(ctx.handle, 7, *none, None, 10, C.byref(out), None, None, None) != N.NBG_OK


def test_resident_step_replays_from_a_cuda_graph(ctx):
    """nbg_step_device allocates nothing and synchronises nothing, so a caller can capture it (side
    streams included) once and replay the step: same outputs as the direct call."""
    import torch
    devmod = importlib.import_module("odh-kubeflow_b200.device")
    n = 10_000
    fleet = host.Fleet(host.FLEET_KUBEFLOW, host.SEED_BASE + 2, n, 200)
    sts, svc, vs = fleet.pack()
    cf = host.CullFleet(host.SEED_BASE + 3, n)
    cb = cf.pack()
    d = [devmod.DeviceBatch.from_host(b) for b in (sts.sparse(), svc.gen(), vs.gen(), cb)]
    step = devmod.ResidentStep(ctx, *d, now=(cf.now_sec, cf.now_nsec))
    s = torch.cuda.Stream()
    with torch.cuda.stream(s):
        step.run()
    torch.cuda.synchronize()
    graph = torch.cuda.CUDAGraph()
    with torch.cuda.graph(graph, stream=s):
        step.run()
    for b in d:
        for o in b.outs:
            o.fill_(0xA5)
    torch.cuda.synchronize()
    graph.replay()
    graph.replay()
    torch.cuda.synchronize()
    want = (O.packed_diff_sts(sts), O.packed_diff_svc(svc), O.packed_diff_vs(vs), O.packed_cull(cb, (cf.now_sec, cf.now_nsec), 1440, 1, 1))
    for b, w in zip(d, want):
        assert np.array_equal(b.outs[0].cpu().numpy(), w[0])
        assert np.array_equal(_masked(b.outs[-1].cpu().numpy(), n), _masked(w[len(b.outs) - 1], n))


def test_ten_million_notebooks_repeat_the_checked_tile(ctx):
    """BASELINE config #5 at its full size (10 M notebooks on one GPU, ~11 GB resident) through the size-independent
    property the layout offers: a fleet of 156 250 notebooks checked against the oracle, tiled 64x on the device,
    must give the same 156 250 answers 64 times — for all four predicates of the step."""
    import torch
    devmod = importlib.import_module("odh-kubeflow_b200.device")
    base, tiles = 156_250, 64
    fleet = host.Fleet(host.FLEET_KUBEFLOW, host.SEED_BASE + 5, base, 3000)
    sts, svc, vs = fleet.pack()
    cf = host.CullFleet(host.SEED_BASE + 6, base)
    cb = cf.pack()
    now = (cf.now_sec, cf.now_nsec)
    want = (O.packed_diff_sts(sts, threads=4)[0], O.packed_diff_svc(svc, threads=4)[0], O.packed_diff_vs(vs, threads=4)[0],
            O.packed_cull(cb, now, 1440, 1, 1, threads=4))
    d = [devmod.DeviceBatch.from_host(b, tile=tiles) for b in (sts.sparse(), svc.gen(), vs.gen(), cb)]
    assert d[0].n == 10_000_000
    devmod.ResidentStep(ctx, *d, now=now).run()
    torch.cuda.synchronize()
    for k in range(3):
        assert (d[k].outs[0].cpu().numpy().reshape(tiles, base) == want[k][None, :]).all(), k
    assert (d[3].outs[0].cpu().numpy().reshape(tiles, base) == want[3][0][None, :]).all()
    assert (d[3].outs[1].cpu().numpy().view(np.uint32).reshape(tiles, base) == want[3][1][None, :]).all()
    # bitmaps: 156 250 is not a multiple of 64, so tiles are not word-aligned: compare bit by bit
    bits = pkg.engine.bitmap_to_bools(d[0].outs[1].cpu().numpy().view(np.uint64), 10_000_000).reshape(tiles, base)
    assert (bits == ((want[0] & 0x0F) != 0)[None, :]).all()
    del d
    torch.cuda.empty_cache()


def test_config5_distinct_notebooks_whole_and_in_balanced_shards(ctx):
    """BASELINE config #5 with DISTINCT notebooks (1.2 M here; bench.py runs and checks all 10 M): lean-packed, one
    nbg_step_device over the whole fleet against the packed oracle on every object and against the object-level
    oracle on a 20 k subset; then the same fleet cut into 8 shards by the size-aware namespace table
    (nbg_balance_shards) — every shard's results, scattered back through its index list, must reproduce the
    unsharded results, and no namespace may span two shards."""
    import torch
    devmod = importlib.import_module("odh-kubeflow_b200.device")
    n = 1_200_000
    fleet = host.Fleet(host.FLEET_KUBEFLOW, host.SEED_BASE + 5, n, n // 50)
    cf = host.CullFleet(host.SEED_BASE + 5, n)
    now = (cf.now_sec, cf.now_nsec)
    sts, svc, vs = fleet.pack(lean=True)
    cb = cf.pack()
    d = [devmod.DeviceBatch.from_host(b) for b in (sts.sparse(), svc.gen(), vs.gen(), cb)]
    devmod.ResidentStep(ctx, *d, now=now).run()
    torch.cuda.synchronize()
    whole = [b.outs[0].cpu().numpy() for b in d] + [d[3].outs[1].cpu().numpy().view(np.uint32)]
    want = (O.packed_diff_sts_sparse(sts.sparse(), threads=8)[0], O.packed_diff_svc(svc, threads=8)[0],
            O.packed_diff_vs(vs, threads=8)[0]) + O.packed_cull(cb, now, 1440, 1, 1, threads=8)[:2]
    for k, (g, w) in enumerate(zip((whole[0], whole[1], whole[2], whole[3], whole[4]), want)):
        assert np.array_equal(g, w), k
    sub = (np.arange(20_000, dtype=np.uint64) * np.uint64(n // 20_000))
    l1 = O.l1_diff_all(fleet, idx=sub)
    a1, am1 = O.l1_cull_all(cf, idx=sub)
    s = sub.astype(np.int64)
    assert np.array_equal(whole[0][s], l1["sts"]) and np.array_equal(whole[1][s], l1["svc"]) and np.array_equal(whole[2][s], l1["vs"])
    assert np.array_equal(whole[3][s], a1) and np.array_equal(whole[4][s], am1)
    del d
    torch.cuda.empty_cache()
    # the same fleet in 8 balanced shards, each through the host-buffer step
    table, load = fleet.shard_table(8, balanced=True)
    t_hash, l_hash = fleet.shard_table(8, balanced=False)
    assert load.max() / load.mean() < l_hash.max() / l_hash.mean()
    out = [np.full(n, 255, np.uint8) for _ in range(4)]
    consts = host.vs_gen_consts()
    for shard in range(8):
        idx = fleet.shard_indices_by_table(table, shard)
        assert len(idx) == int(load[shard])
        a, b, c = fleet.pack(idx=idx, lean=True)
        r = ctx.step(a.sparse(), b.gen(), c.gen(), consts, cf.pack(idx=idx), now[0], now[1])
        for k, key in enumerate(("sts", "svc", "vs", "cull")):
            out[k][idx.astype(np.int64)] = r[key][0]
    for k in range(4):
        assert np.array_equal(out[k], whole[k]), k
