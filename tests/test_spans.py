"""Spans instead of CSR (ABI v2, include/nbgraft.h `end` pointers): every variable-length part of a batch
scattered through a heap in random order, with gaps full of garbage.  Same objects => same answers, from the
packed oracle (CPU) and from every kernel (GPU, device-pointer entry points)."""
import importlib

import numpy as np
import pytest

import oracle as O
import spans

pkg = O.pkg
host, N = pkg.host, pkg._native


def _batches(n=3001):
    fleet = host.Fleet(host.FLEET_ODH, host.SEED_BASE + 31, n, max(n // 40, 1))
    sts, svc, vs = fleet.pack()
    cf = host.CullFleet(host.SEED_BASE + 32, n)
    gb, rb = host.GenBatch(N.GEN_DEEPEQUAL, "metadata.labels", "spec.template.spec"), host.RestartBatch()
    for k, o in enumerate(fleet.objects(0, 257)):
        gb.add(o["sts_from"], o["sts_to"])
        m = o["notebook"].clone()
        if k % 3:
            host.odh_mutate(m, 2 | 4 | 8 | 16)
        rb.add(o["notebook"], o["notebook"] if k % 5 else m, m)
    return {"sts": sts, "sts_sparse": sts.sparse(), "svc": svc, "vs": vs, "svc_gen": svc.gen(), "vs_gen": vs.gen(),
            "cull": cf.pack(), "generic": gb, "restart": rb}, (cf.now_sec, cf.now_nsec)


ORACLE = {"sts": O.packed_diff_sts, "sts_sparse": O.packed_diff_sts_sparse, "svc": O.packed_diff_svc, "vs": O.packed_diff_vs,
          "generic": O.packed_diff_generic, "restart": O.packed_restart_block}


def test_packed_oracle_reads_spans_like_csr():
    b, now = _batches()
    for name, fn in ORACLE.items():
        want = fn(b[name])
        got = fn(spans.for_oracle(spans.SpanBatch(b[name], seed=7)))
        assert np.array_equal(got[0], want[0]) and np.array_equal(got[1], want[1]), name
    for mode in (0, 1):
        want = O.packed_cull(b["cull"], now, 1440, 1, mode)
        got = O.packed_cull(spans.for_oracle(spans.SpanBatch(b["cull"], seed=8)), now, 1440, 1, mode)
        assert all(np.array_equal(got[k], want[k]) for k in range(4))


@pytest.fixture(scope="module")
def ctx():
    c = pkg.Context()
    yield c
    c.close()


@pytest.mark.gpu
@pytest.mark.parametrize("n", [1, 33, 3001])
def test_kernels_read_spans_like_csr(ctx, n):
    import torch
    devmod = importlib.import_module("odh-kubeflow_b200.device")
    b, now = _batches(n)
    want = {name: fn(b[name]) for name, fn in ORACLE.items()}
    want["svc_gen"], want["vs_gen"] = want["svc"], want["vs"]
    for name in ("sts", "sts_sparse", "svc", "vs", "svc_gen", "vs_gen", "generic", "restart"):
        d = devmod.DeviceBatch.from_host(spans.SpanBatch(b[name], seed=11 + n))
        outs = d.run(ctx)
        torch.cuda.synchronize()
        assert np.array_equal(outs[0].cpu().numpy(), want[name][0]), name
    for mode in (0, 1):
        d = devmod.DeviceBatch.from_host(spans.SpanBatch(b["cull"], seed=5 + n))
        outs = d.run(ctx, now=now, mode=mode)
        torch.cuda.synchronize()
        w = O.packed_cull(b["cull"], now, 1440, 1, mode)
        assert np.array_equal(outs[0].cpu().numpy(), w[0])
        assert np.array_equal(outs[1].cpu().numpy().view(np.uint32), w[1])


@pytest.mark.gpu
def test_host_entry_points_reject_spans(ctx):
    import ctypes as C
    b, now = _batches(65)
    sb = spans.SpanBatch(b["svc"], seed=3)
    view = spans.struct_over(sb)
    reason, bm = np.zeros(65, np.uint8), np.zeros(2, np.uint64)
    assert N.nbg().nbg_diff_svc(ctx.handle, C.byref(view), 65, reason.ctypes.data_as(C.c_void_p),
                                bm.ctypes.data_as(C.c_void_p)) == N.NBG_E_INVAL
    view = spans.struct_over(spans.SpanBatch(b["cull"], seed=3))
    p = N.nbg_cull_params(now[0], now[1], 1440, 1, 1)
    a, am = np.zeros(65, np.uint8), np.zeros(65, np.uint32)
    assert N.nbg().nbg_cull(ctx.handle, C.byref(view), 65, C.byref(p), a.ctypes.data_as(C.c_void_p), am.ctypes.data_as(C.c_void_p),
                            bm.ctypes.data_as(C.c_void_p), bm.ctypes.data_as(C.c_void_p)) == N.NBG_E_INVAL
