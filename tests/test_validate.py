"""nbg_check_step_device: the host-buffer calls' argument checks, run on the device for resident batches (the
*_device entry points only enqueue, so they cannot look).  Good batches pass; a batch with one bad word anywhere is
refused with the kind of fault named."""
import ctypes as C
import importlib

import numpy as np
import pytest

import oracle as O

pkg = O.pkg
host, N = pkg.host, pkg._native
pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ctx():
    c = pkg.Context()
    yield c
    c.close()


@pytest.fixture(scope="module")
def batches():
    devmod = importlib.import_module("odh-kubeflow_b200.device")
    n = 5_003
    fleet = host.Fleet(host.FLEET_ODH, host.SEED_BASE + 51, n, 40)
    sts, svc, vs = fleet.pack(lean=True)
    cb = host.CullFleet(host.SEED_BASE + 52, n).pack()
    keep = (sts, svc, vs, cb)
    return devmod, [devmod.DeviceBatch.from_host(b) for b in (sts.sparse(), svc.gen(), vs.gen(), cb)], keep


def _poke(t, word_index, value, dtype=np.uint32):
    """Overwrite one word of a device tensor (viewed as `dtype`); returns the old value."""
    import torch
    v = t.view(torch.int32 if dtype == np.uint32 else torch.int64)
    old = int(v[word_index].item())
    v[word_index] = int(np.array(value, dtype=dtype).view(np.int32 if dtype == np.uint32 else np.int64))
    return old


def test_good_batches_pass_whole_and_one_at_a_time(ctx, batches):
    devmod, d, _ = batches
    assert devmod.check_step(ctx, *d) == 0
    for k in range(4):
        only = [d[j] if j == k else None for j in range(4)]
        assert devmod.check_step(ctx, *only) == 0


@pytest.mark.parametrize("which,tensor,index,value,bit", [
    (0, "pod_from.voff", 700, 0xFFFFFFF0, 2),       # a run far past the value array (also start > next start)
    (0, "pod_to.voff", 1, None, 8),                 # run shorter than its mask says (set below: off[1] += 2)
    (0, "labels_to.off", 4000, 0x7FFFFFFF, 2),      # a label run past its entries
    (1, "sel_to.off", 3, 0x0FFFFFFF, 2),
    (2, "hdr_from.off", 2500, 0x0FFFFFFF, 2),
    (3, "k_off", 100, 0x0FFFFFFF, 2),               # kernel records past the record arrays
    (3, "t_off", 5003, 0x0FFFFFFF, 2),              # the LAST offset (end of the last run)
])
def test_one_bad_word_is_found(ctx, batches, which, tensor, index, value, bit):
    devmod, d, _ = batches
    t = d[which].tensors[tensor]
    if value is None:
        value = (int(t.view(__import__("torch").int32)[index].item()) + 2) & 0xFFFFFFFF
    old = _poke(t, index, value)
    try:
        what = devmod.check_step(ctx, *d)
        assert what & bit, (tensor, what)
        only = [d[j] if j == which else None for j in range(4)]
        assert devmod.check_step(ctx, *only) & bit
    finally:
        _poke(t, index, np.uint32(old & 0xFFFFFFFF))
    assert devmod.check_step(ctx, *d) == 0


def test_misaligned_sparse_run_and_mask_mismatch(ctx, batches):
    devmod, d, _ = batches
    t = d[0].tensors["pod_from.voff"]
    import torch
    cur = int(t.view(torch.int32)[10].item())
    old = _poke(t, 10, cur + 1)                      # odd start: run 9 one longer, run 10 one shorter and misaligned
    try:
        what = devmod.check_step(ctx, d[0])
        assert what & 4 and what & 8
    finally:
        _poke(t, 10, np.uint32(old))
    m = d[0].tensors["pod_to.mask"]
    oldm = _poke(m, 3 * 77, 0, dtype=np.uint64)      # object 77: mask says fewer slots than its run holds
    try:
        assert devmod.check_step(ctx, d[0]) & 8
    finally:
        _poke(m, 3 * 77, np.uint64(oldm & 0xFFFFFFFFFFFFFFFF), dtype=np.uint64)
    assert devmod.check_step(ctx, d[0]) == 0


def test_extents_shorter_than_the_offsets_claim(ctx, batches):
    """The same good batch checked against payload lengths one element too short."""
    devmod, d, _ = batches
    lib = N.nbg()
    ext = N.nbg_step_extents()
    ext.cull_kernels = d[3].tensors["k_sec"].numel() // 8 - 1
    ext.cull_terminals = d[3].tensors["t_sec"].numel() // 8
    what = C.c_uint(0)
    rc = lib.nbg_check_step_device(ctx.handle, 0, None, None, None, C.byref(d[3]._struct), d[3].n, C.byref(ext), None, C.byref(what))
    assert rc == N.NBG_E_RANGE and what.value == 2
    ext.cull_kernels += 1
    assert lib.nbg_check_step_device(ctx.handle, 0, None, None, None, C.byref(d[3]._struct), d[3].n, C.byref(ext), None, None) == N.NBG_OK
    assert lib.nbg_check_step_device(ctx.handle, 0, None, None, None, C.byref(d[3]._struct), d[3].n, None, None, None) == N.NBG_E_INVAL
    assert lib.nbg_check_step_device(ctx.handle, 3, None, None, None, C.byref(d[3]._struct), d[3].n, C.byref(ext), None, None) == N.NBG_E_INVAL


def test_the_store_hands_out_batches_that_pass(ctx):
    """Spans form: the resident store's own view of itself after puts in scrambled order."""
    n = 3000
    fleet = host.Fleet(host.FLEET_KUBEFLOW, host.SEED_BASE + 53, n, 30)
    cf = host.CullFleet(host.SEED_BASE + 54, n)
    store = pkg.engine.Store(ctx, n)
    try:
        perm = np.random.default_rng(1).permutation(n).astype(np.uint64)
        for part in (perm[: n // 2], perm[n // 2:], perm[:500]):
            a, b, c = fleet.pack(idx=part, lean=True)
            store.put(part, a.sparse(), b.gen(), c.gen(), cf.pack(idx=part))
        assert store.check(n) == 0
    finally:
        store.close()
