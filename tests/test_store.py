"""The resident fleet (nbg_store_*): rows put into slots in any order, replaced in place many times, heaps
compacted along the way — a step over the store must always equal the packed oracle applied to the rows
the slots currently hold (every predicate is per object, so the expectation is assembled row by row)."""
import ctypes as C

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


class Source:
    """A fleet + culler fleet and the oracle's verdict on each of their rows."""

    def __init__(self, kind, seed, n):
        self.fleet = host.Fleet(kind, seed, n, max(n // 40, 1))
        self.cf = host.CullFleet(seed + 1, n)
        sts, svc, vs = self.fleet.pack()
        self.now = (self.cf.now_sec, self.cf.now_nsec)
        cull = O.packed_cull(self.cf.pack(), self.now, 1440, 1, 1)
        self.want = {"sts": O.packed_diff_sts(sts)[0], "svc": O.packed_diff_svc(svc)[0], "vs": O.packed_diff_vs(vs)[0],
                     "action": cull[0], "argmax": cull[1]}

    def rows(self, idx):
        sts, svc, vs = self.fleet.pack(idx=idx)
        return sts.sparse(), svc.gen(), vs.gen(), self.cf.pack(idx=idx), (sts, svc, vs)


def check(store, n, want, now):
    r = store.step(n, host.vs_gen_consts(), now[0], now[1])
    assert np.array_equal(r["sts"][0], want["sts"][:n])
    assert np.array_equal(r["svc"][0], want["svc"][:n])
    assert np.array_equal(r["vs"][0], want["vs"][:n])
    assert np.array_equal(r["cull"][0], want["action"][:n])
    assert np.array_equal(r["cull"][1], want["argmax"][:n])
    bits = pkg.engine.bitmap_to_bools
    assert np.array_equal(bits(r["sts"][1], n), (want["sts"][:n] & 0x0F) != 0)     # NBG_STS_UPDATE_MASK
    assert np.array_equal(bits(r["cull"][2], n), (want["action"][:n] & N.ACT_CULL) != 0)
    assert np.array_equal(bits(r["cull"][3], n), (want["action"][:n] & N.ACT_PERIOD) != 0)


def test_store_follows_replacements(ctx):
    n = 20_011
    a, b = Source(host.FLEET_KUBEFLOW, host.SEED_BASE + 41, n), Source(host.FLEET_ODH, host.SEED_BASE + 43, n)
    assert a.now == b.now
    rng = np.random.default_rng(5)
    store = pkg.engine.Store(ctx, n)
    try:
        want = {k: np.zeros(n, v.dtype) for k, v in a.want.items()}
        order = rng.permutation(n).astype(np.uint64)
        for chunk in np.array_split(order, 3):                    # initial load, slots in random order
            s1, s2, s3, s4, keep = a.rows(chunk)
            store.put(chunk, s1, s2, s3, s4)
            for k in want:
                want[k][chunk] = a.want[k][chunk]
        check(store, n, want, a.now)
        check(store, 4097, want, a.now)                           # a prefix of the slots
        for rnd in range(8):                                      # replacements: heaps fill with garbage and get compacted
            src = b if rnd % 2 == 0 else a
            sub = rng.choice(n, size=n // 3, replace=False).astype(np.uint64)
            s1, s2, s3, s4, keep = src.rows(sub)
            if rnd == 3:                                          # only the culler rows change (a poll, no object event)
                store.put(sub, cull=s4)
                for k in ("action", "argmax"):
                    want[k][sub] = src.want[k][sub]
            else:
                store.put(sub, s1, s2, s3, s4)
                for k in want:
                    want[k][sub] = src.want[k][sub]
            check(store, n, want, a.now)
        st = store.stats()
        assert st["compactions"] > 0 and st["rows_put"] == n + 8 * (n // 3) and st["puts"] == 11
        assert st["live_bytes"] + st["garbage_bytes"] <= st["heap_bytes"]
    finally:
        store.close()


def test_store_rejects_bad_arguments(ctx):
    lib = N.nbg()
    h = C.c_void_p()
    assert lib.nbg_store_create(ctx.handle, 0, 0, C.byref(h)) == N.NBG_E_INVAL
    assert lib.nbg_store_create(ctx.handle, 9, 16, C.byref(h)) == N.NBG_E_INVAL
    store = pkg.engine.Store(ctx, 64)
    try:
        src = Source(host.FLEET_KUBEFLOW, 3, 64)
        s1, s2, s3, s4, keep = src.rows(np.arange(8, dtype=np.uint64))
        with pytest.raises(pkg.engine.NbgError):
            store.put(np.arange(60, 68), s1, s2, s3, s4)          # slot 64.. out of range
        with pytest.raises(pkg.engine.NbgError):
            store.step(65, host.vs_gen_consts(), *src.now)
        store.put(np.arange(8), s1, s2, s3, s4)
        r = store.step(8, host.vs_gen_consts(), *src.now)
        assert np.array_equal(r["sts"][0], src.want["sts"][:8])
        # a slot named twice in one put is refused (two rows would race for it), and nothing is changed by the call
        before = store.stats()
        with pytest.raises(pkg.engine.NbgError) as e:
            store.put(np.array([1, 2, 3, 4, 5, 6, 7, 1]), s1, s2, s3, s4)
        assert e.value.status == N.NBG_E_INVAL
        assert store.stats() == before
        r = store.step(8, host.vs_gen_consts(), *src.now)
        assert np.array_equal(r["sts"][0], src.want["sts"][:8])
        store.put(np.array([7, 6, 5, 4, 3, 2, 1, 0]), s1, s2, s3, s4)         # the same slots again in another order: fine
        r = store.step(8, host.vs_gen_consts(), *src.now)
        assert np.array_equal(r["sts"][0], src.want["sts"][:8][::-1])
    finally:
        store.close()


def test_a_large_put_runs_as_several_chunks(ctx):
    """A put of more rows than one staging slot takes (> 65 536 rows, or > 48 MB) is cut into chunks that
    alternate between the two slots; slots are written in a scrambled order so that every chunk scatters all
    over the store.  The step over the store must equal the oracle on the rows the slots hold."""
    n = 150_001
    src = Source(host.FLEET_KUBEFLOW, host.SEED_BASE + 47, n)
    perm = np.random.default_rng(11).permutation(n).astype(np.uint64)
    s1, s2, s3, s4, keep = src.rows(np.arange(n, dtype=np.uint64))
    store = pkg.engine.Store(ctx, n)
    try:
        store.put(perm, s1, s2, s3, s4)                                       # row k -> slot perm[k]
        want = {k: np.zeros(n, v.dtype) for k, v in src.want.items()}
        for k, v in src.want.items():
            want[k][perm.astype(np.int64)] = v
        check(store, n, want, src.now)
        st = store.stats()
        assert st["puts"] == 1 and st["rows_put"] == n and st["garbage_bytes"] == 0
    finally:
        store.close()


def test_put_right_behind_an_asynchronous_step(ctx):
    """nbg_store_step_device only enqueues; a put issued straight after it (heap growth and compaction included)
    must neither disturb that step's results nor free what it is reading."""
    import torch
    n = 30_000
    a, b = Source(host.FLEET_KUBEFLOW, host.SEED_BASE + 51, n), Source(host.FLEET_ODH, host.SEED_BASE + 53, n)
    store = pkg.engine.Store(ctx, n)
    try:
        all_slots = np.arange(n, dtype=np.uint64)
        s1, s2, s3, s4, keep = a.rows(all_slots)
        store.put(all_slots, s1, s2, s3, s4)
        words = (n + 63) // 64
        outs = [torch.full((sz,), 0xA5, dtype=torch.uint8, device="cuda") for sz in
                (n, words * 8, n, words * 8, n, words * 8, n, n * 4, words * 8, words * 8)]
        o = N.nbg_step_outputs(*[t.data_ptr() for t in outs])
        p = N.nbg_cull_params(a.now[0], a.now[1], 1440, 1, 1)
        big = b.rows(all_slots)                                   # ODH rows are larger: every heap has to grow
        stream = torch.cuda.Stream()
        for _ in range(3):
            store.step_device(n, host.vs_gen_consts(), p, o, stream.cuda_stream)
        store.put(all_slots, *big[:4])                            # returns after ITS copies; the steps were queued first
        stream.synchronize()
        assert np.array_equal(outs[0].cpu().numpy(), a.want["sts"])
        assert np.array_equal(outs[6].cpu().numpy(), a.want["action"])
        r = store.step(n, host.vs_gen_consts(), *a.now)
        assert np.array_equal(r["sts"][0], b.want["sts"]) and np.array_equal(r["cull"][1], b.want["argmax"])
    finally:
        store.close()


def test_put_and_step_from_two_threads(ctx):
    """A controller's informer thread puts rows while its tick thread steps: whatever the interleaving, every slot's
    answer must be the oracle's answer for one of the two rows that slot can hold — never a torn mixture."""
    import threading
    n = 8192
    a, b = Source(host.FLEET_KUBEFLOW, host.SEED_BASE + 71, n), Source(host.FLEET_ODH, host.SEED_BASE + 73, n)
    store = pkg.engine.Store(ctx, n)
    errors, stop = [], threading.Event()
    try:
        all_slots = np.arange(n, dtype=np.uint64)
        rows = {"a": a.rows(all_slots), "b": b.rows(all_slots)}
        store.put(all_slots, *rows["a"][:4])

        def writer():
            rng = np.random.default_rng(1)
            try:
                for k in range(40):
                    src = "b" if k % 2 == 0 else "a"
                    lo = int(rng.integers(0, n - 512))
                    sub = np.arange(lo, lo + 512, dtype=np.uint64)
                    s = (a if src == "a" else b).rows(sub)
                    store.put(sub, *s[:4])
            except Exception as e:                                   # noqa: BLE001
                errors.append(e)
            finally:
                stop.set()

        t = threading.Thread(target=writer)
        t.start()
        steps = 0
        while not stop.is_set() or steps < 3:
            r = store.step(n, host.vs_gen_consts(), *a.now)
            for key, got in (("sts", r["sts"][0]), ("svc", r["svc"][0]), ("vs", r["vs"][0]), ("action", r["cull"][0])):
                ok = (got == a.want[key]) | (got == b.want[key])
                assert ok.all(), (key, int((~ok).sum()))
            steps += 1
        t.join()
        assert not errors, errors
        assert steps >= 3
    finally:
        store.close()
