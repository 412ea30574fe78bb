"""Hand-made sparse PodSpec tiles that the fleets never produce: empty tiles (zero-length runs), full 192-slot
tiles (runs of 96 vectors), differences in the first / last / only vector of a run, masks that differ with equal
values — against a numpy reconstruction of the dense tiles, the packed oracle and (GPU) both kernel forms."""
import ctypes as C
import importlib

import numpy as np
import pytest

import oracle as O
import spans

pkg = O.pkg
N = pkg._native


class Synthetic:
    kind = "StsSparseBatch"
    mode = 0

    def __init__(self, n, seed):
        rng = np.random.default_rng(seed)
        self.n = n
        dense_f = np.zeros((n, 192), dtype=np.uint64)
        for i in range(n):
            cls = rng.integers(0, 5)
            fill = [0, 1, 7, 60, 192][cls]
            slots = rng.choice(192, size=fill, replace=False)
            dense_f[i, slots] = rng.integers(1, 2 ** 63, size=fill, dtype=np.uint64)
        dense_t = dense_f.copy()
        self.want_pod = np.zeros(n, dtype=bool)
        for i in range(n):
            nz = np.flatnonzero(dense_f[i])
            m = rng.integers(0, 7)
            if m == 1 and len(nz):                       # first value of the run
                dense_t[i, nz[0]] ^= np.uint64(1)
            elif m == 2 and len(nz):                     # last value of the run
                dense_t[i, nz[-1]] ^= np.uint64(1 << 40)
            elif m == 3 and len(nz):                     # a middle one
                dense_t[i, nz[len(nz) // 2]] += np.uint64(2)
            elif m == 4:                                 # a slot appears (mask differs)
                z = np.flatnonzero(dense_f[i] == 0)
                if len(z):
                    dense_t[i, z[0]] = np.uint64(5)
            elif m == 5 and len(nz):                     # a slot disappears
                dense_t[i, nz[0]] = np.uint64(0)
            self.want_pod[i] = not np.array_equal(dense_f[i], dense_t[i])
        self._a = {"replicas_from": np.ones(n, np.int32), "replicas_to": np.ones(n, np.int32),
                   "tmpl_from": np.zeros(n, np.uint64), "tmpl_to": np.zeros(n, np.uint64)}
        for side, dense in (("pod_from", dense_f), ("pod_to", dense_t)):
            mask = np.zeros((n, 3), dtype=np.uint64)
            off, vals = [0], []
            for i in range(n):
                nz = np.flatnonzero(dense[i])
                for s in nz:
                    mask[i, s >> 6] |= np.uint64(1) << np.uint64(s & 63)
                run = list(dense[i, nz])
                if len(run) & 1:
                    run.append(0)                        # runs are zero-padded to an even count
                vals += run
                off.append(off[-1] + len(run))
            self._a[side + ".mask"] = mask.reshape(-1)
            self._a[side + ".voff"] = np.array(off, dtype=np.uint32)
            self._a[side + ".val"] = np.array(vals + [0, 0], dtype=np.uint64)      # 16-byte padded
        for name in ("labels_from", "labels_to", "ann_from", "ann_to"):
            self._a[name + ".off"] = np.zeros(n + 1, dtype=np.uint32)
            self._a[name + ".ent"] = np.zeros(2, dtype=np.uint64)
        self.view = N.batch_struct(self.kind, lambda nm, ct: self._a[nm].ctypes.data_as(C.POINTER(ct)))

    def __len__(self):
        return self.n

    def arrays(self):
        return self._a


@pytest.mark.parametrize("n,seed", [(1, 1), (31, 2), (32, 3), (33, 4), (1000, 5)])
def test_packed_oracle_on_synthetic_tiles(n, seed):
    b = Synthetic(n, seed)
    reason = O.packed_diff_sts_sparse(b)[0]
    assert np.array_equal((reason & N.STS_PODSPEC) != 0, b.want_pod)
    assert not (reason & ~np.uint8(N.STS_PODSPEC)).any()


@pytest.fixture(scope="module")
def ctx():
    c = pkg.Context()
    yield c
    c.close()


@pytest.mark.gpu
@pytest.mark.parametrize("n,seed", [(1, 1), (31, 2), (32, 3), (33, 4), (64, 6), (4099, 5)])
def test_kernels_on_synthetic_tiles(ctx, n, seed):
    devmod = importlib.import_module("odh-kubeflow_b200.device")
    b = Synthetic(n, seed)
    want = np.where(b.want_pod, N.STS_PODSPEC, 0).astype(np.uint8)
    got = devmod.DeviceBatch.from_host(b).run(ctx)[0].cpu().numpy()           # CSR form: block sweep
    assert np.array_equal(got, want)
    got = devmod.DeviceBatch.from_host(spans.SpanBatch(b, seed=seed)).run(ctx)[0].cpu().numpy()   # spans form
    assert np.array_equal(got, want)
    r, _ = ctx.diff_sts_sparse(b)                                             # host-buffer call (chunked)
    assert np.array_equal(r, want)
