"""Test helper: the same batch with every variable-length part stored as SPANS instead of CSR
(include/nbgraft.h: `end` pointers) — runs scattered through a larger heap in random order with
garbage-filled gaps.  Results must not change."""
import ctypes as C

import numpy as np

import oracle as O

N = O.pkg._native

KV, SPARSE = N._KV_FIELDS, N._SPARSE_FIELDS


def _scatter(off, payloads, rng, align, width):
    """off: n+1 CSR offsets (in elements of `width` payload words); payloads: arrays sharing that index space.
    -> starts, ends, new payloads (runs in random order, random gaps, gaps filled with garbage)."""
    n = len(off) - 1
    lens = (off[1:] - off[:-1]).astype(np.int64)
    order = rng.permutation(n)
    starts = np.zeros(n, dtype=np.int64)
    at = int(rng.integers(0, 4)) * align
    for i in order:
        starts[i] = at
        at += int(lens[i]) + int(rng.integers(0, 3)) * align
        at = (at + align - 1) // align * align
    total = at + align
    outs = []
    for p in payloads:
        w = width if p.dtype == np.uint64 and width > 1 else 1
        heap = rng.integers(1, 2 ** 31, size=total * w).astype(p.dtype)       # garbage everywhere first
        for i in range(n):
            heap[starts[i] * w:(starts[i] + lens[i]) * w] = p[int(off[i]) * w:int(off[i + 1]) * w]
        outs.append(heap)
    return starts.astype(np.uint32), (starts + lens).astype(np.uint32), outs


class SpanBatch:
    """Looks like a host batch to device.DeviceBatch and to struct_over()."""

    def __init__(self, batch, seed=1):
        rng = np.random.default_rng(seed)
        self.kind = getattr(batch, "kind", type(batch).__name__)
        self.mode = getattr(batch, "mode", 0)
        self.n = len(batch)
        a = {k: np.array(v, copy=True) for k, v in batch.arrays().items()}
        base0 = lambda off: off - off[0]
        for name in KV[self.kind]:
            off = a[name + ".off"].astype(np.int64)
            s, e, (ent,) = _scatter(off, [a[name + ".ent"]], rng, 1, 2)
            a[name + ".off"], a[name + ".end"], a[name + ".ent"] = s, e, ent
        for side in SPARSE.get(self.kind, ()):
            off = a[side + ".voff"].astype(np.int64)
            s, e, (val,) = _scatter(off, [a[side + ".val"]], rng, 2, 1)        # runs stay 16-byte aligned
            a[side + ".voff"], a[side + ".vend"], a[side + ".val"] = s, e, val
        if self.kind == "CullBatch":
            for p in ("k", "t"):
                off = a[p + "_off"].astype(np.int64)
                s, e, (sec, nsf) = _scatter(off, [a[p + "_sec"], a[p + "_nsf"]], rng, 1, 1)
                a[p + "_off"], a[p + "_end"], a[p + "_sec"], a[p + "_nsf"] = s, e, sec, nsf
        self._a = a

    def __len__(self):
        return self.n

    def arrays(self):
        return self._a


def struct_over(sb):
    """The nbg_* batch struct over the HOST arrays of a SpanBatch (for the packed oracle)."""
    a = sb.arrays()
    return N.batch_struct(sb.kind, lambda name, ct: a[name].ctypes.data_as(C.POINTER(ct)), lambda name: name in a)


class _Viewed:
    """Adapter so that tests/oracle.py's packed_* wrappers accept a SpanBatch."""

    def __init__(self, sb):
        self._sb, self.n, self.view = sb, sb.n, struct_over(sb)
        self.mode = sb.mode


def for_oracle(sb):
    return _Viewed(sb)
