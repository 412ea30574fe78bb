"""Test helper: the same batch with every variable-length part stored as SPANS instead of CSR
(include/nbgraft.h: `end` pointers) — runs scattered through a larger heap in random order with
garbage-filled gaps.  Results must not change."""
import ctypes as C

import numpy as np

import oracle as O

N = O.pkg._native

KV = {"StsBatch": ("labels_from", "labels_to", "ann_from", "ann_to"),
      "StsSparseBatch": ("labels_from", "labels_to", "ann_from", "ann_to"),
      "SvcBatch": ("labels_from", "labels_to", "ann_from", "ann_to", "sel_from", "sel_to"),
      "VsBatch": ("hdr_from", "hdr_to"), "svc_gen": ("labels_to", "ann_to", "sel_to"), "vs_gen": ("hdr_from", "hdr_to"),
      "GenBatch": ("labels_a", "labels_b", "spec_a", "spec_b"), "RestartBatch": (), "CullBatch": ()}
SPARSE = {"StsSparseBatch": ("pod_from", "pod_to"), "RestartBatch": ("old_spec", "updated_spec", "mutated_spec")}


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
    a, k = sb.arrays(), sb.kind

    def p(name, ct):
        return a[name].ctypes.data_as(C.POINTER(ct))

    def kv(prefix):
        m = N.nbg_kvmap()
        m.off, m.end = p(prefix + ".off", C.c_uint32), p(prefix + ".end", C.c_uint32)
        m.ent = C.cast(a[prefix + ".ent"].ctypes.data, C.POINTER(N.nbg_kv))
        return m

    def sparse(side):
        ps = N.nbg_podsparse()
        ps.mask, ps.off, ps.end, ps.val = (p(side + ".mask", C.c_uint64), p(side + ".voff", C.c_uint32),
                                           p(side + ".vend", C.c_uint32), p(side + ".val", C.c_uint64))
        return ps
    if k == "StsSparseBatch":
        s = N.nbg_sts_sparse_batch()
        s.replicas_from, s.replicas_to = p("replicas_from", C.c_int32), p("replicas_to", C.c_int32)
        s.tmpl_from, s.tmpl_to = p("tmpl_from", C.c_uint64), p("tmpl_to", C.c_uint64)
        s.pod_from, s.pod_to = sparse("pod_from"), sparse("pod_to")
    elif k == "StsBatch":
        s = N.nbg_sts_batch()
        s.replicas_from, s.replicas_to = p("replicas_from", C.c_int32), p("replicas_to", C.c_int32)
        s.tmpl_from, s.tmpl_to = p("tmpl_from", C.c_uint64), p("tmpl_to", C.c_uint64)
        s.pod_from, s.pod_to = p("pod_from", C.c_uint64), p("pod_to", C.c_uint64)
    elif k == "SvcBatch":
        s = N.nbg_svc_batch()
        s.flags, s.ports_from, s.ports_to = p("flags", C.c_uint8), p("ports_from", C.c_uint64), p("ports_to", C.c_uint64)
    elif k == "VsBatch":
        s = N.nbg_vs_batch()
        s.flags, s.spec_from, s.spec_to = p("flags", C.c_uint8), p("spec_from", C.c_uint64), p("spec_to", C.c_uint64)
    elif k == "GenBatch":
        s = N.nbg_generic_batch()
    elif k == "RestartBatch":
        s = N.nbg_restart_batch()
        s.flags = p("flags", C.c_uint8)
        for side in SPARSE[k]:
            setattr(s, side, sparse(side))
    elif k == "CullBatch":
        s = N.nbg_cull_batch()
        for f, ct in (("k_off", C.c_uint32), ("k_end", C.c_uint32), ("k_sec", C.c_int64), ("k_nsf", C.c_uint32),
                      ("t_off", C.c_uint32), ("t_end", C.c_uint32), ("t_sec", C.c_int64), ("t_nsf", C.c_uint32),
                      ("ann_sec", C.c_int64), ("ann_nsf", C.c_uint32), ("chk_sec", C.c_int64), ("chk_nsf", C.c_uint32),
                      ("nb_flags", C.c_uint8)):
            setattr(s, f, p(f, ct))
    else:
        raise ValueError(k)
    for name in KV[k]:
        setattr(s, name, kv(name))
    return s


class _Viewed:
    """Adapter so that tests/oracle.py's packed_* wrappers accept a SpanBatch."""

    def __init__(self, sb):
        self._sb, self.n, self.view = sb, sb.n, struct_over(sb)
        self.mode = sb.mode


def for_oracle(sb):
    return _Viewed(sb)
