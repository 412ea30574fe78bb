"""Thin Python face of the C ABI (include/nbgraft.h).  Every method ends in a
CUDA kernel of libnbgraft.so; nothing here computes a predicate itself."""
import ctypes as C

import numpy as np

from . import _native as N


class NbgError(RuntimeError):
    def __init__(self, status, where):
        lib = N.nbg()
        msg = lib.nbg_strerror(status).decode()
        cuda = lib.nbg_last_cuda_error().decode()
        super().__init__(f"{where}: {msg} (status {status})" + (f" [{cuda}]" if cuda and status == N.NBG_E_CUDA else ""))
        self.status = status


def _check(status, where):
    if status != N.NBG_OK:
        raise NbgError(status, where)


def hash64(s):
    b = s.encode() if isinstance(s, str) else bytes(s)
    return N.nbg().nbg_hash64(b, len(b))


def shard_of(ns_hash, n_shards):
    return N.nbg().nbg_shard_of(ns_hash, n_shards)


class Context:
    """nbg_ctx: one or more CUDA devices.  Raises NbgError(NBG_E_NODEVICE) when
    there is no GPU — there is deliberately no CPU fallback."""

    def __init__(self, devices=None):
        self._lib = N.nbg()
        self._h = C.c_void_p()
        if devices is None:
            rc = self._lib.nbg_init(None, 0, C.byref(self._h))
        else:
            arr = (C.c_int * len(devices))(*devices)
            rc = self._lib.nbg_init(arr, len(devices), C.byref(self._h))
        _check(rc, "nbg_init")

    def close(self):
        if self._h:
            self._lib.nbg_destroy(self._h)
            self._h = C.c_void_p()

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    @property
    def handle(self):
        return self._h

    def launch_count(self):
        return int(self._lib.nbg_launch_count(self._h))

    def index_of(self, cuda_ordinal):
        """Position of CUDA device `cuda_ordinal` in this context (what the nbg_*_device calls take as
        `device`); raises when the context was not created on that device."""
        cache = self.__dict__.setdefault("_index_of", {})
        if cuda_ordinal not in cache:
            for k in range(self._lib.nbg_device_count(self._h)):
                cache[self._lib.nbg_device_id(self._h, k)] = k
        if cuda_ordinal not in cache:
            raise ValueError("CUDA device %d is not part of this context" % cuda_ordinal)
        return cache[cuda_ordinal]

    # ---- host-buffer calls (H2D + kernel + D2H inside) -------------------
    def _diff(self, fn, batch, allow_nil=False):
        view = batch.view
        n = batch.n
        reason = np.zeros(n, dtype=np.uint8)
        bitmap = np.zeros((n + 63) // 64, dtype=np.uint64)
        rc = fn(self._h, C.byref(view), n, reason.ctypes.data_as(C.c_void_p), bitmap.ctypes.data_as(C.c_void_p))
        if not (allow_nil and rc == N.NBG_E_NILREPLICAS):
            _check(rc, fn.__name__)
        return reason, bitmap

    def diff_sts(self, batch, allow_nil=False):
        """Batched CopyStatefulSetFields (COMMON:107-134) + NBCTL:188-194."""
        return self._diff(self._lib.nbg_diff_sts, batch, allow_nil)

    def diff_sts_gen(self, batch, consts, allow_nil=False):
        """CopyStatefulSetFields against a desired StatefulSet generated on the device from the Notebook
        (batch = a host.StsBatch filled with add_gen; consts = host.sts_gen_consts())."""
        view, gen = batch.view, batch.gen_inputs
        n = batch.n
        reason = np.zeros(n, dtype=np.uint8)
        bitmap = np.zeros((n + 63) // 64, dtype=np.uint64)
        rc = self._lib.nbg_diff_sts_gen(self._h, C.byref(view), C.byref(gen), C.byref(consts), n,
                                        reason.ctypes.data_as(C.c_void_p), bitmap.ctypes.data_as(C.c_void_p))
        if not (allow_nil and rc == N.NBG_E_NILREPLICAS):
            _check(rc, "nbg_diff_sts_gen")
        return reason, bitmap

    def diff_deploy(self, batch):
        """Batched CopyDeploymentSetFields (COMMON:136-163) on a StsBatch of Deployment pairs: Replicas compared as
        pointers (different unless both nil), no panic, no template-label pre-step."""
        return self._diff(self._lib.nbg_diff_deploy, batch, True)

    def diff_sts_sparse(self, batch, allow_nil=False):
        """Same predicate on sparsely stored PodSpec tiles (batch = StsBatch.sparse())."""
        return self._diff(self._lib.nbg_diff_sts_sparse, batch, allow_nil)

    def diff_generic(self, batch):
        """The ODH comparator family on a host.GenBatch (reason bit0 labels, bit1 spec/data differ)."""
        view = batch.view
        n = batch.n
        reason = np.zeros(n, dtype=np.uint8)
        bitmap = np.zeros((n + 63) // 64, dtype=np.uint64)
        _check(self._lib.nbg_diff_generic(self._h, C.byref(view), n, batch.mode, reason.ctypes.data_as(C.c_void_p),
                                          bitmap.ctypes.data_as(C.c_void_p)), "nbg_diff_generic")
        return reason, bitmap

    def diff_svc_gen(self, gen_view):
        """CopyServiceFields against a desired Service generated on the device (batch = SvcBatch.gen())."""
        return self._diff(self._lib.nbg_diff_svc_gen, gen_view)

    def diff_vs_gen(self, gen_view, consts):
        """CopyVirtualService against a desired VirtualService generated on the device (VsBatch.gen())."""
        view = gen_view.view
        n = gen_view.n
        reason = np.zeros(n, dtype=np.uint8)
        bitmap = np.zeros((n + 63) // 64, dtype=np.uint64)
        _check(self._lib.nbg_diff_vs_gen(self._h, C.byref(view), C.byref(consts), n, reason.ctypes.data_as(C.c_void_p),
                                         bitmap.ctypes.data_as(C.c_void_p)), "nbg_diff_vs_gen")
        return reason, bitmap

    def step(self, sts_sparse, svc_gen, vs_gen, vs_consts, cull, now_sec=0, now_nsec=0, cull_idle_min=1440,
             check_period_min=1, mode=N.CULL_RECONCILE):
        """nbg_step: all four predicates over the same notebooks as one pipeline (cull=None: the three
        Copy*Fields predicates only).
        Returns {"sts": (reason, bitmap), "svc": ..., "vs": ..., "cull": (action, argmax, cull_bm, period_bm)}."""
        n = len(sts_sparse)
        words = (n + 63) // 64
        r = {k: (np.zeros(n, np.uint8), np.zeros(words, np.uint64)) for k in ("sts", "svc", "vs")}
        action, argmax = np.zeros(n, np.uint8), np.zeros(n, np.uint32)
        cb, pb = np.zeros(words, np.uint64), np.zeros(words, np.uint64)
        o = N.nbg_step_outputs(*[a.ctypes.data for a in (r["sts"][0], r["sts"][1], r["svc"][0], r["svc"][1], r["vs"][0],
                                                         r["vs"][1], action, argmax, cb, pb)])
        p = N.nbg_cull_params(now_sec, now_nsec, cull_idle_min, check_period_min, mode)
        v1, v2, v3 = sts_sparse.view, svc_gen.view, vs_gen.view
        v4 = cull.view if cull is not None else None
        _check(self._lib.nbg_step(self._h, C.byref(v1), C.byref(v2), C.byref(v3), C.byref(vs_consts),
                                  C.byref(v4) if v4 is not None else None, C.byref(p), n, C.byref(o)), "nbg_step")
        if cull is not None:
            r["cull"] = (action, argmax, cb, pb)
        return r

    def step3(self, sts_sparse, svc_gen, vs_gen, vs_consts):
        """The three Copy*Fields predicates through nbg_step (no culler batch)."""
        return self.step(sts_sparse, svc_gen, vs_gen, vs_consts, None)

    def restart_block(self, batch):
        """Batched maybeRestartRunningNotebook (ODHWH:505-564) on a host.RestartBatch."""
        view = batch.view
        n = batch.n
        decision = np.zeros(n, dtype=np.uint8)
        bitmap = np.zeros((n + 63) // 64, dtype=np.uint64)
        _check(self._lib.nbg_restart_block(self._h, C.byref(view), n, decision.ctypes.data_as(C.c_void_p),
                                           bitmap.ctypes.data_as(C.c_void_p)), "nbg_restart_block")
        return decision, bitmap

    def diff_svc(self, batch):
        """Batched CopyServiceFields (COMMON:166-195)."""
        return self._diff(self._lib.nbg_diff_svc, batch)

    def diff_vs(self, batch):
        """Batched CopyVirtualService (COMMON:199-219)."""
        return self._diff(self._lib.nbg_diff_vs, batch)

    def cull(self, batch, now_sec, now_nsec, cull_idle_min=1440, check_period_min=1, mode=N.CULL_RECONCILE):
        """Batched culler decision (CULL:86-241,315-428)."""
        view = batch.view
        n = batch.n
        p = N.nbg_cull_params(now_sec, now_nsec, cull_idle_min, check_period_min, mode)
        action = np.zeros(n, dtype=np.uint8)
        argmax = np.zeros(n, dtype=np.uint32)
        words = (n + 63) // 64
        cull_bm = np.zeros(words, dtype=np.uint64)
        period_bm = np.zeros(words, dtype=np.uint64)
        _check(self._lib.nbg_cull(self._h, C.byref(view), n, C.byref(p), action.ctypes.data_as(C.c_void_p),
                                  argmax.ctypes.data_as(C.c_void_p), cull_bm.ctypes.data_as(C.c_void_p),
                                  period_bm.ctypes.data_as(C.c_void_p)), "nbg_cull")
        return action, argmax, cull_bm, period_bm


def _owner_join(self, nb_keys, sts_keys):
    """Batched NBCTL:157-170: match[i] = index of the first StatefulSet controlled by Notebook i
    (0xFFFFFFFF = none yet).  Keys from host.join_key_owner / join_key_controlled."""
    nb = np.ascontiguousarray(nb_keys, dtype=np.uint64)
    st = np.ascontiguousarray(sts_keys, dtype=np.uint64)
    match = np.empty(len(nb), dtype=np.uint32)
    _check(self._lib.nbg_owner_join(self._h, nb.ctypes.data_as(C.c_void_p), len(nb), st.ctypes.data_as(C.c_void_p),
                                    len(st), match.ctypes.data_as(C.c_void_p)), "nbg_owner_join")
    return match


Context.owner_join = _owner_join


def _count_running(self, ns_id, name_hash, label_hash, flags, n_namespaces):
    """Metrics.scrape (metrics.go:82-98) as a segmented count -> uint32[n_namespaces]."""
    ns = np.ascontiguousarray(ns_id, np.uint32)
    nm = np.ascontiguousarray(name_hash, np.uint64)
    lb = np.ascontiguousarray(label_hash, np.uint64)
    fl = np.ascontiguousarray(flags, np.uint8)
    counts = np.zeros(n_namespaces, np.uint32)
    _check(self._lib.nbg_count_running(self._h, ns.ctypes.data_as(C.c_void_p), nm.ctypes.data_as(C.c_void_p),
                                       lb.ctypes.data_as(C.c_void_p), fl.ctypes.data_as(C.c_void_p), len(ns), n_namespaces,
                                       counts.ctypes.data_as(C.c_void_p)), "nbg_count_running")
    return counts


Context.count_running = _count_running


class StringTable:
    """R strings as CSR (off u32[R+1], bytes padded to 16) — the layout of nbg_strings."""

    def __init__(self, strings):
        enc = [s if isinstance(s, (bytes, bytearray)) else s.encode() for s in strings]
        self.n = len(enc)
        lens = np.fromiter((len(b) for b in enc), dtype=np.int64, count=self.n)
        self.off = np.zeros(self.n + 1, dtype=np.uint32)
        np.cumsum(lens, out=self.off[1:])
        total = int(self.off[-1])
        raw = np.zeros((total + 15) // 16 * 16 + 16, dtype=np.uint8)
        # 16-byte aligned start inside the numpy buffer
        shift = (-raw.ctypes.data) % 16
        self._raw = raw
        self.bytes = raw[shift:shift + (total + 15) // 16 * 16]
        self.bytes[:total] = np.frombuffer(b"".join(enc), dtype=np.uint8)
        self.view = N.nbg_strings(self.off.ctypes.data_as(N.u32p), self.bytes.ctypes.data_as(N.u8p))


def _parse_records(self, last_activity, states=None):
    """time.Parse(RFC3339) + execution_state == "idle" for every record, on the GPU.
    Returns (sec int64[R], nsf uint32[R], zone int32[R])."""
    la = last_activity if isinstance(last_activity, StringTable) else StringTable(last_activity)
    st = None if states is None else (states if isinstance(states, StringTable) else StringTable(states))
    R = la.n
    sec, nsf, zone = np.zeros(R, np.int64), np.zeros(R, np.uint32), np.zeros(R, np.int32)
    _check(self._lib.nbg_parse_records(self._h, C.byref(la.view), C.byref(st.view) if st else None, R,
                                       sec.ctypes.data_as(C.c_void_p), nsf.ctypes.data_as(C.c_void_p),
                                       zone.ctypes.data_as(C.c_void_p)), "nbg_parse_records")
    return sec, nsf, zone


def _hash_strings(self, strings):
    """nbg_hash64 of every string, on the GPU."""
    tab = strings if isinstance(strings, StringTable) else StringTable(strings)
    out = np.zeros(tab.n, np.uint64)
    _check(self._lib.nbg_hash_strings(self._h, C.byref(tab.view), tab.n, out.ctypes.data_as(C.c_void_p)), "nbg_hash_strings")
    return out


LA_KEY = "notebooks.kubeflow.org/last-activity"
CHK_KEY = "notebooks.kubeflow.org/last_activity_check_timestamp"
STOP_KEY = "kubeflow-resource-stopped"


class CullTextBatch:
    """Culler inputs as the reference handles them — strings — laid out for nbg_cull_text.
    notebooks: iterable of {"annotations": {...}|None, "kernels": [...], "terminals": [...], "pod_missing": bool}."""

    def __init__(self, notebooks):
        k_off, t_off, kla, kst, tla, ann, chk, flags = [0], [0], [], [], [], [], [], []
        for nb in notebooks:
            a = nb.get("annotations")
            f = 0
            la = ck = ""
            if a is not None:
                if LA_KEY in a:
                    f |= N.NB_LA_PRESENT
                    la = a[LA_KEY]
                if CHK_KEY in a:
                    f |= N.NBT_CHK_PRESENT
                    ck = a[CHK_KEY]
                if STOP_KEY in a:
                    f |= N.NBT_STOP_SET
            if nb.get("pod_missing"):
                f |= N.NB_POD_MISSING
            for k in nb.get("kernels") or []:
                kla.append(k.get("last_activity", ""))
                kst.append(k.get("execution_state", ""))
            for t in nb.get("terminals") or []:
                tla.append(t.get("last_activity", ""))
            k_off.append(len(kla))
            t_off.append(len(tla))
            ann.append(la)
            chk.append(ck)
            flags.append(f)
        self.n = len(flags)
        self.k_off, self.t_off = np.array(k_off, np.uint32), np.array(t_off, np.uint32)
        self.flags = np.array(flags, np.uint8)
        self.tables = [StringTable(x) for x in (kla, kst, tla, ann, chk)]
        self.kernel_texts, self.terminal_texts = kla, tla
        v = N.nbg_cull_text_batch()
        v.k_off, v.t_off = self.k_off.ctypes.data_as(N.u32p), self.t_off.ctypes.data_as(N.u32p)
        v.k_last_activity, v.k_state, v.t_last_activity, v.ann_text, v.chk_text = (t.view for t in self.tables)
        v.nb_flags = self.flags.ctypes.data_as(N.u8p)
        self.view = v

    def input_bytes(self):
        return (self.k_off.nbytes + self.t_off.nbytes + self.flags.nbytes +
                sum(t.off.nbytes + int(t.off[-1]) for t in self.tables))


def _cull_text(self, batch, now_sec, now_nsec, cull_idle_min=1440, check_period_min=1, mode=N.CULL_RECONCILE):
    """nbg_cull_text: the culler decision from raw strings; returns
    (action, argmax, argmax_zone, cull_bitmap, period_bitmap)."""
    n = batch.n
    p = N.nbg_cull_params(now_sec, now_nsec, cull_idle_min, check_period_min, mode)
    action, argmax, zone = np.zeros(n, np.uint8), np.zeros(n, np.uint32), np.zeros(n, np.int32)
    words = (n + 63) // 64
    cb, pb = np.zeros(words, np.uint64), np.zeros(words, np.uint64)
    _check(self._lib.nbg_cull_text(self._h, C.byref(batch.view), n, C.byref(p), action.ctypes.data_as(C.c_void_p),
                                   argmax.ctypes.data_as(C.c_void_p), zone.ctypes.data_as(C.c_void_p),
                                   cb.ctypes.data_as(C.c_void_p), pb.ctypes.data_as(C.c_void_p)), "nbg_cull_text")
    return action, argmax, zone, cb, pb


Context.cull_text = _cull_text
Context.parse_records = _parse_records
Context.hash_strings = _hash_strings


def bitmap_to_bools(bitmap, n):
    """uint64 words -> bool[n] (bit i of word i//64)."""
    bits = np.unpackbits(bitmap.view(np.uint8), bitorder="little")
    return bits[:n].astype(bool)


class StepBuffers:
    """Host result buffers of one step over n notebooks: {"sts"|"svc"|"vs": (reason, bitmap), "cull": (action, argmax,
    cull bitmap, period bitmap)}.  pinned=True takes them from nbg_alloc_pinned (the read-back then runs at PCIe speed)."""

    def __init__(self, n, pinned=True):
        words = (n + 63) // 64
        self._lib, self._ptrs = N.nbg(), []
        spec = [("sts_r", n, np.uint8), ("sts_b", words, np.uint64), ("svc_r", n, np.uint8), ("svc_b", words, np.uint64),
                ("vs_r", n, np.uint8), ("vs_b", words, np.uint64), ("act", n, np.uint8), ("amax", n, np.uint32),
                ("cb", words, np.uint64), ("pb", words, np.uint64)]
        a = {}
        for name, count, dt in spec:
            if pinned:
                nbytes = max(count * np.dtype(dt).itemsize, 1)
                p = self._lib.nbg_alloc_pinned(nbytes)
                if not p:
                    raise MemoryError("nbg_alloc_pinned(%d)" % nbytes)
                self._ptrs.append(p)
                a[name] = np.frombuffer((C.c_uint8 * nbytes).from_address(p), dtype=np.uint8, count=count * np.dtype(dt).itemsize).view(dt)
                a[name][...] = 0
            else:
                a[name] = np.zeros(count, dt)
        self.arrays = {"sts": (a["sts_r"], a["sts_b"]), "svc": (a["svc_r"], a["svc_b"]), "vs": (a["vs_r"], a["vs_b"]),
                       "cull": (a["act"], a["amax"], a["cb"], a["pb"])}

    def close(self):
        self.arrays = None
        for p in self._ptrs:
            self._lib.nbg_free_pinned(p)
        self._ptrs = []

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class Store:
    """A resident fleet (nbg_store_*): packed rows live in device memory, put() replaces the rows of given
    slots from host batches of as many objects, step() runs the four predicates over slots [0, n)."""

    def __init__(self, ctx, capacity, device=0):
        self._lib, self.ctx, self.capacity = N.nbg(), ctx, capacity
        h = C.c_void_p()
        _check(self._lib.nbg_store_create(ctx.handle, device, capacity, C.byref(h)), "nbg_store_create")
        self._h = h

    def close(self):
        if self._h:
            self._lib.nbg_store_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def put(self, idx, sts_sparse=None, svc_gen=None, vs_gen=None, cull=None):
        idx = np.ascontiguousarray(idx, dtype=np.uint64)
        views = [b.view if b is not None else None for b in (sts_sparse, svc_gen, vs_gen, cull)]   # keep alive over the call
        for b in (sts_sparse, svc_gen, vs_gen, cull):
            assert b is None or b.n == len(idx)            # .view has just set .n
        _check(self._lib.nbg_store_put(self._h, idx.ctypes.data_as(N.u64p), len(idx),
                                       *[C.byref(v) if v is not None else None for v in views]), "nbg_store_put")

    def step(self, n, vs_consts, now_sec, now_nsec, cull_idle_min=1440, check_period_min=1, mode=N.CULL_RECONCILE, out=None):
        """-> the same dict as Context.step().  out: a StepBuffers to write into (a caller that steps every tick keeps
        one, pinned, instead of allocating 9 bytes per notebook each time)."""
        r = out.arrays if out is not None else StepBuffers(n, pinned=False).arrays
        o = N.nbg_step_outputs(*[a.ctypes.data for a in (r["sts"][0], r["sts"][1], r["svc"][0], r["svc"][1], r["vs"][0],
                                                         r["vs"][1], r["cull"][0], r["cull"][1], r["cull"][2], r["cull"][3])])
        p = N.nbg_cull_params(now_sec, now_nsec, cull_idle_min, check_period_min, mode)
        _check(self._lib.nbg_store_step(self._h, n, C.byref(vs_consts), C.byref(p), C.byref(o)), "nbg_store_step")
        return r

    def step_device(self, n, vs_consts, params, outputs, stream):
        """Enqueue only: `outputs` is an nbg_step_outputs of DEVICE pointers, `stream` a cudaStream_t value."""
        _check(self._lib.nbg_store_step_device(self._h, n, C.byref(vs_consts), C.byref(params), C.byref(outputs),
                                               C.c_void_p(stream)), "nbg_store_step_device")

    def check(self, n):
        """nbg_store_check: fault bits of the store's own spans over slots [0, n) (0 = consistent)."""
        what = C.c_uint(0)
        rc = self._lib.nbg_store_check(self._h, n, C.byref(what))
        if rc not in (N.NBG_OK, N.NBG_E_RANGE):
            _check(rc, "nbg_store_check")
        return int(what.value)

    def stats(self):
        s = N.nbg_store_stats()
        _check(self._lib.nbg_store_get_stats(self._h, C.byref(s)), "nbg_store_get_stats")
        return {k: getattr(s, k) for k, _ in s._fields_}
