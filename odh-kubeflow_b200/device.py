"""Batches resident in HBM.  torch is used here for what it is good at — device
memory and streams; the compute is the nbg_*_device entry points of the C ABI,
launched on torch's current stream so torch.cuda.Event brackets them."""
import ctypes as C

import numpy as np
import torch

from . import _native as N
from . import host


def _to_dev(a, device, tile=1, add=None):
    t = torch.from_numpy(np.ascontiguousarray(a).view(np.uint8).copy()).to(f"cuda:{device}")
    if tile > 1:
        t = t.repeat(tile)
    return t


class DeviceBatch:
    """Uploads a host batch (optionally tiled `tile` times to build a larger fleet out of a
    base fleet) and launches the matching kernel with device pointers."""

    def __init__(self):
        self.tensors = {}

    @classmethod
    def from_host(cls, batch, device=0, tile=1):
        self = cls()
        self.device, self.kind = device, getattr(batch, "kind", type(batch).__name__)
        self.mode = getattr(batch, "mode", 0)
        arrays = batch.arrays()
        n0 = len(batch)
        self.n = n0 * tile
        self.in_bytes = 0
        for name, a in arrays.items():
            if tile > 1 and (name.endswith(".off") or name.endswith(".voff") or name in ("k_off", "t_off")):
                # CSR offsets of tile k are the base offsets shifted by k * (#entries)
                base = a.astype(np.int64)
                total = int(base[-1])
                parts = [base[:-1] + k * total for k in range(tile)] + [np.array([tile * total])]
                off = np.concatenate(parts)
                assert off[-1] < 2 ** 32
                t = torch.from_numpy(off.astype(np.uint32).view(np.uint8).copy()).to(f"cuda:{device}")
            else:
                t = _to_dev(a, device, tile)
            self.tensors[name] = t
            self.in_bytes += t.numel()
        words = (self.n + 63) // 64
        dev = f"cuda:{device}"
        if self.kind == "CullBatch":
            self.outs = [torch.empty(self.n, dtype=torch.uint8, device=dev), torch.empty(self.n * 4, dtype=torch.uint8, device=dev),
                         torch.empty(words * 8, dtype=torch.uint8, device=dev), torch.empty(words * 8, dtype=torch.uint8, device=dev)]
        else:
            self.outs = [torch.empty(self.n, dtype=torch.uint8, device=dev), torch.empty(words * 8, dtype=torch.uint8, device=dev)]
        self.out_bytes = sum(o.numel() for o in self.outs)
        self._struct = self._build_struct()
        return self

    @property
    def algorithmic_bytes(self):
        return self.in_bytes + self.out_bytes

    def _build_struct(self):
        def ptr(name, ctype):
            return C.cast(self.tensors[name].data_ptr(), C.POINTER(ctype))
        return N.batch_struct(self.kind, ptr, lambda name: name in self.tensors)

    def run(self, ctx, now=None, cull_idle_min=1440, check_period_min=1, mode=N.CULL_RECONCILE, stream=None):
        """Enqueue the kernel (asynchronous).  Returns the device output tensors."""
        lib = N.nbg()
        if stream is None:
            stream = torch.cuda.current_stream(self.device).cuda_stream
        di = ctx.index_of(self.device)        # the C ABI wants the context's index of the device the batch lives on
        o = [C.c_void_p(t.data_ptr()) for t in self.outs]
        if self.kind == "CullBatch":
            now = now or (host.NOW_SEC, host.NOW_NSEC)
            p = N.nbg_cull_params(now[0], now[1], cull_idle_min, check_period_min, mode)
            rc = lib.nbg_cull_device(ctx.handle, di, C.byref(self._struct), self.n, C.byref(p), o[0], o[1], o[2], o[3],
                                     C.c_void_p(stream))
        elif self.kind == "svc_gen":
            rc = lib.nbg_diff_svc_gen_device(ctx.handle, di, C.byref(self._struct), self.n, o[0], o[1], C.c_void_p(stream))
        elif self.kind == "vs_gen":
            from . import host as _host
            if not hasattr(self, "_consts"):
                self._consts = _host.vs_gen_consts()
            rc = lib.nbg_diff_vs_gen_device(ctx.handle, di, C.byref(self._struct), C.byref(self._consts), self.n, o[0], o[1],
                                            C.c_void_p(stream))
        elif self.kind == "sts_gen":
            if not hasattr(self, "_consts"):
                self._consts = host.sts_gen_consts()
                self._gen = N.nbg_sts_gen_inputs(C.cast(self.tensors["gen.prefix_hash"].data_ptr(), N.u64p),
                                                 C.cast(self.tensors["gen.flags"].data_ptr(), N.u8p))
            rc = lib.nbg_diff_sts_gen_device(ctx.handle, di, C.byref(self._struct), C.byref(self._gen), C.byref(self._consts),
                                             self.n, o[0], o[1], C.c_void_p(stream))
        elif self.kind == "GenBatch":
            rc = lib.nbg_diff_generic_device(ctx.handle, di, C.byref(self._struct), self.n, self.mode, o[0], o[1], C.c_void_p(stream))
        elif self.kind == "RestartBatch":
            rc = lib.nbg_restart_block_device(ctx.handle, di, C.byref(self._struct), self.n, o[0], o[1], C.c_void_p(stream))
        else:
            fn = {"StsBatch": lib.nbg_diff_sts_device, "StsSparseBatch": lib.nbg_diff_sts_sparse_device,
                  "SvcBatch": lib.nbg_diff_svc_device,
                  "VsBatch": lib.nbg_diff_vs_device}[self.kind]
            rc = fn(ctx.handle, di, C.byref(self._struct), self.n, o[0], o[1], C.c_void_p(stream))
        if rc != N.NBG_OK:
            from .engine import NbgError
            raise NbgError(rc, "nbg_*_device")
        return self.outs


class ResidentStep:
    """One reconcile step over resident batches (nbg_step_device): `sts`, `svc`, `vs`, `cull` are
    DeviceBatches of kinds StsSparseBatch, svc_gen, vs_gen and CullBatch over the same notebooks (any
    may be None).  run() enqueues the step on `stream` (default: torch's current stream); results land
    in each batch's own output tensors.  The argument structs are built once, so run() is one C call."""

    def __init__(self, ctx, sts=None, svc=None, vs=None, cull=None, now=None, cull_idle_min=1440, check_period_min=1,
                 mode=N.CULL_RECONCILE, time_sts=False):
        given = [b for b in (sts, svc, vs, cull) if b is not None]
        assert given and len({b.n for b in given}) == 1 and len({b.device for b in given}) == 1
        assert sts is None or sts.kind == "StsSparseBatch"
        assert svc is None or svc.kind == "svc_gen"
        assert vs is None or vs.kind == "vs_gen"
        assert cull is None or cull.kind == "CullBatch"
        self.ctx, self.n, self.device = ctx, given[0].n, given[0].device
        self.batches = {"sts": sts, "svc": svc, "vs": vs, "cull": cull}
        self._consts = host.vs_gen_consts() if vs is not None else None
        now = now or (host.NOW_SEC, host.NOW_NSEC)
        self._p = N.nbg_cull_params(now[0], now[1], cull_idle_min, check_period_min, mode)
        o = N.nbg_step_outputs()
        if sts is not None:
            o.sts_reason, o.sts_bitmap = sts.outs[0].data_ptr(), sts.outs[1].data_ptr()
        if svc is not None:
            o.svc_reason, o.svc_bitmap = svc.outs[0].data_ptr(), svc.outs[1].data_ptr()
        if vs is not None:
            o.vs_reason, o.vs_bitmap = vs.outs[0].data_ptr(), vs.outs[1].data_ptr()
        if cull is not None:
            o.action, o.argmax, o.cull_bitmap, o.period_bitmap = (t.data_ptr() for t in cull.outs)
        self._o = o
        self._refs = [C.byref(b._struct) if b is not None else None for b in (sts, svc, vs, cull)]
        self.launches = len(given)
        self._di = ctx.index_of(self.device)  # the context's index of the device the batches live on
        # optional timing of the StatefulSet kernel inside the running step: pairs of events the
        # library records around that launch, read back with sts_ms()
        self._pairs, self._free = [], []
        self._time_sts = time_sts and sts is not None

    def reserve(self, k):
        """Create k event pairs ahead of time, so that timed run() calls create nothing."""
        with torch.cuda.device(self.device):
            while len(self._free) < k:
                a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                a.record(); b.record()                     # materialise the handles (the library re-records them)
                self._free.append((a, b))

    def run(self, stream=None):
        if stream is None:
            stream = torch.cuda.current_stream(self.device).cuda_stream
        e0 = e1 = None
        if self._time_sts:
            if not self._free:
                self.reserve(1)
            a, b = self._free.pop()
            self._pairs.append((a, b))
            e0, e1 = C.c_void_p(a.cuda_event), C.c_void_p(b.cuda_event)
        rc = N.nbg().nbg_step_device(self.ctx.handle, self._di, self._refs[0], self._refs[1], self._refs[2],
                                     C.byref(self._consts) if self._consts is not None else None, self._refs[3],
                                     C.byref(self._p), self.n, C.byref(self._o), C.c_void_p(stream), e0, e1)
        if rc != N.NBG_OK:
            from .engine import NbgError
            raise NbgError(rc, "nbg_step_device")
        return {k: b.outs for k, b in self.batches.items() if b is not None}

    def sts_ms(self):
        """Durations (ms) of the StatefulSet launches of the timed run() calls so far; call after a sync."""
        out = [a.elapsed_time(b) for a, b in self._pairs]
        self._free.extend(self._pairs)
        self._pairs = []
        return out


def check_step(ctx, sts=None, svc=None, vs=None, cull=None, stream=None):
    """nbg_check_step_device on resident batches: the argument checks a host-buffer call would have made (offsets
    monotone and inside their payload arrays, sparse runs aligned and as long as their masks say).  The payload
    lengths are taken from the tensors the batches own.  Returns the fault bits (0 = fine)."""
    given = [b for b in (sts, svc, vs, cull) if b is not None]
    assert given and len({b.n for b in given}) == 1 and len({b.device for b in given}) == 1
    device = given[0].device
    if stream is None:
        stream = torch.cuda.current_stream(device).cuda_stream
    ext = N.nbg_step_extents()

    def elems(b, name, size):
        return b.tensors[name].numel() // size if b is not None else 0
    ext.sts_val_from, ext.sts_val_to = elems(sts, "pod_from.val", 8), elems(sts, "pod_to.val", 8)
    for field, (b, name) in {"sts_labels_from": (sts, "labels_from.ent"), "sts_labels_to": (sts, "labels_to.ent"),
                             "sts_ann_from": (sts, "ann_from.ent"), "sts_ann_to": (sts, "ann_to.ent"),
                             "svc_labels_to": (svc, "labels_to.ent"), "svc_ann_to": (svc, "ann_to.ent"),
                             "svc_sel_to": (svc, "sel_to.ent"), "vs_hdr_from": (vs, "hdr_from.ent"),
                             "vs_hdr_to": (vs, "hdr_to.ent")}.items():
        setattr(ext, field, elems(b, name, 16))
    ext.cull_kernels, ext.cull_terminals = elems(cull, "k_sec", 8), elems(cull, "t_sec", 8)
    what = C.c_uint(0)
    refs = [C.byref(b._struct) if b is not None else None for b in (sts, svc, vs, cull)]
    rc = N.nbg().nbg_check_step_device(ctx.handle, ctx.index_of(device), refs[0], refs[1], refs[2], refs[3], given[0].n,
                                       C.byref(ext), C.c_void_p(stream), C.byref(what))
    if rc not in (N.NBG_OK, N.NBG_E_RANGE):
        from .engine import NbgError
        raise NbgError(rc, "nbg_check_step_device")
    return int(what.value)
