"""ctypes binding of oracle/_build/liboracle.so — the CPU checker (test
infrastructure: only tests/, smoke() and bench.py's CPU legs import this)."""
import ctypes as C
import importlib
import json
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
pkg = importlib.import_module("odh-kubeflow_b200")
N = pkg._native
host = pkg.host

_lib = None


def build():
    subprocess.run(["make", "-s", "-C", os.path.join(ROOT, "oracle")], check=True)


def lib():
    global _lib
    if _lib is None:
        path = os.path.join(ROOT, "oracle", "_build", "liboracle.so")
        if not os.path.exists(path):
            build()
        L = C.CDLL(path)
        vp, cp = C.c_void_p, C.c_char_p
        L.orc_deep_equal.argtypes = [vp, vp]
        L.orc_strict_equal.argtypes = [vp, vp]
        L.orc_copy_statefulset_fields.argtypes = [vp, vp, C.c_int]
        L.orc_copy_service_fields.argtypes = [vp, vp]
        L.orc_copy_deployment_fields.argtypes = [vp, vp]
        L.orc2_diff_deploy.argtypes = [C.POINTER(N.nbg_sts_batch), C.c_uint64, vp, vp]
        L.orc2_diff_deploy.restype = None
        L.orc_copy_virtualservice.argtypes = [vp, vp]
        L.orc_bench_pairs.argtypes = [C.c_int, C.POINTER(vp), C.POINTER(vp), C.c_uint64, C.c_int, vp]
        L.orc_bench_pairs.restype = C.c_double
        L.orc_parse_rfc3339.argtypes = [cp, C.POINTER(C.c_int64), C.POINTER(C.c_int32), C.POINTER(C.c_int32)]
        L.orc_format_rfc3339.argtypes = [C.c_int64, C.c_int32, cp, C.c_size_t]
        L.orc_all_kernels_are_idle.argtypes = [cp]
        L.orc_stop_annotation_is_set.argtypes = [cp]
        L.orc_notebook_is_idle.argtypes = [cp, C.c_int64, C.c_int32, C.c_int]
        L.orc_culling_check_period_has_passed.argtypes = [cp, C.c_int64, C.c_int32, C.c_int]
        L.orc_cull.argtypes = [cp, C.c_int64, C.c_int32, C.c_int32, C.c_int, C.c_int, C.c_int, C.POINTER(C.c_uint32),
                               cp, C.c_size_t]
        L.orc_bench_cull.argtypes = [C.POINTER(cp), C.c_uint64, C.c_int64, C.c_int32, C.c_int, C.c_int, C.c_int,
                                     C.c_int, vp, vp]
        L.orc_bench_cull.restype = C.c_double
        for name, bt in (("orc2_diff_sts", N.nbg_sts_batch), ("orc2_diff_svc", N.nbg_svc_batch),
                         ("orc2_diff_vs", N.nbg_vs_batch), ("orc2_diff_sts_sparse", N.nbg_sts_sparse_batch)):
            f = getattr(L, name)
            f.argtypes = [C.POINTER(bt), C.c_uint64, C.c_int, vp, vp]
            f.restype = C.c_double
        L.orc2_generate_sts_tile.argtypes = [vp, C.c_uint64, C.c_uint, C.POINTER(N.nbg_sts_gen_consts), vp]
        L.orc2_generate_sts_tile.restype = None
        L.orc2_diff_sts_gen.argtypes = [C.POINTER(N.nbg_sts_batch), C.POINTER(N.nbg_sts_gen_inputs), C.POINTER(N.nbg_sts_gen_consts),
                                        C.c_uint64, vp, vp]
        L.orc2_diff_sts_gen.restype = None
        L.orc_counts_as_running.argtypes = [vp]
        L.orc_semantic_equal.argtypes = [vp, vp]
        L.orc_restart_block.argtypes = [vp, vp, vp, C.c_int]
        L.orc2_restart_block.argtypes = [C.POINTER(N.nbg_restart_batch), C.c_uint64, vp, vp]
        L.orc2_restart_block.restype = None
        L.orc_compare_generic.argtypes = [vp, vp, cp, cp, C.c_int]
        L.orc2_diff_generic.argtypes = [C.POINTER(N.nbg_generic_batch), C.c_uint64, C.c_int, vp, vp]
        L.orc2_diff_generic.restype = None
        L.orc_find_controlled.argtypes = [vp, C.POINTER(vp), C.c_uint64]
        L.orc_find_controlled.restype = C.c_long
        L.orc2_owner_join.argtypes = [vp, C.c_uint64, vp, C.c_uint64, vp]
        L.orc2_owner_join.restype = C.c_double
        L.orc2_cull.argtypes = [C.POINTER(N.nbg_cull_batch), C.c_uint64, C.POINTER(N.nbg_cull_params), C.c_int, vp, vp,
                                vp, vp]
        L.orc2_cull.restype = C.c_double
        _lib = L
    return _lib


# ---- L1: semantic oracle on objects -----------------------------------------
def deep_equal(a, b):
    return bool(lib().orc_deep_equal(a._h, b._h))


def copy_statefulset_fields(from_, to, prestep=1):
    """returns reason bits (NBG_STS_*); mutates `to` (a host.Obj) like the Go code."""
    return lib().orc_copy_statefulset_fields(from_._h, to._h, prestep)


def copy_deployment_fields(from_, to):
    """CopyDeploymentSetFields (COMMON:136-163) on two Deployment objects; mutates `to`."""
    return lib().orc_copy_deployment_fields(from_._h, to._h)


def packed_diff_deploy(batch):
    view = batch.view
    n = batch.n
    reason = np.zeros(n, dtype=np.uint8)
    bitmap = np.zeros((n + 63) // 64, dtype=np.uint64)
    lib().orc2_diff_deploy(C.byref(view), n, reason.ctypes.data_as(C.c_void_p), bitmap.ctypes.data_as(C.c_void_p))
    return reason, bitmap


def copy_service_fields(from_, to):
    return lib().orc_copy_service_fields(from_._h, to._h)


def copy_virtualservice(from_, to):
    return lib().orc_copy_virtualservice(from_._h, to._h)


def parse_rfc3339(s):
    sec, nsec, off = C.c_int64(), C.c_int32(), C.c_int32()
    if lib().orc_parse_rfc3339(s.encode(), C.byref(sec), C.byref(nsec), C.byref(off)):
        return sec.value, nsec.value, off.value
    return None


def format_rfc3339(sec, off=0):
    buf = C.create_string_buffer(64)
    lib().orc_format_rfc3339(sec, off, buf, len(buf))
    return buf.value.decode()


def _j(nb):
    return (nb if isinstance(nb, str) else json.dumps(nb)).encode()


def all_kernels_are_idle(kernels):
    return bool(lib().orc_all_kernels_are_idle(_j({"kernels": kernels})))


def stop_annotation_is_set(meta):
    return bool(lib().orc_stop_annotation_is_set(_j({"annotations": (meta or {}).get("annotations")})))


def notebook_is_idle(meta, now, cull_idle_min):
    return bool(lib().orc_notebook_is_idle(_j({"annotations": (meta or {}).get("annotations")}), now[0], now[1],
                                           cull_idle_min))


def culling_check_period_has_passed(meta, now, period_min):
    return bool(lib().orc_culling_check_period_has_passed(_j({"annotations": (meta or {}).get("annotations")}),
                                                          now[0], now[1], period_min))


def cull(nb_json, now, cull_idle_min, period_min, mode, now_offset=0):
    """-> (action, argmax, new_last_activity_string)"""
    am = C.c_uint32()
    buf = C.create_string_buffer(64)
    a = lib().orc_cull(_j(nb_json), now[0], now[1], now_offset, cull_idle_min, period_min, mode, C.byref(am), buf,
                       len(buf))
    if a < 0:
        raise ValueError("bad culler input")
    return a, am.value, buf.value.decode()


def l1_diff_all(fleet, idx=None, chunk=20000, threads=8):
    """The object-level oracle (CopyStatefulSetFields / CopyServiceFields / CopyVirtualService restated on Go-shaped
    trees) on EVERY listed notebook of a fleet: {"sts"|"svc"|"vs": reason bytes}.  Objects are materialised a chunk at
    a time, so memory stays bounded at any fleet size."""
    idx = np.arange(fleet.n, dtype=np.uint64) if idx is None else np.ascontiguousarray(idx, dtype=np.uint64)
    out = {k: np.zeros(len(idx), np.uint8) for k in ("sts", "svc", "vs")}
    L, vp = lib(), C.c_void_p
    for lo in range(0, len(idx), chunk):
        rows = fleet.objects(idx=idx[lo:lo + chunk])
        m = len(rows)
        for kind, (key, a, b) in enumerate((("sts", "sts_from", "sts_to"), ("svc", "svc_from", "svc_to"), ("vs", "vs_from", "vs_to"))):
            fa = (vp * m)(*[r[a]._h.value for r in rows])
            ta = (vp * m)(*[r[b]._h.value for r in rows])
            L.orc_bench_pairs(kind, fa, ta, m, threads, out[key][lo:lo + m].ctypes.data_as(vp))
    return out


def l1_cull_all(cfleet, idx=None, mode=1, cull_idle_min=1440, period_min=1, threads=8):
    """The object-level culler oracle (decision flow on RFC3339 strings) on every listed notebook: (action, argmax)."""
    idx = np.arange(cfleet.n, dtype=np.uint64) if idx is None else np.ascontiguousarray(idx, dtype=np.uint64)
    texts = [cfleet.json(int(i)).encode() for i in idx]
    arr = (C.c_char_p * len(texts))(*texts)
    action, argmax = np.zeros(len(idx), np.uint8), np.zeros(len(idx), np.uint32)
    t = lib().orc_bench_cull(arr, len(texts), cfleet.now_sec, cfleet.now_nsec, cull_idle_min, period_min, mode, threads,
                             action.ctypes.data_as(C.c_void_p), argmax.ctypes.data_as(C.c_void_p))
    assert t >= 0, "a culler input did not decode"
    return action, argmax


# ---- L2: packed oracle on the SoA batches -------------------------------------
def _diff2(fn, batch, threads):
    view = batch.view
    n = batch.n
    reason = np.zeros(n, dtype=np.uint8)
    bitmap = np.zeros((n + 63) // 64, dtype=np.uint64)
    secs = fn(C.byref(view), n, threads, reason.ctypes.data_as(C.c_void_p), bitmap.ctypes.data_as(C.c_void_p))
    return reason, bitmap, secs


def packed_diff_sts(batch, threads=1):
    return _diff2(lib().orc2_diff_sts, batch, threads)


def packed_diff_sts_sparse(batch, threads=1):
    return _diff2(lib().orc2_diff_sts_sparse, batch, threads)


def generate_sts_tile(nb_tile, prefix_hash, flags, consts):
    """The packed oracle's generateStatefulSet: the Notebook's dense PodSpec tile -> the desired StatefulSet's."""
    nb_tile = np.ascontiguousarray(nb_tile, dtype=np.uint64)
    out = np.zeros(N.NBG_POD_SLOTS, np.uint64)
    lib().orc2_generate_sts_tile(nb_tile.ctypes.data_as(C.c_void_p), int(prefix_hash), int(flags), C.byref(consts),
                                 out.ctypes.data_as(C.c_void_p))
    return out


def packed_diff_sts_gen(batch, consts):
    view, gen = batch.view, batch.gen_inputs
    n = batch.n
    reason = np.zeros(n, dtype=np.uint8)
    bitmap = np.zeros((n + 63) // 64, dtype=np.uint64)
    lib().orc2_diff_sts_gen(C.byref(view), C.byref(gen), C.byref(consts), n, reason.ctypes.data_as(C.c_void_p),
                            bitmap.ctypes.data_as(C.c_void_p))
    return reason, bitmap


def packed_diff_svc(batch, threads=1):
    return _diff2(lib().orc2_diff_svc, batch, threads)


def packed_diff_vs(batch, threads=1):
    return _diff2(lib().orc2_diff_vs, batch, threads)


def packed_cull(batch, now, cull_idle_min=1440, period_min=1, mode=1, threads=1):
    view = batch.view
    n = batch.n
    p = N.nbg_cull_params(now[0], now[1], cull_idle_min, period_min, mode)
    action = np.zeros(n, dtype=np.uint8)
    argmax = np.zeros(n, dtype=np.uint32)
    words = (n + 63) // 64
    cb, pb = np.zeros(words, dtype=np.uint64), np.zeros(words, dtype=np.uint64)
    secs = lib().orc2_cull(C.byref(view), n, C.byref(p), threads, action.ctypes.data_as(C.c_void_p),
                           argmax.ctypes.data_as(C.c_void_p), cb.ctypes.data_as(C.c_void_p),
                           pb.ctypes.data_as(C.c_void_p))
    return action, argmax, cb, pb, secs


def find_controlled(owner, sts_list):
    """notebook_controller.go:157-170 on objects: index of the first StatefulSet in the owner's
    namespace whose controller reference carries the owner's UID, or -1."""
    arr = (C.c_void_p * len(sts_list))(*[o._h for o in sts_list])
    return lib().orc_find_controlled(owner._h, arr, len(sts_list))


def packed_owner_join(nb_keys, sts_keys):
    nb = np.ascontiguousarray(nb_keys, dtype=np.uint64)
    st = np.ascontiguousarray(sts_keys, dtype=np.uint64)
    match = np.empty(len(nb), dtype=np.uint32)
    secs = lib().orc2_owner_join(nb.ctypes.data_as(C.c_void_p), len(nb), st.ctypes.data_as(C.c_void_p), len(st),
                                 match.ctypes.data_as(C.c_void_p))
    return match, secs


def compare_generic(desired, found, labels_path="metadata.labels", spec_path="spec", mode=0):
    """L1: the ODH comparators on objects -> reason bits (1 labels differ, 2 spec/data differ)."""
    return lib().orc_compare_generic(desired._h, found._h, labels_path.encode(), spec_path.encode(), mode)


def packed_diff_generic(batch):
    view = batch.view
    n = batch.n
    reason = np.zeros(n, dtype=np.uint8)
    bitmap = np.zeros((n + 63) // 64, dtype=np.uint64)
    lib().orc2_diff_generic(C.byref(view), n, batch.mode, reason.ctypes.data_as(C.c_void_p), bitmap.ctypes.data_as(C.c_void_p))
    return reason, bitmap


def semantic_equal(a, b):
    """equality.Semantic.DeepEqual on objects (L1)."""
    return bool(lib().orc_semantic_equal(a._h, b._h))


def restart_block(old_nb, updated_nb, mutated_nb, create=False):
    """maybeRestartRunningNotebook on objects (L1) -> decision byte."""
    return lib().orc_restart_block(old_nb._h, updated_nb._h, mutated_nb._h, 1 if create else 0)


def packed_restart_block(batch):
    view = batch.view
    n = batch.n
    decision = np.zeros(n, dtype=np.uint8)
    bitmap = np.zeros((n + 63) // 64, dtype=np.uint64)
    lib().orc2_restart_block(C.byref(view), n, decision.ctypes.data_as(C.c_void_p), bitmap.ctypes.data_as(C.c_void_p))
    return decision, bitmap


def counts_as_running(sts):
    """metrics.go:90-93 on one StatefulSet object."""
    return bool(lib().orc_counts_as_running(sts._h))
