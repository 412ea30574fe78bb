"""Host side of the graft: objects, desired-object builders, packed batches and
synthetic fleets, all implemented in libnbgraft_host.so (C++).

Mirrors what the reference does in Go before its predicates run: decode objects,
build the desired StatefulSet/Service/VirtualService (NBCTL:433-658) and — new in
this build — flatten desired/live pairs into the structure-of-arrays batches of
include/nbgraft.h.
"""
import ctypes as C
import json

import numpy as np

from . import _native as N


class HostError(RuntimeError):
    pass


def _err():
    return C.create_string_buffer(512)


class Obj:
    """One decoded object (a Go value tree held by the native library)."""

    def __init__(self, handle):
        self._h = C.c_void_p(handle)

    @classmethod
    def from_json(cls, type_name, obj):
        text = obj if isinstance(obj, str) else json.dumps(obj)
        out, err = C.c_void_p(), _err()
        if N.nbh().nbh_obj_from_json(type_name.encode(), text.encode(), C.byref(out), err, len(err)) != 0:
            raise HostError(err.value.decode())
        return cls(out.value)

    def to_python(self):
        p = N.nbh().nbh_obj_to_json(self._h)
        try:
            return json.loads(C.string_at(p).decode())
        finally:
            N.nbh().nbh_str_free(p)

    def digest(self):
        return N.nbh().nbh_obj_digest(self._h)

    def get(self, dotted_path):
        """Copy of the sub-object at e.g. "spec.template.spec" (None if absent)."""
        out = C.c_void_p()
        if N.nbh().nbh_obj_get(self._h, dotted_path.encode(), C.byref(out)) != 0:
            return None
        return Obj(out.value)

    def clone(self):
        return Obj(N.nbh().nbh_obj_clone(self._h))

    def erase(self, dotted_path):
        if N.nbh().nbh_obj_erase(self._h, dotted_path.encode()) != 0:
            raise HostError("no such path: " + dotted_path)
        return self

    def set_str(self, dotted_path, value):
        if N.nbh().nbh_obj_set_str(self._h, dotted_path.encode(), value.encode()) != 0:
            raise HostError("no such path: " + dotted_path)
        return self

    def __del__(self):
        try:
            if self._h:
                N.nbh().nbh_obj_free(self._h)
                self._h = None
        except Exception:
            pass


def Notebook(obj):
    return Obj.from_json("Notebook", obj)


def StatefulSet(obj):
    return Obj.from_json("StatefulSet", obj)


def Deployment(obj):
    return Obj.from_json("Deployment", obj)


def Service(obj):
    return Obj.from_json("Service", obj)


def Unstructured(obj):
    return Obj.from_json("Unstructured", obj)


def _env(env):
    if env is None:
        return None
    e = N.nbh_env()
    e.add_fsgroup = 1 if env.get("ADD_FSGROUP", "true") == "true" else 0
    e.cluster_domain = env["CLUSTER_DOMAIN"].encode() if "CLUSTER_DOMAIN" in env else None
    e.istio_host = env.get("ISTIO_HOST", "").encode()
    e.istio_gateway = env.get("ISTIO_GATEWAY", "").encode()
    return C.byref(e)


def generate_statefulset(notebook, env=None):
    """generateStatefulSet (NBCTL:433-523). `env` = dict of the process env vars it reads."""
    out, err = C.c_void_p(), _err()
    if N.nbh().nbh_generate_statefulset(notebook._h, _env(env), C.byref(out), err, len(err)) != 0:
        raise HostError("panic: " + err.value.decode())
    return Obj(out.value)


def generate_service(notebook):
    """generateService (NBCTL:525-552)."""
    out, err = C.c_void_p(), _err()
    if N.nbh().nbh_generate_service(notebook._h, C.byref(out), err, len(err)) != 0:
        raise HostError("panic: " + err.value.decode())
    return Obj(out.value)


def generate_virtualservice(notebook, env=None):
    """generateVirtualService (NBCTL:558-658)."""
    out, err = C.c_void_p(), _err()
    if N.nbh().nbh_generate_virtualservice(notebook._h, _env(env), C.byref(out), err, len(err)) != 0:
        raise HostError(err.value.decode())
    return Obj(out.value)


ODH_LOCK, ODH_CA_BUNDLE, ODH_RUNTIME_IMAGES, ODH_ELYRA_SECRET, ODH_RBAC_PROXY, ODH_FEAST = 1, 2, 4, 8, 16, 32


def odh_mutate(notebook, flags=2 | 4 | 8 | 16 | 32, proxy=None):
    """ODH mutating-webhook injections (ODHWH:105-114,177-326,431-444,666-781), in place.
    proxy = (http_proxy, https_proxy, no_proxy): also InjectProxyConfigEnvVars (ODHWH:566-616)."""
    if proxy is not None:
        rc = N.nbh().nbh_odh_mutate_proxy(notebook._h, flags, *[p.encode() for p in proxy])
    else:
        rc = N.nbh().nbh_odh_mutate(notebook._h, flags)
    if rc == 1:        # InjectKubeRbacProxy refused the auth-sidecar resource annotations: no proxy was injected
        raise HostError("invalid kube-rbac-proxy resource configuration")
    if rc != 0:
        raise HostError("odh_mutate")
    return notebook


def odh_build(notebook, what, central_namespace="redhat-ods-applications", gateway_name="", gateway_namespace="",
              rolebinding_name="", role_kind="", role_name=""):
    """The ODH controller's desired object for a Notebook: what = "HTTPRoute" | "NetworkPolicy" |
    "KubeRbacProxyNetworkPolicy" | "ReferenceGrant" | "RoleBinding" (the `from` side of its comparators)."""
    out = C.c_void_p()
    rc = N.nbh().nbh_odh_build(notebook._h, what.encode(), central_namespace.encode(), gateway_name.encode(),
                               gateway_namespace.encode(), rolebinding_name.encode(), role_kind.encode(), role_name.encode(),
                               C.byref(out))
    if rc != 0:
        raise HostError("odh_build: unknown kind " + what)
    return Obj(out.value)


def _arr(ptr, count, dtype):
    if count == 0:
        return np.empty(0, dtype=dtype)
    ct = {np.uint8: C.c_uint8, np.uint32: C.c_uint32, np.uint64: C.c_uint64, np.int32: C.c_int32,
          np.int64: C.c_int64}[dtype]
    return np.ctypeslib.as_array(C.cast(ptr, C.POINTER(ct)), shape=(count,))


def _kv_arrays(m, n, prefix, out):
    off = _arr(m.off, n + 1, np.uint32)
    out[prefix + ".off"] = off
    cnt = int(off[n]) if n else 0
    out[prefix + ".ent"] = _arr(m.ent, cnt * 2, np.uint64)


class _Batch:
    """A packed batch owned by the native library.  `.view` is the C struct of
    include/nbgraft.h, `.arrays()` numpy views of every array it points to."""
    _new = _free = _add = _view = None
    _struct = None

    def __init__(self, handle=None):
        lib = N.nbh()
        self._h = C.c_void_p(handle if handle is not None else getattr(lib, self._new)())
        self._keep = []

    def add(self, from_, to):
        if getattr(N.nbh(), self._add)(self._h, from_._h, to._h) != 0:
            raise HostError("add failed")
        return self

    @property
    def view(self):
        v = self._struct()
        self.n = int(getattr(N.nbh(), self._view)(self._h, C.byref(v)))
        return v

    def __len__(self):
        self.view
        return self.n

    def algorithmic_bytes(self):
        """sum of sizeof(input arrays) + sizeof(output arrays): SURVEY.md 8(d)."""
        n = len(self)
        return sum(a.nbytes for a in self.arrays().values()) + self.output_bytes(n)

    def __del__(self):
        try:
            if self._h:
                getattr(N.nbh(), self._free)(self._h)
                self._h = None
        except Exception:
            pass


class StsBatch(_Batch):
    _new, _free, _add, _view, _struct = "nbh_sts_new", "nbh_sts_free", "nbh_sts_add", "nbh_sts_view", N.nbg_sts_batch

    @property
    def lean(self):
        return bool(N.nbh().nbh_sts_is_lean(self._h))

    def arrays(self):
        if self.lean:
            raise HostError("a lean StsBatch has no dense tiles: use .sparse()")
        v, out = self.view, {}
        n = self.n
        out["replicas_from"] = _arr(v.replicas_from, n, np.int32)
        out["replicas_to"] = _arr(v.replicas_to, n, np.int32)
        out["tmpl_from"] = _arr(v.tmpl_from, n, np.uint64)
        out["tmpl_to"] = _arr(v.tmpl_to, n, np.uint64)
        out["pod_from"] = _arr(v.pod_from, n * N.NBG_POD_SLOTS, np.uint64)
        out["pod_to"] = _arr(v.pod_to, n * N.NBG_POD_SLOTS, np.uint64)
        for name in ("labels_from", "labels_to", "ann_from", "ann_to"):
            _kv_arrays(getattr(v, name), n, name, out)
        return out

    @staticmethod
    def output_bytes(n):
        return n + (n + 63) // 64 * 8


    def sparse(self):
        """The same pairs with sparsely stored PodSpec tiles (include/nbgraft.h: nbg_sts_sparse_batch)."""
        return StsSparseBatch(self)

    def add_gen(self, notebook, to, env=None):
        """Append a pair whose desired side will be generated ON THE DEVICE from the Notebook (nbg_diff_sts_gen):
        the batch then holds the Notebook's own PodSpec tile and labels instead of a generated StatefulSet's."""
        if N.nbh().nbh_sts_add_gen(self._h, notebook._h, to._h, _env(env)) != 0:
            raise HostError("add_gen: generateStatefulSet would fail or leave its common path for this Notebook")
        return self

    def add_notebook(self, notebook, to, env=None):
        """Append the pair add(generate_statefulset(notebook), to) would, without building the StatefulSet: the
        desired tile is made on the host from the Notebook's own (nbh_sts_add_notebook).  For the sparse step."""
        if N.nbh().nbh_sts_add_notebook(self._h, notebook._h, to._h, _env(env)) != 0:
            raise HostError("add_notebook: generateStatefulSet would fail or leave its common path for this Notebook")
        return self

    @property
    def gen_inputs(self):
        g = N.nbg_sts_gen_inputs()
        n = int(N.nbh().nbh_sts_view_gen(self._h, C.byref(g)))
        if n != len(self):
            raise HostError("not a batch filled with add_gen")
        return g


class StsGenView:
    """Non-owning view of a StsBatch filled with add_gen: the arrays nbg_diff_sts_gen reads (the Notebook's tile and
    labels, the live side, one hash and one flag byte per Notebook)."""
    kind = "sts_gen"

    def __init__(self, parent):
        self.parent = parent
        parent.gen_inputs                                   # raises unless it is a gen batch

    def __len__(self):
        return len(self.parent)

    def arrays(self):
        a = self.parent.arrays()
        out = {k: v for k, v in a.items() if k != "replicas_from" and not k.startswith("ann_from")}
        g, n = self.parent.gen_inputs, len(self.parent)
        out["gen.prefix_hash"] = _arr(g.prefix_hash, n, np.uint64)
        out["gen.flags"] = _arr(g.flags, n, np.uint8)
        return out

    output_bytes = staticmethod(StsBatch.output_bytes)

    def algorithmic_bytes(self):
        return sum(a.nbytes for a in self.arrays().values()) + self.output_bytes(len(self))


def sts_gen_consts(env=None):
    c = N.nbg_sts_gen_consts()
    if N.nbh().nbh_sts_gen_consts(_env(env), C.byref(c)) != 0:
        raise HostError("sts_gen_consts")
    return c


class StsSparseBatch:
    """Non-owning sparse view of a StsBatch (keeps the parent alive)."""

    def __init__(self, parent):
        self.parent = parent

    @property
    def view(self):
        v = N.nbg_sts_sparse_batch()
        self.n = int(N.nbh().nbh_sts_view_sparse(self.parent._h, C.byref(v)))
        return v

    def __len__(self):
        self.view
        return self.n

    def arrays(self):
        v, out = self.view, {}
        n = self.n
        out["replicas_from"] = _arr(v.replicas_from, n, np.int32)
        out["replicas_to"] = _arr(v.replicas_to, n, np.int32)
        out["tmpl_from"] = _arr(v.tmpl_from, n, np.uint64)
        out["tmpl_to"] = _arr(v.tmpl_to, n, np.uint64)
        for side in ("pod_from", "pod_to"):
            ps = getattr(v, side)
            off = _arr(ps.off, n + 1, np.uint32)
            out[side + ".mask"] = _arr(ps.mask, n * 3, np.uint64)
            out[side + ".voff"] = off
            out[side + ".val"] = _arr(ps.val, int(off[n]) if n else 0, np.uint64)
        for name in ("labels_from", "labels_to", "ann_from", "ann_to"):
            _kv_arrays(getattr(v, name), n, name, out)
        return out

    @staticmethod
    def output_bytes(n):
        return n + (n + 63) // 64 * 8

    def algorithmic_bytes(self):
        n = len(self)
        return sum(a.nbytes for a in self.arrays().values()) + self.output_bytes(n)


class SvcBatch(_Batch):
    _new, _free, _add, _view, _struct = "nbh_svc_new", "nbh_svc_free", "nbh_svc_add", "nbh_svc_view", N.nbg_svc_batch

    def gen(self):
        """Desired side generated on the device (valid when every `from` came from generate_service)."""
        g = _GenView(self, "nbh_svc_view_gen", N.nbg_svc_gen_batch,
                     [("name_hash", np.uint64, 1), ("target_port", np.int32, 1), ("flags", np.uint8, 1), ("ports_to", np.uint64, N.NBG_SVC_SLOTS)],
                     ["labels_to", "ann_to", "sel_to"])
        g.kind = "svc_gen"
        return g

    def arrays(self):
        v, out = self.view, {}
        n = self.n
        out["flags"] = _arr(v.flags, n, np.uint8)
        out["ports_from"] = _arr(v.ports_from, n * N.NBG_SVC_SLOTS, np.uint64)
        out["ports_to"] = _arr(v.ports_to, n * N.NBG_SVC_SLOTS, np.uint64)
        for name in ("labels_from", "labels_to", "ann_from", "ann_to", "sel_from", "sel_to"):
            _kv_arrays(getattr(v, name), n, name, out)
        return out

    @staticmethod
    def output_bytes(n):
        return n + (n + 63) // 64 * 8


class _GenView:
    """Non-owning view of a Svc/Vs batch with the desired side reduced to the few Notebook words
    generateService / generateVirtualService depend on (nbg_*_gen_batch)."""

    def __init__(self, parent, fn, struct, names_fixed, names_kv):
        self.parent, self._fn, self._struct, self._fixed, self._kv = parent, fn, struct, names_fixed, names_kv

    @property
    def view(self):
        v = self._struct()
        self.n = int(getattr(N.nbh(), self._fn)(self.parent._h, C.byref(v)))
        return v

    def __len__(self):
        self.view
        return self.n

    def arrays(self):
        v, out = self.view, {}
        for name, dt, per in self._fixed:
            out[name] = _arr(getattr(v, name), self.n * per, dt)
        for name in self._kv:
            _kv_arrays(getattr(v, name), self.n, name, out)
        return out

    @staticmethod
    def output_bytes(n):
        return n + (n + 63) // 64 * 8

    def algorithmic_bytes(self):
        n = len(self)
        return sum(a.nbytes for a in self.arrays().values()) + self.output_bytes(n)


def vs_gen_consts(env=None):
    c = N.nbg_vs_gen_consts()
    if N.nbh().nbh_vs_gen_consts(_env(env), C.byref(c)) != 0:
        raise HostError("vs_gen_consts")
    return c


class VsBatch(_Batch):
    _new, _free, _add, _view, _struct = "nbh_vs_new", "nbh_vs_free", "nbh_vs_add", "nbh_vs_view", N.nbg_vs_batch

    def gen(self):
        """Desired side generated on the device (valid when every `from` came from generate_virtualservice)."""
        g = _GenView(self, "nbh_vs_view_gen", N.nbg_vs_gen_batch,
                     [("prefix_hash", np.uint64, 1), ("rewrite_hash", np.uint64, 1), ("host_hash", np.uint64, 1),
                      ("flags", np.uint8, 1), ("spec_to", np.uint64, N.NBG_VS_SLOTS)], ["hdr_from", "hdr_to"])
        g.kind = "vs_gen"
        return g

    def arrays(self):
        v, out = self.view, {}
        n = self.n
        out["flags"] = _arr(v.flags, n, np.uint8)
        out["spec_from"] = _arr(v.spec_from, n * N.NBG_VS_SLOTS, np.uint64)
        out["spec_to"] = _arr(v.spec_to, n * N.NBG_VS_SLOTS, np.uint64)
        for name in ("hdr_from", "hdr_to"):
            _kv_arrays(getattr(v, name), n, name, out)
        return out

    @staticmethod
    def output_bytes(n):
        return n + (n + 63) // 64 * 8


class GenBatch(_Batch):
    """Pairs for the ODH comparator family (nbg_generic_batch)."""
    _new, _free, _view, _struct = "nbh_gen_new", "nbh_gen_free", "nbh_gen_view", N.nbg_generic_batch

    def __init__(self, mode=N.GEN_DEEPEQUAL, labels_path="metadata.labels", spec_path="spec"):
        super().__init__()
        self.mode, self.labels_path, self.spec_path = mode, labels_path, spec_path

    def add(self, desired, found):
        if N.nbh().nbh_gen_add(self._h, desired._h, found._h, self.labels_path.encode(), self.spec_path.encode(), self.mode) != 0:
            raise HostError("gen add failed")
        return self

    def arrays(self):
        v, out = self.view, {}
        for name in ("labels_a", "labels_b", "spec_a", "spec_b"):
            _kv_arrays(getattr(v, name), self.n, name, out)
        return out

    @staticmethod
    def output_bytes(n):
        return n + (n + 63) // 64 * 8


class RestartBatch(_Batch):
    """Webhook restart-block inputs (nbg_restart_batch)."""
    _new, _free, _view, _struct = "nbh_rb_new", "nbh_rb_free", "nbh_rb_view", N.nbg_restart_batch

    def add(self, old_nb, updated_nb, mutated_nb, create=False):
        if N.nbh().nbh_rb_add(self._h, old_nb._h, updated_nb._h, mutated_nb._h, 1 if create else 0) != 0:
            raise HostError("rb add failed")
        return self

    def arrays(self):
        v, out = self.view, {}
        n = self.n
        out["flags"] = _arr(v.flags, n, np.uint8)
        for side in ("old_spec", "updated_spec", "mutated_spec"):
            ps = getattr(v, side)
            off = _arr(ps.off, n + 1, np.uint32)
            out[side + ".mask"] = _arr(ps.mask, n * 3, np.uint64)
            out[side + ".voff"] = off
            out[side + ".val"] = _arr(ps.val, int(off[n]) if n else 0, np.uint64)
        return out

    @staticmethod
    def output_bytes(n):
        return n + (n + 63) // 64 * 8


class CullBatch(_Batch):
    _new, _free, _view, _struct = "nbh_cull_new", "nbh_cull_free", "nbh_cull_view", N.nbg_cull_batch

    def add(self, notebook):
        """notebook: {"annotations": {...}|None, "kernels": [...], "terminals": [...], "pod_missing": bool}"""
        text = notebook if isinstance(notebook, str) else json.dumps(notebook)
        err = _err()
        if N.nbh().nbh_cull_add_json(self._h, text.encode(), err, len(err)) != 0:
            raise HostError(err.value.decode())
        return self

    def records(self):
        k, t = C.c_uint64(), C.c_uint64()
        N.nbh().nbh_cull_records(self._h, C.byref(k), C.byref(t))
        return int(k.value), int(t.value)

    def arrays(self):
        v, out = self.view, {}
        n = self.n
        nk, nt = self.records()
        out["k_off"] = _arr(v.k_off, n + 1, np.uint32)
        out["k_sec"] = _arr(v.k_sec, nk, np.int64)
        out["k_nsf"] = _arr(v.k_nsf, nk, np.uint32)
        out["t_off"] = _arr(v.t_off, n + 1, np.uint32)
        out["t_sec"] = _arr(v.t_sec, nt, np.int64)
        out["t_nsf"] = _arr(v.t_nsf, nt, np.uint32)
        out["ann_sec"] = _arr(v.ann_sec, n, np.int64)
        out["ann_nsf"] = _arr(v.ann_nsf, n, np.uint32)
        out["chk_sec"] = _arr(v.chk_sec, n, np.int64)
        out["chk_nsf"] = _arr(v.chk_nsf, n, np.uint32)
        out["nb_flags"] = _arr(v.nb_flags, n, np.uint8)
        return out

    @staticmethod
    def output_bytes(n):
        return n + 4 * n + 2 * ((n + 63) // 64 * 8)

    def new_last_activity(self, i, action, argmax, now_sec, now_offset=0):
        """The string CULL:386/399/425 would have written for notebook i ('' = unchanged)."""
        buf = C.create_string_buffer(64)
        r = N.nbh().nbh_cull_new_last_activity(self._h, i, int(action), int(argmax), now_sec, now_offset, buf, len(buf))
        if r < 0:
            raise HostError("bad argmax for notebook %d" % i)
        return buf.value.decode()


def parse_rfc3339(s):
    """time.Parse(time.RFC3339, s) -> (sec, nsec, offset) or None."""
    sec, nsec, off = C.c_int64(), C.c_int32(), C.c_int32()
    if N.nbh().nbh_parse_rfc3339(s.encode(), C.byref(sec), C.byref(nsec), C.byref(off)):
        return sec.value, nsec.value, off.value
    return None


def format_rfc3339(sec, offset=0):
    buf = C.create_string_buffer(64)
    N.nbh().nbh_format_rfc3339(sec, offset, buf, len(buf))
    return buf.value.decode()


def join_key_owner(obj):
    """Join key of an owner (Notebook): namespace + metadata.uid."""
    return N.nbh().nbh_join_key_owner(obj._h)


def join_key_controlled(obj):
    """Join key of a dependent (StatefulSet): namespace + UID of its controller owner reference; 0 = none."""
    return N.nbh().nbh_join_key_controlled(obj._h)


FLEET_KUBEFLOW, FLEET_ODH = 2, 4
SEED_BASE = 0x6E62677261667400  # seed = base + config number


class Fleet:
    """Seeded synthetic Notebook fleet (desired/live StatefulSet, Service, VirtualService pairs)."""
    WHICH = {"notebook": 0, "sts_from": 1, "sts_to": 2, "svc_from": 3, "svc_to": 4, "vs_from": 5, "vs_to": 6}

    def __init__(self, kind, seed, n, n_namespaces):
        self.kind, self.seed, self.n, self.n_namespaces = kind, seed, n, n_namespaces
        self._h = C.c_void_p(N.nbh().nbh_fleet_new(kind, seed, n, n_namespaces))
        if not self._h:
            raise HostError("bad fleet kind")

    def get(self, i, which):
        out = C.c_void_p()
        if N.nbh().nbh_fleet_get(self._h, i, self.WHICH[which], C.byref(out)) != 0:
            raise HostError("fleet_get")
        return Obj(out.value)

    def drift_class(self, i):
        return N.nbh().nbh_fleet_drift_class(self._h, i)

    def ns_hash(self, i):
        return N.nbh().nbh_fleet_ns_hash(self._h, i)

    def pack(self, i0=0, m=None, idx=None, sts=True, svc=True, vs=True, lean=False):
        """Pack objects [i0, i0+m) (or the explicit index list) on all host threads.
        lean=True keeps only the sparse PodSpec tiles of the StatefulSet pairs (what the step reads):
        the StsBatch then has no dense arrays (use .sparse())."""
        if idx is not None:
            idx = np.ascontiguousarray(idx, dtype=np.uint64)
            m, ip = len(idx), idx.ctypes.data_as(N.u64p)
        else:
            m, ip = (self.n - i0 if m is None else m), None
        a, b, c = C.c_void_p(), C.c_void_p(), C.c_void_p()
        fn = N.nbh().nbh_fleet_pack_lean if lean else N.nbh().nbh_fleet_pack
        if fn(self._h, ip, i0, m, C.byref(a) if sts else None, C.byref(b) if svc else None,
              C.byref(c) if vs else None) != 0:
            raise HostError("fleet_pack")
        return (StsBatch(a.value) if sts else None, SvcBatch(b.value) if svc else None,
                VsBatch(c.value) if vs else None)

    @staticmethod
    def pack_timing():
        """(synthesis, packing) CPU seconds, summed over threads, of the last pack() call."""
        a, b = C.c_double(), C.c_double()
        N.nbh().nbh_fleet_pack_timing(C.byref(a), C.byref(b))
        return a.value, b.value

    def shard_indices(self, n_shards, shard=None):
        """Bucket object indices by nbg_shard_of(hash(namespace)) — SURVEY.md 8(e).
        Returns the list of all shards' index arrays, or one shard's array."""
        def one(s):
            cnt = N.nbh().nbh_fleet_shard_indices(self._h, n_shards, s, None, 0)
            out = np.empty(cnt, dtype=np.uint64)
            N.nbh().nbh_fleet_shard_indices(self._h, n_shards, s, out.ctypes.data_as(N.u64p), cnt)
            return out
        return one(shard) if shard is not None else [one(s) for s in range(n_shards)]

    def ns_table(self):
        """(ns_hash[K], ns_size[K]) of the fleet's K namespaces: the input of nbg_shard_of / nbg_balance_shards."""
        K = self.n_namespaces
        h, sz = np.empty(K, np.uint64), np.empty(K, np.uint64)
        got = N.nbh().nbh_fleet_ns_table(self._h, h.ctypes.data_as(N.u64p), sz.ctypes.data_as(N.u64p))
        assert got == K
        return h, sz

    def shard_table(self, n_shards, balanced=False):
        """namespace -> shard table (uint32[K]) and the shards' sizes: nbg_shard_of per namespace, or the
        size-aware assignment of nbg_balance_shards (still one namespace -> one shard, NBCTL:159)."""
        h, sz = self.ns_table()
        table = np.empty(len(h), np.uint32)
        load = np.zeros(n_shards, np.uint64)
        if balanced:
            rc = N.nbg().nbg_balance_shards(h.ctypes.data, sz.ctypes.data, len(h), n_shards, table.ctypes.data, load.ctypes.data)
            if rc != 0:
                raise HostError("nbg_balance_shards: %d" % rc)
        else:
            lib = N.nbg()
            for k in range(len(h)):
                table[k] = lib.nbg_shard_of(int(h[k]), n_shards)
            load = np.bincount(table, weights=sz.astype(np.float64), minlength=n_shards).astype(np.uint64)
        return table, load

    def shard_indices_by_table(self, table, shard):
        table = np.ascontiguousarray(table, dtype=np.uint32)
        assert len(table) == self.n_namespaces
        cnt = N.nbh().nbh_fleet_shard_indices_by_table(self._h, table.ctypes.data_as(N.u32p), shard, None, 0)
        out = np.empty(cnt, dtype=np.uint64)
        N.nbh().nbh_fleet_shard_indices_by_table(self._h, table.ctypes.data_as(N.u32p), shard, out.ctypes.data_as(N.u64p), cnt)
        return out

    def join_lists(self, n):
        """Synthetic StatefulSet list for notebooks [0,n): (nb_keys[n], sts_keys[m], sts_owner[m])."""
        m = N.nbh().nbh_fleet_join_lists(self._h, n, None, None, None, 0)
        nb, st, ow = np.empty(n, np.uint64), np.empty(m, np.uint64), np.empty(m, np.int64)
        N.nbh().nbh_fleet_join_lists(self._h, n, nb.ctypes.data_as(N.u64p), st.ctypes.data_as(N.u64p),
                                     ow.ctypes.data_as(N.i64p), m)
        return nb, st, ow

    def objects(self, i0=0, m=None, idx=None):
        """All seven objects of m notebooks at once -> list of dicts name -> Obj."""
        if idx is not None:
            idx = np.ascontiguousarray(idx, dtype=np.uint64)
            m, ip = len(idx), idx.ctypes.data_as(N.u64p)
        else:
            m, ip = (self.n - i0 if m is None else m), None
        out = (C.c_void_p * (m * 7))()
        if N.nbh().nbh_fleet_objects(self._h, ip, i0, m, out) != 0:
            raise HostError("fleet_objects")
        names = sorted(self.WHICH, key=self.WHICH.get)
        return [{nm: Obj(out[k * 7 + w]) for w, nm in enumerate(names)} for k in range(m)]

    def __del__(self):
        try:
            if self._h:
                N.nbh().nbh_fleet_free(self._h)
                self._h = None
        except Exception:
            pass


def pack_objects(rows, lean=False):
    """Pack notebooks whose objects are already in memory: rows = list of dicts as Fleet.objects() returns
    (sts_from, sts_to, svc_from, svc_to, vs_from, vs_to; the Notebook is not read).  All host threads."""
    arr, m = rows if isinstance(rows, tuple) else object_rows(rows)
    a, b, c = C.c_void_p(), C.c_void_p(), C.c_void_p()
    if N.nbh().nbh_pack_objects(arr, m, 1 if lean else 0, C.byref(a), C.byref(b), C.byref(c)) != 0:
        raise HostError("pack_objects")
    return StsBatch(a.value), SvcBatch(b.value), VsBatch(c.value)


def object_rows(rows):
    """The handle table nbh_pack_objects takes, built once from Fleet.objects() rows (so that a timed
    pack_objects call contains no Python loop)."""
    names = sorted(Fleet.WHICH, key=Fleet.WHICH.get)
    return (C.c_void_p * (len(rows) * 7))(*[r[nm]._h.value for r in rows for nm in names]), len(rows)


NOW_SEC, NOW_NSEC = 1767225600, 123456789  # 2026-01-01T00:00:00.123456789Z


class CullFleet:
    def __init__(self, seed, n, now_sec=NOW_SEC, now_nsec=NOW_NSEC, mean_kernels=4.0):
        self.seed, self.n, self.now_sec, self.now_nsec = seed, n, now_sec, now_nsec
        self._h = C.c_void_p(N.nbh().nbh_cullfleet_new(seed, n, now_sec, now_nsec, mean_kernels))

    def json(self, i):
        p = N.nbh().nbh_cullfleet_json(self._h, i)
        try:
            return C.string_at(p).decode()
        finally:
            N.nbh().nbh_str_free(p)

    def pack(self, i0=0, m=None, idx=None):
        if idx is not None:
            idx = np.ascontiguousarray(idx, dtype=np.uint64)
            m, ip = len(idx), idx.ctypes.data_as(N.u64p)
        else:
            m, ip = (self.n - i0 if m is None else m), None
        out = C.c_void_p()
        if N.nbh().nbh_cullfleet_pack(self._h, ip, i0, m, C.byref(out)) != 0:
            raise HostError("cullfleet_pack")
        return CullBatch(out.value)

    def __del__(self):
        try:
            if self._h:
                N.nbh().nbh_cullfleet_free(self._h)
                self._h = None
        except Exception:
            pass
