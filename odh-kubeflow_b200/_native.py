"""ctypes bindings of the two in-tree shared libraries.

libnbgraft.so       kernels + C ABI (include/nbgraft.h)
libnbgraft_host.so  host packer / builders / fleets (include/nbgraft_host.h)

There is no Python or CPU fallback for the kernels: if libnbgraft.so is missing
the import fails loudly, and without a CUDA device ``nbg_init`` returns
NBG_E_NODEVICE, which ``Context`` turns into an exception.
"""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))


class NativeLibraryMissing(ImportError):
    pass


def _load(name):
    # NBGRAFT_LIB points at an alternative build of libnbgraft.so (kernel variants under tools/)
    path = os.environ.get("NBGRAFT_LIB") if name == "libnbgraft.so" and os.environ.get("NBGRAFT_LIB") else os.path.join(_HERE, name)
    if not os.path.exists(path):
        raise NativeLibraryMissing(
            f"{path} is not built. Run `python -c 'import __graft_entry__ as g; g.build()'` "
            f"(or `make -C odh-kubeflow_b200/csrc`). There is no Python fallback for the CUDA path.")
    return C.CDLL(path)


u8p, u32p, u64p = C.POINTER(C.c_uint8), C.POINTER(C.c_uint32), C.POINTER(C.c_uint64)
i32p, i64p = C.POINTER(C.c_int32), C.POINTER(C.c_int64)


class nbg_kv(C.Structure):
    _fields_ = [("key", C.c_uint64), ("val", C.c_uint64)]


class nbg_kvmap(C.Structure):
    _fields_ = [("off", u32p), ("ent", C.POINTER(nbg_kv)), ("end", u32p)]


class nbg_sts_batch(C.Structure):
    _fields_ = [("replicas_from", i32p), ("replicas_to", i32p), ("tmpl_from", u64p), ("tmpl_to", u64p),
                ("pod_from", u64p), ("pod_to", u64p),
                ("labels_from", nbg_kvmap), ("labels_to", nbg_kvmap),
                ("ann_from", nbg_kvmap), ("ann_to", nbg_kvmap)]


class nbg_podsparse(C.Structure):
    _fields_ = [("mask", u64p), ("off", u32p), ("val", u64p), ("end", u32p)]


class nbg_sts_sparse_batch(C.Structure):
    _fields_ = [("replicas_from", i32p), ("replicas_to", i32p), ("tmpl_from", u64p), ("tmpl_to", u64p),
                ("pod_from", nbg_podsparse), ("pod_to", nbg_podsparse),
                ("labels_from", nbg_kvmap), ("labels_to", nbg_kvmap),
                ("ann_from", nbg_kvmap), ("ann_to", nbg_kvmap)]


class nbg_svc_batch(C.Structure):
    _fields_ = [("flags", u8p), ("ports_from", u64p), ("ports_to", u64p),
                ("labels_from", nbg_kvmap), ("labels_to", nbg_kvmap),
                ("ann_from", nbg_kvmap), ("ann_to", nbg_kvmap),
                ("sel_from", nbg_kvmap), ("sel_to", nbg_kvmap)]


class nbg_vs_batch(C.Structure):
    _fields_ = [("flags", u8p), ("spec_from", u64p), ("spec_to", u64p),
                ("hdr_from", nbg_kvmap), ("hdr_to", nbg_kvmap)]


class nbg_svc_gen_batch(C.Structure):
    _fields_ = [("name_hash", u64p), ("target_port", i32p), ("flags", u8p), ("ports_to", u64p),
                ("labels_to", nbg_kvmap), ("ann_to", nbg_kvmap), ("sel_to", nbg_kvmap)]


class nbg_vs_gen_batch(C.Structure):
    _fields_ = [("prefix_hash", u64p), ("rewrite_hash", u64p), ("host_hash", u64p), ("hdr_from", nbg_kvmap),
                ("flags", u8p), ("spec_to", u64p), ("hdr_to", nbg_kvmap)]


class nbg_vs_gen_consts(C.Structure):
    _fields_ = [("istio_host_hash", C.c_uint64), ("gateway_hash", C.c_uint64), ("residual", C.c_uint64)]


class nbg_step_outputs(C.Structure):
    _fields_ = [(k, C.c_void_p) for k in ("sts_reason", "sts_bitmap", "svc_reason", "svc_bitmap", "vs_reason", "vs_bitmap",
                                           "action", "argmax", "cull_bitmap", "period_bitmap")]


class nbg_store_stats(C.Structure):
    _fields_ = [(k, C.c_uint64) for k in ("capacity", "fixed_bytes", "heap_bytes", "live_bytes", "garbage_bytes", "compactions",
                                          "puts", "rows_put")]


class nbg_restart_batch(C.Structure):
    _fields_ = [("flags", u8p), ("old_spec", nbg_podsparse), ("updated_spec", nbg_podsparse), ("mutated_spec", nbg_podsparse)]


RB_CREATE, RB_STOPPED, RB_RESTARTING = 1, 2, 4
RB_BLOCKED, RB_EXTERNAL_CHANGE, RB_NO_CHANGE = 1, 2, 4


class nbg_generic_batch(C.Structure):
    _fields_ = [("labels_a", nbg_kvmap), ("labels_b", nbg_kvmap), ("spec_a", nbg_kvmap), ("spec_b", nbg_kvmap)]


GEN_DEEPEQUAL, GEN_CONFIGMAP = 0, 1


class nbg_strings(C.Structure):
    _fields_ = [("off", u32p), ("bytes", u8p)]


class nbg_cull_text_batch(C.Structure):
    _fields_ = [("k_off", u32p), ("k_last_activity", nbg_strings), ("k_state", nbg_strings), ("t_off", u32p),
                ("t_last_activity", nbg_strings), ("ann_text", nbg_strings), ("chk_text", nbg_strings), ("nb_flags", u8p)]


NB_LA_PRESENT, NB_POD_MISSING, NBT_STOP_SET, NBT_CHK_PRESENT = 1, 2, 0x10, 0x20


class nbg_cull_batch(C.Structure):
    _fields_ = [("k_off", u32p), ("k_sec", i64p), ("k_nsf", u32p),
                ("t_off", u32p), ("t_sec", i64p), ("t_nsf", u32p),
                ("ann_sec", i64p), ("ann_nsf", u32p), ("chk_sec", i64p), ("chk_nsf", u32p),
                ("nb_flags", u8p), ("k_end", u32p), ("t_end", u32p)]


class nbg_cull_params(C.Structure):
    _fields_ = [("now_sec", C.c_int64), ("now_nsec", C.c_int32), ("cull_idle_min", C.c_int32),
                ("check_period_min", C.c_int32), ("mode", C.c_int32)]


class nbh_env(C.Structure):
    _fields_ = [("add_fsgroup", C.c_int), ("cluster_domain", C.c_char_p), ("istio_host", C.c_char_p),
                ("istio_gateway", C.c_char_p)]


# constants of include/nbgraft.h
NBG_OK, NBG_E_NODEVICE, NBG_E_INVAL, NBG_E_NOMEM, NBG_E_CUDA = 0, -1, -2, -3, -4
NBG_E_NILREPLICAS, NBG_E_RANGE, NBG_E_TOOBIG = -5, -6, -7
NBG_POD_SLOTS, NBG_SVC_SLOTS, NBG_VS_SLOTS = 192, 8, 8
NBG_REPLICAS_NIL = -(2 ** 31)
NBG_JOIN_STATIC_MAX_STS = 131072
NBG_ARGMAX_NONE = 0xFFFFFFFF
STS_LABELS, STS_ANNOTATIONS, STS_REPLICAS, STS_PODSPEC, STS_TMPL_LABELS, STS_NILREPLICAS = 1, 2, 4, 8, 16, 0x80
SVC_LABELS, SVC_ANNOTATIONS, SVC_SELECTOR, SVC_PORTS = 1, 2, 4, 8
VS_SPEC_DIFFERS, VS_TO_MISSING, VS_FROM_MISSING = 1, 2, 4
ACT_SRC_MASK, ACT_PERIOD, ACT_CULL, ACT_STOPPED, ACT_INIT, ACT_NOPOD = 3, 4, 8, 16, 32, 64
CULL_PREDICATES, CULL_RECONCILE = 0, 1
GEN_STOPPED, GEN_PREPATCHED = 1, 2
STS_UPDATE_MASK = 0x0F

# ---- one place that knows which array goes into which field of which batch struct -----------------
# Array names are those of host.*Batch.arrays(): "<map>.off/.ent[/.end]" for string maps,
# "<side>.mask/.voff/.val[/.vend]" for sparse PodSpec tiles, plain field names otherwise.  `ptr(name, ctype)`
# returns a ctypes pointer to that array wherever it lives (numpy, pinned memory, a device tensor), `has(name)`
# says whether the optional span arrays exist.
_KV_FIELDS = {"StsBatch": ("labels_from", "labels_to", "ann_from", "ann_to"),
              "StsSparseBatch": ("labels_from", "labels_to", "ann_from", "ann_to"),
              "SvcBatch": ("labels_from", "labels_to", "ann_from", "ann_to", "sel_from", "sel_to"),
              "sts_gen": ("labels_from", "labels_to", "ann_to"),
              "VsBatch": ("hdr_from", "hdr_to"), "svc_gen": ("labels_to", "ann_to", "sel_to"), "vs_gen": ("hdr_from", "hdr_to"),
              "GenBatch": ("labels_a", "labels_b", "spec_a", "spec_b"), "RestartBatch": (), "CullBatch": ()}
_SPARSE_FIELDS = {"StsSparseBatch": ("pod_from", "pod_to"), "RestartBatch": ("old_spec", "updated_spec", "mutated_spec")}
_PLAIN_FIELDS = {
    "StsBatch": (("replicas_from", C.c_int32), ("replicas_to", C.c_int32), ("tmpl_from", C.c_uint64), ("tmpl_to", C.c_uint64),
                 ("pod_from", C.c_uint64), ("pod_to", C.c_uint64)),
    "sts_gen": (("replicas_to", C.c_int32), ("tmpl_from", C.c_uint64), ("tmpl_to", C.c_uint64), ("pod_from", C.c_uint64),
                ("pod_to", C.c_uint64)),
    "StsSparseBatch": (("replicas_from", C.c_int32), ("replicas_to", C.c_int32), ("tmpl_from", C.c_uint64), ("tmpl_to", C.c_uint64)),
    "SvcBatch": (("flags", C.c_uint8), ("ports_from", C.c_uint64), ("ports_to", C.c_uint64)),
    "VsBatch": (("flags", C.c_uint8), ("spec_from", C.c_uint64), ("spec_to", C.c_uint64)),
    "svc_gen": (("name_hash", C.c_uint64), ("target_port", C.c_int32), ("flags", C.c_uint8), ("ports_to", C.c_uint64)),
    "vs_gen": (("prefix_hash", C.c_uint64), ("rewrite_hash", C.c_uint64), ("host_hash", C.c_uint64), ("flags", C.c_uint8),
               ("spec_to", C.c_uint64)),
    "GenBatch": (), "RestartBatch": (("flags", C.c_uint8),),
    "CullBatch": (("k_off", C.c_uint32), ("k_sec", C.c_int64), ("k_nsf", C.c_uint32), ("t_off", C.c_uint32), ("t_sec", C.c_int64),
                  ("t_nsf", C.c_uint32), ("ann_sec", C.c_int64), ("ann_nsf", C.c_uint32), ("chk_sec", C.c_int64),
                  ("chk_nsf", C.c_uint32), ("nb_flags", C.c_uint8)),
}


def batch_struct(kind, ptr, has=lambda name: False):
    cls = {"StsBatch": nbg_sts_batch, "sts_gen": nbg_sts_batch, "StsSparseBatch": nbg_sts_sparse_batch, "SvcBatch": nbg_svc_batch, "VsBatch": nbg_vs_batch,
           "svc_gen": nbg_svc_gen_batch, "vs_gen": nbg_vs_gen_batch, "GenBatch": nbg_generic_batch,
           "RestartBatch": nbg_restart_batch, "CullBatch": nbg_cull_batch}[kind]
    s = cls()
    for name, ct in _PLAIN_FIELDS[kind]:
        setattr(s, name, ptr(name, ct))
    for name in _KV_FIELDS[kind]:
        m = nbg_kvmap()
        m.off = ptr(name + ".off", C.c_uint32)
        m.ent = C.cast(ptr(name + ".ent", C.c_uint64), C.POINTER(nbg_kv))
        if has(name + ".end"):
            m.end = ptr(name + ".end", C.c_uint32)
        setattr(s, name, m)
    for side in _SPARSE_FIELDS.get(kind, ()):
        ps = nbg_podsparse()
        ps.mask, ps.off, ps.val = ptr(side + ".mask", C.c_uint64), ptr(side + ".voff", C.c_uint32), ptr(side + ".val", C.c_uint64)
        if has(side + ".vend"):
            ps.end = ptr(side + ".vend", C.c_uint32)
        setattr(s, side, ps)
    if kind == "CullBatch" and has("k_end"):
        s.k_end, s.t_end = ptr("k_end", C.c_uint32), ptr("t_end", C.c_uint32)
    return s


NBG_SYMBOLS = [
    "nbg_init", "nbg_destroy", "nbg_strerror", "nbg_last_cuda_error", "nbg_abi_version", "nbg_device_count",
    "nbg_alloc_pinned", "nbg_free_pinned", "nbg_hash64", "nbg_shard_of", "nbg_balance_shards",
    "nbg_diff_sts", "nbg_diff_svc", "nbg_diff_vs", "nbg_cull", "nbg_diff_deploy", "nbg_diff_deploy_device",
    "nbg_diff_sts_device", "nbg_diff_svc_device", "nbg_diff_vs_device", "nbg_cull_device",
    "nbg_sync", "nbg_launch_count", "nbg_ctx_stream",
    "nbg_owner_join", "nbg_join_key", "nbg_owner_join_scratch_bytes", "nbg_owner_join_device",
    "nbg_owner_join_static", "nbg_diff_sts_sparse", "nbg_diff_sts_sparse_device",
    "nbg_step", "nbg_step_device", "nbg_store_create", "nbg_store_destroy", "nbg_store_put", "nbg_store_step",
    "nbg_store_step_device", "nbg_store_get_stats", "nbg_store_check", "nbg_device_id", "nbg_diff_svc_gen", "nbg_diff_svc_gen_device", "nbg_diff_vs_gen", "nbg_diff_vs_gen_device",
    "nbg_count_running", "nbg_count_running_device", "nbg_diff_generic", "nbg_diff_generic_device", "nbg_restart_block", "nbg_restart_block_device",
    "nbg_cull_text", "nbg_parse_records", "nbg_parse_records_device", "nbg_hash_strings", "nbg_hash_strings_device", "nbg_check_step_device", "nbg_diff_sts_gen", "nbg_diff_sts_gen_device",
]

_vp = C.c_void_p


class nbg_sts_gen_inputs(C.Structure):
    _fields_ = [("prefix_hash", u64p), ("flags", u8p)]


class nbg_sts_gen_consts(C.Structure):
    _fields_ = [("working_dir", C.c_uint64), ("port_name", C.c_uint64), ("port_rest", C.c_uint64), ("env_name", C.c_uint64),
                ("add_fsgroup", C.c_uint32), ("pad", C.c_uint32)]


class nbg_step_extents(C.Structure):
    _fields_ = [(k, C.c_uint64) for k in ("sts_val_from", "sts_val_to", "sts_labels_from", "sts_labels_to", "sts_ann_from",
                                          "sts_ann_to", "svc_labels_to", "svc_ann_to", "svc_sel_to", "vs_hdr_from", "vs_hdr_to",
                                          "cull_kernels", "cull_terminals")]


def _bind_nbg(lib):
    lib.nbg_init.argtypes = [C.POINTER(C.c_int), C.c_int, C.POINTER(_vp)]
    lib.nbg_init.restype = C.c_int
    lib.nbg_destroy.argtypes = [_vp]
    lib.nbg_destroy.restype = None
    lib.nbg_strerror.argtypes = [C.c_int]
    lib.nbg_strerror.restype = C.c_char_p
    lib.nbg_last_cuda_error.restype = C.c_char_p
    lib.nbg_abi_version.restype = C.c_int
    lib.nbg_device_count.argtypes = [_vp]
    lib.nbg_alloc_pinned.argtypes = [C.c_size_t]
    lib.nbg_alloc_pinned.restype = _vp
    lib.nbg_free_pinned.argtypes = [_vp]
    lib.nbg_free_pinned.restype = None
    lib.nbg_hash64.argtypes = [C.c_char_p, C.c_size_t]
    lib.nbg_hash64.restype = C.c_uint64
    lib.nbg_shard_of.argtypes = [C.c_uint64, C.c_uint32]
    lib.nbg_shard_of.restype = C.c_uint32
    lib.nbg_balance_shards.argtypes = [_vp, _vp, C.c_uint64, C.c_uint32, _vp, _vp]
    lib.nbg_balance_shards.restype = C.c_int
    lib.nbg_check_step_device.argtypes = [_vp, C.c_int, _vp, _vp, _vp, _vp, C.c_uint64, C.POINTER(nbg_step_extents), _vp,
                                          C.POINTER(C.c_uint)]
    lib.nbg_check_step_device.restype = C.c_int
    lib.nbg_diff_sts_gen.argtypes = [_vp, C.POINTER(nbg_sts_batch), C.POINTER(nbg_sts_gen_inputs), C.POINTER(nbg_sts_gen_consts),
                                     C.c_uint64, _vp, _vp]
    lib.nbg_diff_sts_gen.restype = C.c_int
    lib.nbg_diff_sts_gen_device.argtypes = [_vp, C.c_int, C.POINTER(nbg_sts_batch), C.POINTER(nbg_sts_gen_inputs),
                                            C.POINTER(nbg_sts_gen_consts), C.c_uint64, _vp, _vp, _vp]
    lib.nbg_diff_sts_gen_device.restype = C.c_int
    lib.nbg_store_check.argtypes = [_vp, C.c_uint64, C.POINTER(C.c_uint)]
    lib.nbg_store_check.restype = C.c_int
    for name, bt in (("nbg_diff_sts", nbg_sts_batch), ("nbg_diff_svc", nbg_svc_batch), ("nbg_diff_vs", nbg_vs_batch),
                     ("nbg_diff_sts_sparse", nbg_sts_sparse_batch), ("nbg_diff_deploy", nbg_sts_batch)):
        f = getattr(lib, name)
        f.argtypes = [_vp, C.POINTER(bt), C.c_uint64, _vp, _vp]
        f.restype = C.c_int
        g = getattr(lib, name + "_device")
        g.argtypes = [_vp, C.c_int, C.POINTER(bt), C.c_uint64, _vp, _vp, _vp]
        g.restype = C.c_int
    lib.nbg_cull.argtypes = [_vp, C.POINTER(nbg_cull_batch), C.c_uint64, C.POINTER(nbg_cull_params), _vp, _vp, _vp, _vp]
    lib.nbg_cull.restype = C.c_int
    lib.nbg_cull_device.argtypes = [_vp, C.c_int, C.POINTER(nbg_cull_batch), C.c_uint64, C.POINTER(nbg_cull_params),
                                    _vp, _vp, _vp, _vp, _vp]
    lib.nbg_cull_device.restype = C.c_int
    lib.nbg_owner_join.argtypes = [_vp, _vp, C.c_uint64, _vp, C.c_uint64, _vp]
    lib.nbg_owner_join.restype = C.c_int
    lib.nbg_join_key.argtypes = [C.c_uint64, C.c_uint64]
    lib.nbg_join_key.restype = C.c_uint64
    lib.nbg_owner_join_scratch_bytes.argtypes = [C.c_uint64]
    lib.nbg_owner_join_scratch_bytes.restype = C.c_size_t
    lib.nbg_owner_join_device.argtypes = [_vp, C.c_int, _vp, C.c_uint64, _vp, C.c_uint64, _vp, _vp, C.c_size_t, _vp]
    lib.nbg_owner_join_device.restype = C.c_int
    lib.nbg_owner_join_static.argtypes = [_vp, C.c_uint64, _vp, C.c_uint64, _vp, _vp]
    lib.nbg_owner_join_static.restype = C.c_int
    lib.nbg_diff_generic.argtypes = [_vp, C.POINTER(nbg_generic_batch), C.c_uint64, C.c_int, _vp, _vp]
    lib.nbg_diff_generic.restype = C.c_int
    lib.nbg_diff_generic_device.argtypes = [_vp, C.c_int, C.POINTER(nbg_generic_batch), C.c_uint64, C.c_int, _vp, _vp, _vp]
    lib.nbg_diff_generic_device.restype = C.c_int
    lib.nbg_restart_block.argtypes = [_vp, C.POINTER(nbg_restart_batch), C.c_uint64, _vp, _vp]
    lib.nbg_restart_block.restype = C.c_int
    lib.nbg_restart_block_device.argtypes = [_vp, C.c_int, C.POINTER(nbg_restart_batch), C.c_uint64, _vp, _vp, _vp]
    lib.nbg_restart_block_device.restype = C.c_int
    lib.nbg_diff_svc_gen.argtypes = [_vp, C.POINTER(nbg_svc_gen_batch), C.c_uint64, _vp, _vp]
    lib.nbg_diff_svc_gen.restype = C.c_int
    lib.nbg_diff_svc_gen_device.argtypes = [_vp, C.c_int, C.POINTER(nbg_svc_gen_batch), C.c_uint64, _vp, _vp, _vp]
    lib.nbg_diff_svc_gen_device.restype = C.c_int
    lib.nbg_diff_vs_gen.argtypes = [_vp, C.POINTER(nbg_vs_gen_batch), C.POINTER(nbg_vs_gen_consts), C.c_uint64, _vp, _vp]
    lib.nbg_diff_vs_gen.restype = C.c_int
    lib.nbg_diff_vs_gen_device.argtypes = [_vp, C.c_int, C.POINTER(nbg_vs_gen_batch), C.POINTER(nbg_vs_gen_consts), C.c_uint64, _vp, _vp, _vp]
    lib.nbg_diff_vs_gen_device.restype = C.c_int
    lib.nbg_step.argtypes = [_vp, C.POINTER(nbg_sts_sparse_batch), C.POINTER(nbg_svc_gen_batch), C.POINTER(nbg_vs_gen_batch),
                             C.POINTER(nbg_vs_gen_consts), C.POINTER(nbg_cull_batch), C.POINTER(nbg_cull_params), C.c_uint64,
                             C.POINTER(nbg_step_outputs)]
    lib.nbg_step.restype = C.c_int
    lib.nbg_step_device.argtypes = [_vp, C.c_int, C.POINTER(nbg_sts_sparse_batch), C.POINTER(nbg_svc_gen_batch),
                                    C.POINTER(nbg_vs_gen_batch), C.POINTER(nbg_vs_gen_consts), C.POINTER(nbg_cull_batch),
                                    C.POINTER(nbg_cull_params), C.c_uint64, C.POINTER(nbg_step_outputs), _vp, _vp, _vp]
    lib.nbg_step_device.restype = C.c_int
    lib.nbg_store_create.argtypes = [_vp, C.c_int, C.c_uint64, C.POINTER(_vp)]
    lib.nbg_store_create.restype = C.c_int
    lib.nbg_store_destroy.argtypes = [_vp]
    lib.nbg_store_destroy.restype = None
    lib.nbg_store_put.argtypes = [_vp, u64p, C.c_uint64, C.POINTER(nbg_sts_sparse_batch), C.POINTER(nbg_svc_gen_batch),
                                  C.POINTER(nbg_vs_gen_batch), C.POINTER(nbg_cull_batch)]
    lib.nbg_store_put.restype = C.c_int
    lib.nbg_store_step.argtypes = [_vp, C.c_uint64, C.POINTER(nbg_vs_gen_consts), C.POINTER(nbg_cull_params), C.POINTER(nbg_step_outputs)]
    lib.nbg_store_step.restype = C.c_int
    lib.nbg_store_step_device.argtypes = [_vp, C.c_uint64, C.POINTER(nbg_vs_gen_consts), C.POINTER(nbg_cull_params),
                                          C.POINTER(nbg_step_outputs), _vp]
    lib.nbg_store_step_device.restype = C.c_int
    lib.nbg_store_get_stats.argtypes = [_vp, C.POINTER(nbg_store_stats)]
    lib.nbg_store_get_stats.restype = C.c_int
    lib.nbg_device_id.argtypes = [_vp, C.c_int]
    lib.nbg_device_id.restype = C.c_int
    lib.nbg_count_running.argtypes = [_vp, _vp, _vp, _vp, _vp, C.c_uint64, C.c_uint32, _vp]
    lib.nbg_count_running.restype = C.c_int
    lib.nbg_count_running_device.argtypes = [_vp, C.c_int, _vp, _vp, _vp, _vp, C.c_uint64, C.c_uint32, _vp, _vp]
    lib.nbg_count_running_device.restype = C.c_int
    PS = C.POINTER(nbg_strings)
    lib.nbg_parse_records.argtypes = [_vp, PS, PS, C.c_uint64, _vp, _vp, _vp]
    lib.nbg_parse_records.restype = C.c_int
    lib.nbg_parse_records_device.argtypes = [_vp, C.c_int, PS, PS, C.c_uint64, _vp, _vp, _vp, _vp]
    lib.nbg_parse_records_device.restype = C.c_int
    lib.nbg_cull_text.argtypes = [_vp, C.POINTER(nbg_cull_text_batch), C.c_uint64, C.POINTER(nbg_cull_params), _vp, _vp, _vp, _vp, _vp]
    lib.nbg_cull_text.restype = C.c_int
    lib.nbg_hash_strings.argtypes = [_vp, PS, C.c_uint64, _vp]
    lib.nbg_hash_strings.restype = C.c_int
    lib.nbg_hash_strings_device.argtypes = [_vp, C.c_int, PS, C.c_uint64, _vp, _vp]
    lib.nbg_hash_strings_device.restype = C.c_int
    lib.nbg_ctx_stream.argtypes = [_vp, C.c_int]
    lib.nbg_ctx_stream.restype = _vp
    lib.nbg_sync.argtypes = [_vp, C.c_int]
    lib.nbg_sync.restype = C.c_int
    lib.nbg_launch_count.argtypes = [_vp]
    lib.nbg_launch_count.restype = C.c_uint64
    return lib


def _bind_nbh(lib):
    cp, sz = C.c_char_p, C.c_size_t
    P = C.POINTER
    sig = {
        "nbh_obj_from_json": ([cp, cp, P(_vp), cp, sz], C.c_int),
        "nbh_obj_free": ([_vp], None),
        "nbh_obj_to_json": ([_vp], _vp),
        "nbh_str_free": ([_vp], None),
        "nbh_obj_get": ([_vp, cp, P(_vp)], C.c_int),
        "nbh_obj_erase": ([_vp, cp], C.c_int),
        "nbh_obj_set_str": ([_vp, cp, cp], C.c_int),
        "nbh_obj_digest": ([_vp], C.c_uint64),
        "nbh_generate_statefulset": ([_vp, P(nbh_env), P(_vp), cp, sz], C.c_int),
        "nbh_generate_service": ([_vp, P(_vp), cp, sz], C.c_int),
        "nbh_generate_virtualservice": ([_vp, P(nbh_env), P(_vp), cp, sz], C.c_int),
        "nbh_odh_mutate": ([_vp, C.c_int], C.c_int),
        "nbh_odh_mutate_proxy": ([_vp, C.c_int, C.c_char_p, C.c_char_p, C.c_char_p], C.c_int),
        "nbh_odh_build": ([_vp, C.c_char_p, C.c_char_p, C.c_char_p, C.c_char_p, C.c_char_p, C.c_char_p, C.c_char_p, P(_vp)], C.c_int),
        "nbh_apply_sts_fields": ([_vp, _vp, C.c_uint, C.c_int], C.c_int),
        "nbh_apply_svc_fields": ([_vp, _vp], C.c_int),
        "nbh_apply_vs_fields": ([_vp, _vp, C.c_uint], C.c_int),
        "nbh_sts_new": ([], _vp), "nbh_sts_free": ([_vp], None),
        "nbh_sts_add": ([_vp, _vp, _vp], C.c_int), "nbh_sts_view": ([_vp, P(nbg_sts_batch)], C.c_uint64),
        "nbh_sts_view_sparse": ([_vp, P(nbg_sts_sparse_batch)], C.c_uint64),
        "nbh_svc_new": ([], _vp), "nbh_svc_free": ([_vp], None),
        "nbh_svc_add": ([_vp, _vp, _vp], C.c_int), "nbh_svc_view": ([_vp, P(nbg_svc_batch)], C.c_uint64),
        "nbh_svc_view_gen": ([_vp, P(nbg_svc_gen_batch)], C.c_uint64),
        "nbh_vs_view_gen": ([_vp, P(nbg_vs_gen_batch)], C.c_uint64),
        "nbh_vs_gen_consts": ([P(nbh_env), P(nbg_vs_gen_consts)], C.c_int),
        "nbh_vs_new": ([], _vp), "nbh_vs_free": ([_vp], None),
        "nbh_vs_add": ([_vp, _vp, _vp], C.c_int), "nbh_vs_view": ([_vp, P(nbg_vs_batch)], C.c_uint64),
        "nbh_gen_new": ([], _vp), "nbh_gen_free": ([_vp], None),
        "nbh_gen_add": ([_vp, _vp, _vp, cp, cp, C.c_int], C.c_int), "nbh_gen_view": ([_vp, P(nbg_generic_batch)], C.c_uint64),
        "nbh_rb_new": ([], _vp), "nbh_rb_free": ([_vp], None),
        "nbh_rb_add": ([_vp, _vp, _vp, _vp, C.c_int], C.c_int), "nbh_rb_view": ([_vp, P(nbg_restart_batch)], C.c_uint64),
        "nbh_obj_clone": ([_vp], _vp),
        "nbh_cull_new": ([], _vp), "nbh_cull_free": ([_vp], None),
        "nbh_cull_add_json": ([_vp, cp, cp, sz], C.c_int), "nbh_cull_view": ([_vp, P(nbg_cull_batch)], C.c_uint64),
        "nbh_cull_records": ([_vp, P(C.c_uint64), P(C.c_uint64)], C.c_uint64),
        "nbh_cull_new_last_activity": ([_vp, C.c_uint64, C.c_uint8, C.c_uint32, C.c_int64, C.c_int32, cp, sz], C.c_int),
        "nbh_parse_rfc3339": ([cp, P(C.c_int64), P(C.c_int32), P(C.c_int32)], C.c_int),
        "nbh_format_rfc3339": ([C.c_int64, C.c_int32, cp, sz], C.c_int),
        "nbh_parse_records": ([u32p, u8p, u32p, u8p, C.c_uint64, _vp, _vp, _vp], C.c_int),
        "nbh_fleet_new": ([C.c_int, C.c_uint64, C.c_uint64, C.c_uint32], _vp), "nbh_fleet_free": ([_vp], None),
        "nbh_fleet_get": ([_vp, C.c_uint64, C.c_int, P(_vp)], C.c_int),
        "nbh_fleet_drift_class": ([_vp, C.c_uint64], C.c_int),
        "nbh_fleet_ns_hash": ([_vp, C.c_uint64], C.c_uint64),
        "nbh_fleet_pack": ([_vp, u64p, C.c_uint64, C.c_uint64, P(_vp), P(_vp), P(_vp)], C.c_int),
        "nbh_fleet_pack_lean": ([_vp, u64p, C.c_uint64, C.c_uint64, P(_vp), P(_vp), P(_vp)], C.c_int),
        "nbh_sts_is_lean": ([_vp], C.c_int),
        "nbh_sts_add_gen": ([_vp, _vp, _vp, _vp], C.c_int),
        "nbh_sts_add_notebook": ([_vp, _vp, _vp, _vp], C.c_int),
        "nbh_sts_view_gen": ([_vp, P(nbg_sts_gen_inputs)], C.c_uint64),
        "nbh_sts_gen_consts": ([_vp, P(nbg_sts_gen_consts)], C.c_int),
        "nbh_pack_objects": ([P(_vp), C.c_uint64, C.c_int, P(_vp), P(_vp), P(_vp)], C.c_int),
        "nbh_fleet_ns_table": ([_vp, u64p, u64p], C.c_uint32),
        "nbh_fleet_shard_indices_by_table": ([_vp, u32p, C.c_uint32, u64p, C.c_uint64], C.c_uint64),
        "nbh_fleet_pack_timing": ([P(C.c_double), P(C.c_double)], None),
        "nbh_cullfleet_new": ([C.c_uint64, C.c_uint64, C.c_int64, C.c_int32, C.c_double], _vp),
        "nbh_cullfleet_free": ([_vp], None),
        "nbh_cullfleet_json": ([_vp, C.c_uint64], _vp),
        "nbh_cullfleet_pack": ([_vp, u64p, C.c_uint64, C.c_uint64, P(_vp)], C.c_int),
        "nbh_threads": ([], C.c_int),
        "nbh_join_key_owner": ([_vp], C.c_uint64),
        "nbh_join_key_controlled": ([_vp], C.c_uint64),
        "nbh_fleet_join_lists": ([_vp, C.c_uint64, u64p, u64p, i64p, C.c_uint64], C.c_uint64),
        "nbh_fleet_shard_indices": ([_vp, C.c_uint32, C.c_uint32, u64p, C.c_uint64], C.c_uint64),
        "nbh_fleet_objects": ([_vp, u64p, C.c_uint64, C.c_uint64, P(_vp)], C.c_int),
    }
    for name, (args, res) in sig.items():
        f = getattr(lib, name)
        f.argtypes = args
        f.restype = res
    return lib


NBH_SYMBOLS = [
    "nbh_obj_from_json", "nbh_obj_free", "nbh_obj_to_json", "nbh_str_free", "nbh_obj_get", "nbh_obj_erase", "nbh_obj_set_str", "nbh_obj_digest",
    "nbh_generate_statefulset", "nbh_generate_service", "nbh_generate_virtualservice", "nbh_odh_mutate", "nbh_odh_mutate_proxy", "nbh_odh_build",
    "nbh_sts_new", "nbh_sts_free", "nbh_sts_add", "nbh_sts_view", "nbh_sts_view_sparse", "nbh_svc_new", "nbh_svc_free", "nbh_svc_add",
    "nbh_svc_view", "nbh_svc_view_gen", "nbh_vs_view_gen", "nbh_vs_gen_consts", "nbh_vs_new", "nbh_vs_free", "nbh_vs_add", "nbh_vs_view", "nbh_gen_new", "nbh_gen_free", "nbh_gen_add", "nbh_gen_view", "nbh_rb_new", "nbh_rb_free", "nbh_rb_add", "nbh_rb_view", "nbh_obj_clone", "nbh_cull_new", "nbh_cull_free",
    "nbh_cull_add_json", "nbh_cull_view", "nbh_cull_records", "nbh_cull_new_last_activity", "nbh_parse_rfc3339",
    "nbh_format_rfc3339", "nbh_parse_records", "nbh_fleet_new", "nbh_fleet_free", "nbh_fleet_get", "nbh_fleet_drift_class",
    "nbh_apply_sts_fields", "nbh_apply_svc_fields", "nbh_apply_vs_fields", "nbh_fleet_ns_hash", "nbh_fleet_pack", "nbh_fleet_pack_lean", "nbh_sts_is_lean", "nbh_pack_objects", "nbh_sts_add_gen", "nbh_sts_add_notebook", "nbh_sts_view_gen", "nbh_sts_gen_consts", "nbh_fleet_ns_table", "nbh_fleet_shard_indices_by_table", "nbh_fleet_pack_timing", "nbh_cullfleet_new", "nbh_cullfleet_free", "nbh_cullfleet_json",
    "nbh_cullfleet_pack", "nbh_threads", "nbh_fleet_shard_indices", "nbh_fleet_objects",
    "nbh_join_key_owner", "nbh_join_key_controlled", "nbh_fleet_join_lists",
]

_nbg = None
_nbh = None


def nbg():
    """libnbgraft.so (kernels + C ABI)."""
    global _nbg
    if _nbg is None:
        _nbg = _bind_nbg(_load("libnbgraft.so"))
    return _nbg


def nbh():
    """libnbgraft_host.so (host packer)."""
    global _nbh
    if _nbh is None:
        _nbh = _bind_nbh(_load("libnbgraft_host.so"))
    return _nbh
