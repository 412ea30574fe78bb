/*
 * nbgraft.h — C ABI of libnbgraft.so: batched desired-vs-live "needs update"
 * predicates and the culler's idleness decision, as sm_100a CUDA kernels.
 *
 * The reference (harshad16/odh-kubeflow) is a Go controller with no FFI; the
 * calls this library replaces are ordinary in-process Go calls, one object at
 * a time.  Each entry point below cites the reference function it batches.
 * Path abbreviations (all under the reference tree):
 *   COMMON = components/common/reconcilehelper/util.go
 *   NBCTL  = components/notebook-controller/controllers/notebook_controller.go
 *   CULL   = components/notebook-controller/controllers/culling_controller.go
 *
 * Conventions
 *   - every buffer is caller-allocated, borrowed for the duration of the call,
 *     and is a flat array of fixed-width integers: no pointers inside buffers
 *     (cgo pointer-passing rules are satisfied by construction);
 *   - status is an int: 0 = OK, negative = error (nbg_strerror); nothing
 *     aborts or throws across this boundary;
 *   - there is NO CPU fallback: without a usable CUDA device nbg_init fails
 *     with NBG_E_NODEVICE and nothing else can be called;
 *   - a context may be used from several OS threads at once (the reference
 *     has up to two concurrent callers, the "notebook" and "Culler"
 *     controllers); host-buffer calls serialise per device inside the library.
 *   - strings never cross the boundary: the host packs them as 64-bit hashes
 *     (nbg_hash64; the empty / missing string is 0, no other string is 0).
 */
#ifndef NBGRAFT_H
#define NBGRAFT_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* 4: the Service ports tile is 8 slots (was 16: ports[1:] are one digest now), and the packer of libnbgraft_host.so
 *    spends one word on a Quantity, a container port, a volume mount and a plain volume (the tiles of version 3 and
 *    version 4 packers do not compare with each other; nothing on the device side changed but NBG_SVC_SLOTS);
 *    nbg_diff_sts_gen added.
 * 3: additions only (nbg_balance_shards, nbg_check_step_device, nbg_store_check); nbg_store_put now refuses a slot
 *    named twice (NBG_E_INVAL) instead of leaving the outcome open.  2: spans (end pointers) in the batch structs. */
#define NBG_ABI_VERSION 4

/* ---- status codes ------------------------------------------------------ */
#define NBG_OK              0
#define NBG_E_NODEVICE    (-1) /* no CUDA device / driver: there is no CPU path   */
#define NBG_E_INVAL       (-2) /* null pointer, bad size, bad parameter           */
#define NBG_E_NOMEM       (-3) /* device or pinned-host allocation failed         */
#define NBG_E_CUDA        (-4) /* a CUDA call failed; nbg_last_cuda_error()       */
#define NBG_E_NILREPLICAS (-5) /* an object had a nil Spec.Replicas: the Go code
                                  would panic at COMMON:123; reason bit7 marks it */
#define NBG_E_RANGE       (-6) /* CSR offsets not monotone, a sparse run that disagrees with its
                                  mask, or a slot index past a store's capacity             */
#define NBG_E_TOOBIG      (-7) /* batch exceeds a documented maximum              */

/* ---- schema constants (frozen; tests/test_abi.py checks them) ---------- */
#define NBG_POD_SLOTS   192u  /* u64 slots per PodSpec tile  (1536 B per side)   */
#define NBG_SVC_SLOTS     8u  /* u64 slots per Service ports tile (64 B): ports[0] as six fields (name, protocol,
                               * appProtocol, port|nodePort<<32, targetPort int|type<<32, digest of the rest), the
                               * length of the list (0xFFFFFFFF = nil) and one digest of ports[1:] (0 = none) */
#define NBG_VS_SLOTS      8u  /* u64 slots per VirtualService spec tile (64 B)   */
#define NBG_REPLICAS_NIL INT32_MIN /* marks a nil *int32 Replicas pointer        */
#define NBG_ARGMAX_NONE 0xFFFFFFFFu

/* one map entry: key hash, value hash ("" == 0). Entries of one object are
 * sorted by key; duplicate keys cannot occur (they are Go map keys). */
typedef struct nbg_kv { uint64_t key; uint64_t val; } nbg_kv;

/* Per-object string maps (labels, annotations, selector, headers).
 *   end == NULL (CSR): object i owns ent[off[i] .. off[i+1]).  off has n+1 elements, off[0] may
 *     be non-zero (a slice of a larger batch).  The only form the host-buffer entry points take.
 *   end != NULL (spans; *_device entry points only): object i owns ent[off[i] .. end[i]).  off and
 *     end have n elements; runs may lie anywhere in `ent`, in any order, with gaps between them —
 *     the layout of a resident fleet whose objects are replaced in place (nbg_store_*). */
typedef struct nbg_kvmap { const uint32_t* off; const nbg_kv* ent; const uint32_t* end; } nbg_kvmap;

/* ---- reason bits ------------------------------------------------------- */
/* StatefulSet (COMMON:107-134 + caller pre-step NBCTL:188-194) */
#define NBG_STS_LABELS      0x01u /* L: live label whose desired value differs (COMMON:109-113) */
#define NBG_STS_ANNOTATIONS 0x02u /* A: same on annotations (COMMON:116-120)                    */
#define NBG_STS_REPLICAS    0x04u /* R: *from.Spec.Replicas != *to.Spec.Replicas (COMMON:123)   */
#define NBG_STS_PODSPEC     0x08u /* P: !reflect.DeepEqual(PodSpec, PodSpec) (COMMON:128)       */
#define NBG_STS_TMPL_LABELS 0x10u /* D1': caller copied template labels (NBCTL:190-194); NOT part
                                     of the returned bool                                       */
#define NBG_STS_NILREPLICAS 0x80u /* object would panic in Go; other bits are still valid except R */
#define NBG_STS_UPDATE_MASK 0x0Fu
/* Service (COMMON:166-195) */
#define NBG_SVC_LABELS      0x01u
#define NBG_SVC_ANNOTATIONS 0x02u
#define NBG_SVC_SELECTOR    0x04u /* S: !DeepEqual(to.Spec.Selector, from.Spec.Selector) (COMMON:184) */
#define NBG_SVC_PORTS       0x08u /* Pt: !DeepEqual(to.Spec.Ports, from.Spec.Ports) (COMMON:189)       */
#define NBG_SVC_UPDATE_MASK 0x0Fu
/* VirtualService (COMMON:199-219) */
#define NBG_VS_SPEC_DIFFERS 0x01u /* both specs present and !DeepEqual (COMMON:214)              */
#define NBG_VS_TO_MISSING   0x02u /* to.spec missing or not a map: copied, returns true (:209-212) */
#define NBG_VS_FROM_MISSING 0x04u /* from.spec missing or not a map: returns false (:200-206)     */
/* bool returned by Go = (reason & FROM_MISSING) ? 0 : (reason & 3) != 0 */

/* flags byte shared by the diff batches (one per object) */
#define NBG_F_FROM_SEL_NIL   0x01u /* Service: from.Spec.Selector is a nil map (vs empty)   */
#define NBG_F_TO_SEL_NIL     0x02u
#define NBG_F_FROM_SPEC_MISS 0x04u /* VS: NestedMap(from,"spec") not found / not a map     */
#define NBG_F_TO_SPEC_MISS   0x08u
#define NBG_F_FROM_HDR_NIL   0x10u /* VS: headers.request.set absent (vs empty map)        */
#define NBG_F_TO_HDR_NIL     0x20u

/* ---- batches ----------------------------------------------------------- */
/* StatefulSet desired ("from") vs live ("to") pairs. */
typedef struct nbg_sts_batch {
  const int32_t*  replicas_from;   /* n; NBG_REPLICAS_NIL = nil pointer            */
  const int32_t*  replicas_to;     /* n                                            */
  const uint64_t* tmpl_from;       /* n; digest of Spec.Template.Labels (nil != {}) */
  const uint64_t* tmpl_to;         /* n                                            */
  const uint64_t* pod_from;        /* n * NBG_POD_SLOTS, object-major              */
  const uint64_t* pod_to;          /* n * NBG_POD_SLOTS                            */
  nbg_kvmap labels_from, labels_to; /* ObjectMeta.Labels                            */
  nbg_kvmap ann_from, ann_to;       /* ObjectMeta.Annotations                       */
} nbg_sts_batch;

/* The same StatefulSet pairs with the PodSpec tile stored sparsely.  A synthetic
 * Kubeflow PodSpec fills ~24 of the 192 slots, an ODH one ~68, so the dense tile is
 * mostly zeros; here each side keeps a 192-bit presence mask (bit s = slot s is
 * non-zero) and only the non-zero slots, in slot order:
 *     mask[3*i .. 3*i+3), val[off[i] .. off[i+1])
 * Every run is zero-padded to an EVEN number of values (so every off[] is even and
 * every run starts 16-byte aligned; `val` itself must be 16-byte aligned): a run
 * holds popcount(mask) values, plus one 0 when that count is odd.
 * Dense tiles are equal  <=>  masks are equal and values are equal, exactly (no
 * extra hashing), so the result is bit-identical to nbg_diff_sts on the same pairs. */
typedef struct nbg_podsparse {
  const uint64_t* mask;   /* n * 3                                   */
  const uint32_t* off;    /* n + 1, in u64 units; off[0] may be != 0 */
  const uint64_t* val;    /* off[n] - off[0] values                  */
  const uint32_t* end;    /* NULL (CSR) or n run ends, as in nbg_kvmap: run i = val[off[i] .. end[i]),
                             every off[i] even (runs 16-byte aligned); *_device entry points only */
} nbg_podsparse;
typedef struct nbg_sts_sparse_batch {
  const int32_t*  replicas_from;
  const int32_t*  replicas_to;
  const uint64_t* tmpl_from;
  const uint64_t* tmpl_to;
  nbg_podsparse pod_from, pod_to;
  nbg_kvmap labels_from, labels_to;
  nbg_kvmap ann_from, ann_to;
} nbg_sts_sparse_batch;

typedef struct nbg_svc_batch {
  const uint8_t*  flags;           /* n; NBG_F_*_SEL_NIL                           */
  const uint64_t* ports_from;      /* n * NBG_SVC_SLOTS                            */
  const uint64_t* ports_to;
  nbg_kvmap labels_from, labels_to;
  nbg_kvmap ann_from, ann_to;
  nbg_kvmap sel_from, sel_to;      /* Spec.Selector                                */
} nbg_svc_batch;

typedef struct nbg_vs_batch {
  const uint8_t*  flags;           /* n; NBG_F_*_SPEC_MISS, NBG_F_*_HDR_NIL        */
  const uint64_t* spec_from;       /* n * NBG_VS_SLOTS                             */
  const uint64_t* spec_to;
  nbg_kvmap hdr_from, hdr_to;      /* spec.http[0].headers.request.set             */
} nbg_vs_batch;

/* Culler inputs.  An instant is (sec since Unix epoch, nsec) as Go's time.Time
 * holds it after time.Parse; *_nsf words carry nsec in bits 0..29 and two flags. */
#define NBG_NSEC_MASK     0x3FFFFFFFu
#define NBG_REC_PARSE_OK  0x40000000u /* last_activity parsed as RFC3339 (CULL:332,339) */
#define NBG_REC_IDLE      0x80000000u /* execution_state == "idle" (CULL:320); terminals: unused */
#define NBG_ANN_PARSE_OK  0x40000000u /* last-activity annotation parsed (CULL:229,353)  */
#define NBG_ANN_STOP_SET  0x80000000u /* kubeflow-resource-stopped key present (CULL:493-502) */
#define NBG_CHK_PARSE_OK  0x40000000u /* check-timestamp parsed (CULL:212; failure => zero Time) */
#define NBG_CHK_PRESENT   0x80000000u /* check-timestamp key present (CULL:208)           */
#define NBG_NB_LA_PRESENT  0x01u      /* last-activity key present (CULL:442, annotationsExist) */
#define NBG_NB_POD_MISSING 0x02u      /* pod "<name>-0" not found (CULL:121-135); flow mode only */

typedef struct nbg_cull_batch {
  const uint32_t* k_off;   /* n+1: kernels of notebook i are records k_off[i]..k_off[i+1) */
  const int64_t*  k_sec;   /* K                                                          */
  const uint32_t* k_nsf;   /* K                                                          */
  const uint32_t* t_off;   /* n+1: terminals                                             */
  const int64_t*  t_sec;   /* T                                                          */
  const uint32_t* t_nsf;   /* T                                                          */
  const int64_t*  ann_sec; /* n: last-activity annotation                                */
  const uint32_t* ann_nsf; /* n                                                          */
  const int64_t*  chk_sec; /* n: last_activity_check_timestamp annotation                */
  const uint32_t* chk_nsf; /* n                                                          */
  const uint8_t*  nb_flags;/* n                                                          */
  const uint32_t* k_end;   /* NULL (CSR) or n: kernels of notebook i are k_off[i]..k_end[i) — spans, as in
                              nbg_kvmap; *_device entry points only                       */
  const uint32_t* t_end;   /* same for terminals                                          */
} nbg_cull_batch;

typedef struct nbg_cull_params {
  int64_t now_sec;           /* time.Now(): one instant for the whole batch            */
  int32_t now_nsec;          /* 0..999999999                                            */
  int32_t cull_idle_min;     /* CULL_IDLE_TIME (CULL:37,525-558), minutes               */
  int32_t check_period_min;  /* IDLENESS_CHECK_PERIOD (CULL:39), minutes                */
  int32_t mode;              /* NBG_CULL_PREDICATES or NBG_CULL_RECONCILE               */
} nbg_cull_params;
#define NBG_CULL_PREDICATES 0 /* C4,C5,C7,C8 evaluated unconditionally on every notebook */
#define NBG_CULL_RECONCILE  1 /* the decision flow of CullingReconciler.Reconcile, CULL:86-203 */
#define NBG_CULL_MAX_MINUTES 100000000 /* |minutes| bound so minutes*60e9 ns fits Go's int64 Duration */

/* action byte written per notebook */
#define NBG_ACT_SRC_MASK   0x03u /* new last-activity: 0 unchanged, 1 now (CULL:386), 2 kernels[argmax]
                                    (CULL:399), 3 terminals[argmax] (CULL:425)                      */
#define NBG_ACT_PERIOD     0x04u /* cullingCheckPeriodHasPassed (CULL:207-217)                      */
#define NBG_ACT_CULL       0x08u /* notebookIsIdle after the updates (CULL:220-241) => setStopAnnotation */
#define NBG_ACT_STOPPED    0x10u /* StopAnnotationIsSet at entry (CULL:104): annotations removed     */
#define NBG_ACT_INIT       0x20u /* !annotationsExist (CULL:141): both annotations := now            */
#define NBG_ACT_NOPOD      0x40u /* pod missing (CULL:121): annotations removed                      */

/* ---- context ----------------------------------------------------------- */
typedef struct nbg_ctx nbg_ctx;

/* Create a context on the listed CUDA devices (n_devices >= 1; devices may be
 * NULL meaning {0}).  One stream, one event pair and grow-only staging buffers
 * per device.  Replaces nothing in the reference (there is no device there). */
int nbg_init(const int* devices, int n_devices, nbg_ctx** out);
void nbg_destroy(nbg_ctx* ctx);
const char* nbg_strerror(int status);
/* text of the last CUDA error seen on this thread ("" if none) */
const char* nbg_last_cuda_error(void);
int nbg_abi_version(void);
int nbg_device_count(const nbg_ctx* ctx);

/* Pinned host memory for batch buffers (faster H2D; optional). */
void* nbg_alloc_pinned(size_t bytes);
void  nbg_free_pinned(void* p);

/* ---- hashing / sharding (host side, pure) ------------------------------ */
/* 64-bit string hash used by every packer: "" -> 0, nothing else -> 0. */
uint64_t nbg_hash64(const void* data, size_t len);
/* namespace -> shard (GPU) map, the batch analogue of the per-namespace List
 * at NBCTL:159: shard = mix64(ns_hash) % n_shards. */
uint32_t nbg_shard_of(uint64_t ns_hash, uint32_t n_shards);
/* Size-aware namespace -> shard table for fleets whose namespaces are very unequal (a Zipf fleet hashed into 8
 * shards is 1.9x out of balance, and the step ends when the largest shard does).  The partition unit stays the
 * namespace, as the per-namespace List of NBCTL:159 requires: every namespace goes to exactly one shard.
 * Longest-processing-time rule: namespaces in order of decreasing size (ties: increasing ns_hash) each go to the
 * shard that is lightest so far (ties: lowest shard number), so the result depends only on the (hash, size)
 * pairs, not on their order in the arrays.  shard_out[k] in [0, n_shards) for namespace k; load_out (nullable,
 * n_shards entries) receives the shards' total sizes.  Max load <= mean + largest namespace.  Host side, pure;
 * needs no context. */
int nbg_balance_shards(const uint64_t* ns_hash, const uint64_t* ns_size, uint64_t n_namespaces, uint32_t n_shards,
                       uint32_t* shard_out, uint64_t* load_out);

/* ---- host-buffer entry points (H2D + kernel + D2H inside the call) ----- */
/* Batched CopyStatefulSetFields (COMMON:107-134) with the caller's template-
 * label pre-step (NBCTL:188-194).  reason[i] = NBG_STS_* bits;
 * bitmap bit i (word i/64, bit i%64) = (reason[i] & NBG_STS_UPDATE_MASK) != 0,
 * i.e. the bool the Go function returns.  bitmap has ceil(n/64) words. */
int nbg_diff_sts(nbg_ctx* ctx, const nbg_sts_batch* in, uint64_t n,
                 uint8_t* reason, uint64_t* bitmap);
/* CopyDeploymentSetFields (COMMON:136-163) on the same batch layout: labels, annotations and PodSpec as in
 * CopyStatefulSetFields, but the Go code compares the Replicas POINTERS (COMMON:152), so NBG_STS_REPLICAS is set
 * unless both are nil, nothing panics (NBG_STS_NILREPLICAS and NBG_STS_TMPL_LABELS are never set).  The reference
 * itself has no caller for this function; it is here because the path's file defines it. */
int nbg_diff_deploy(nbg_ctx* ctx, const nbg_sts_batch* in, uint64_t n, uint8_t* reason, uint64_t* bitmap);
int nbg_diff_deploy_device(nbg_ctx* ctx, int device, const nbg_sts_batch* in, uint64_t n,
                           uint8_t* reason, uint64_t* bitmap, void* stream);
/* CopyStatefulSetFields again, on sparsely stored PodSpec tiles (nbg_sts_sparse_batch): outputs identical to
 * nbg_diff_sts, ~4x fewer bytes per object to move over PCIe and to read from HBM.  Every run must be as long
 * as its mask says (popcount, rounded up to even): NBG_E_RANGE otherwise. */
int nbg_diff_sts_sparse(nbg_ctx* ctx, const nbg_sts_sparse_batch* in, uint64_t n,
                        uint8_t* reason, uint64_t* bitmap);
/* Batched CopyServiceFields (COMMON:166-195). */
int nbg_diff_svc(nbg_ctx* ctx, const nbg_svc_batch* in, uint64_t n,
                 uint8_t* reason, uint64_t* bitmap);
/* Batched CopyVirtualService (COMMON:199-219). */
int nbg_diff_vs(nbg_ctx* ctx, const nbg_vs_batch* in, uint64_t n,
                uint8_t* reason, uint64_t* bitmap);
/* Batched culler decision: allKernelsAreIdle, getNotebookRecentTime,
 * compareAnnotationTimeToResource, updateTimestampFrom{Kernels,Terminals}Activity,
 * notebookIsIdle, cullingCheckPeriodHasPassed (CULL:207-241,315-428).
 * action[i] = NBG_ACT_* bits; argmax[i] = index (within the notebook's own
 * list) of the record that supplies the new last-activity for src 2/3, else
 * NBG_ARGMAX_NONE; cull_bitmap bit i = NBG_ACT_CULL; period_bitmap bit i =
 * NBG_ACT_PERIOD. */
int nbg_cull(nbg_ctx* ctx, const nbg_cull_batch* in, uint64_t n,
             const nbg_cull_params* params,
             uint8_t* action, uint32_t* argmax,
             uint64_t* cull_bitmap, uint64_t* period_bitmap);

/* Batched "find the StatefulSet this Notebook controls": the loop over the
 * namespace's StatefulSets with metav1.IsControlledBy at NBCTL:157-170, as a hash
 * join.  nb_key[i] = nbg_join_key(hash64(namespace), hash64(metadata.uid)) of
 * Notebook i; sts_key[j] = the same for StatefulSet j's namespace and the UID of
 * its controller owner reference, or 0 when it has none.  match[i] = the lowest j
 * with sts_key[j] == nb_key[i] (the first list entry, as the reference's `break`
 * picks), or NBG_ARGMAX_NONE => the Notebook has no StatefulSet yet (NBCTL:172-183
 * creates it).  m < 2^31. */
int nbg_owner_join(nbg_ctx* ctx, const uint64_t* nb_key, uint64_t n,
                   const uint64_t* sts_key, uint64_t m, uint32_t* match);
/* never 0 */
uint64_t nbg_join_key(uint64_t ns_hash, uint64_t uid_hash);

/* ---- desired Service / VirtualService generated ON THE DEVICE -------- */
/* generateService (NBCTL:525-552) and generateVirtualService (NBCTL:558-658) are pure functions
 * of a handful of Notebook fields, so their outputs need not be packed and shipped: these two
 * entry points take only those fields for the desired side, synthesise the desired tile in
 * registers and compare it with the live object — identical results to nbg_diff_svc /
 * nbg_diff_vs on (generate*(notebook), live), ~43 % / ~29 % fewer bytes per object.
 * (generateStatefulSet is NOT done this way: its PodSpec is the Notebook's own.) */
typedef struct nbg_svc_gen_batch {
  const uint64_t* name_hash;     /* n: hash64(notebook name) — selector {"statefulset": name}      */
  const int32_t*  target_port;   /* n: containers[0].ports[0].containerPort, or 8888 (NBCTL:527-531) */
  const uint8_t*  flags;         /* n: NBG_F_TO_SEL_NIL                                             */
  const uint64_t* ports_to;      /* n * NBG_SVC_SLOTS                                               */
  nbg_kvmap labels_to, ann_to, sel_to;
} nbg_svc_gen_batch;
typedef struct nbg_vs_gen_batch {
  const uint64_t* prefix_hash;   /* n: hash64("/notebook/<ns>/<name>/")                             */
  const uint64_t* rewrite_hash;  /* n: hash64(rewrite annotation, or the prefix) (NBCTL:569-574)    */
  const uint64_t* host_hash;     /* n: hash64("<name>.<ns>.svc.<CLUSTER_DOMAIN>")                   */
  nbg_kvmap hdr_from;            /* headers.request.set from the annotation JSON (NBCTL:602-618)    */
  const uint8_t*  flags;         /* n: NBG_F_TO_SPEC_MISS, NBG_F_TO_HDR_NIL                         */
  const uint64_t* spec_to;       /* n * NBG_VS_SLOTS                                                */
  nbg_kvmap hdr_to;
} nbg_vs_gen_batch;
/* what does not depend on the notebook: hash64(ISTIO_HOST), hash64(ISTIO_GATEWAY) and the
 * packer's residual digest of the generated spec's shape (nbh_vs_gen_consts computes them) */
typedef struct nbg_vs_gen_consts { uint64_t istio_host_hash, gateway_hash, residual; } nbg_vs_gen_consts;
int nbg_diff_svc_gen(nbg_ctx* ctx, const nbg_svc_gen_batch* in, uint64_t n, uint8_t* reason, uint64_t* bitmap);
int nbg_diff_vs_gen(nbg_ctx* ctx, const nbg_vs_gen_batch* in, const nbg_vs_gen_consts* consts, uint64_t n,
                    uint8_t* reason, uint64_t* bitmap);

/* ---- desired StatefulSet generated ON THE DEVICE (second half) ---------------------------
 * generateStatefulSet (NBCTL:433-523) as a patch of the NOTEBOOK's own PodSpec tile.  The desired PodSpec is a deep
 * copy of the Notebook's with at most four edits — containers[0].workingDir defaulted (NBCTL:495-497), a nil
 * containers[0].ports replaced by the default port (:498-506), NB_PREFIX appended to containers[0].env unless a
 * variable of that name exists (:417-431), securityContext {fsGroup: 100} when it is nil and ADD_FSGROUP (:514-521) —
 * and every one of them lands in named slots of the tile, so the kernel makes the desired tile in registers from
 * the Notebook's and compares it with the live one.  Spec.Replicas is 0 iff the stop annotation is present (:434-437).
 * `in` is an nbg_sts_batch (dense tiles) whose `from` side holds what the NOTEBOOK supplies:
 *     pod_from    = the tile of the Notebook's spec.template.spec (not of the generated StatefulSet)
 *     labels_from = the Notebook's labels (the desired StatefulSet's labels are a copy of them, NBCTL:474-477)
 *     tmpl_from   = digest of the desired template labels ({statefulset, notebook-name, workbenches} + labels)
 *     replicas_from, ann_from: NOT READ (derived from gen->flags / the desired object has no annotations)
 * Same results as nbg_diff_sts on (pack(generateStatefulSet(notebook)), live), bit for bit.
 * This is a capability, not the fast path: it needs DENSE tiles (the sparse sweep cannot shift runs at the
 * insertion points) and saves no byte — the Notebook's tile is as large as the desired one — so nbg_step keeps
 * taking the desired side from the host. */
#define NBG_GEN_STOPPED    0x01u /* annotation kubeflow-resource-stopped present: desired replicas 0              */
#define NBG_GEN_PREPATCHED 0x02u /* pod_from already IS the desired tile: the host generated it (containers[0].env
                                    has 14 or more entries, more than the tile shows by name)                     */
typedef struct nbg_sts_gen_inputs {
  const uint64_t* prefix_hash;   /* n: hash64("/notebook/<namespace>/<name>"), the value of NB_PREFIX */
  const uint8_t*  flags;         /* n: NBG_GEN_*                                                       */
} nbg_sts_gen_inputs;
/* what does not depend on the notebook (nbh_sts_gen_consts computes them with the packer's own functions) */
typedef struct nbg_sts_gen_consts {
  uint64_t working_dir;          /* hash64("/home/jovyan")                                   */
  uint64_t port_name, port_rest; /* hash64("notebook-port"), the packer's digest of the rest of the default port */
  uint64_t env_name;             /* hash64("NB_PREFIX")                                      */
  uint32_t add_fsgroup, pad;     /* ADD_FSGROUP unset or "true" (NBCTL:514)                  */
} nbg_sts_gen_consts;
int nbg_diff_sts_gen(nbg_ctx* ctx, const nbg_sts_batch* in, const nbg_sts_gen_inputs* gen, const nbg_sts_gen_consts* consts,
                     uint64_t n, uint8_t* reason, uint64_t* bitmap);
int nbg_diff_sts_gen_device(nbg_ctx* ctx, int device, const nbg_sts_batch* in, const nbg_sts_gen_inputs* gen,
                            const nbg_sts_gen_consts* consts, uint64_t n, uint8_t* reason, uint64_t* bitmap, void* stream);

/* One reconcile step over n notebooks in ONE call: nbg_diff_sts_sparse + nbg_diff_svc_gen +
 * nbg_diff_vs_gen + nbg_cull (any of the batches may be NULL to skip it), queued as a single
 * copy/compute pipeline so the PCIe link stays busy across the four — the batch analogue of
 * one pass of NotebookReconciler.Reconcile (NBCTL:152-240) and CullingReconciler.Reconcile
 * (CULL:156-195) over the fleet.  Same outputs as the four separate calls. */
typedef struct nbg_step_outputs {
  uint8_t* sts_reason;  uint64_t* sts_bitmap;
  uint8_t* svc_reason;  uint64_t* svc_bitmap;
  uint8_t* vs_reason;   uint64_t* vs_bitmap;
  uint8_t* action;      uint32_t* argmax;     uint64_t* cull_bitmap;  uint64_t* period_bitmap;
} nbg_step_outputs;
int nbg_step(nbg_ctx* ctx, const nbg_sts_sparse_batch* sts, const nbg_svc_gen_batch* svc, const nbg_vs_gen_batch* vs,
             const nbg_vs_gen_consts* vs_consts, const nbg_cull_batch* cull, const nbg_cull_params* params, uint64_t n,
             const nbg_step_outputs* out);
/* The same step over batches that already live in device memory, asynchronous and ordered on
 * `stream` like a single launch (work queued on `stream` afterwards sees every output).  Inside,
 * the StatefulSet kernel runs on `stream` and the three small, latency-bound predicates run
 * beside it on streams the context owns (event fork/join, no host synchronisation).  Any batch may
 * be NULL.  Nil replicas are reported in sts_reason (NBG_STS_NILREPLICAS), not in the status.
 * ev_sts_start / ev_sts_stop: optional cudaEvent_t handles recorded on `stream` right around the
 * StatefulSet kernel, for callers that want that launch's duration inside a running step. */
int nbg_step_device(nbg_ctx* ctx, int device, const nbg_sts_sparse_batch* sts, const nbg_svc_gen_batch* svc,
                    const nbg_vs_gen_batch* vs, const nbg_vs_gen_consts* vs_consts, const nbg_cull_batch* cull,
                    const nbg_cull_params* params, uint64_t n, const nbg_step_outputs* out, void* stream,
                    void* ev_sts_start /*nullable*/, void* ev_sts_stop /*nullable*/);

/* ---- resident fleet ----------------------------------------------------------------------
 * The four batches of a reconcile step for up to `capacity` notebooks, kept in device memory
 * and edited IN PLACE: the caller packs an object when its informer delivers it (the expensive
 * part, once per change), puts the packed row into its slot, and every tick runs the step over
 * the whole resident fleet without moving it again.  The reference has no counterpart: it
 * re-derives and re-compares one object per reconcile (NBCTL:152-240, CULL:86-203); this is the
 * batch form of its informer cache + resync.
 * Layout: fixed-width words in per-notebook arrays; every variable-length part (sparse PodSpec
 * runs, label/annotation/selector/header maps, kernel and terminal records) in an append-only
 * heap addressed by per-notebook spans (off,end) — a replaced run is appended and the old one
 * becomes garbage; a heap that fills up is compacted (live runs copied in slot order) and grown
 * when more than half of it is still live.  A store belongs to one device of the context and must be
 * destroyed before it.  put and step may be called from different threads (the store serialises them), and
 * a put issued while a step enqueued with nbg_store_step_device is still running is ordered behind it. */
typedef struct nbg_store nbg_store;
int  nbg_store_create(nbg_ctx* ctx, int device, uint64_t capacity, nbg_store** out);
void nbg_store_destroy(nbg_store* st);
/* Slots idx[0..m) take rows 0..m of the given HOST batches (CSR form, m objects each; a NULL batch
 * leaves that predicate's rows as they are).  Every idx must be < capacity (NBG_E_RANGE) and named once
 * (NBG_E_INVAL for a duplicate: two rows would race for the slot); both are checked before anything is
 * touched.  The rows travel as ONE pinned staging copy and ONE kernel launch per chunk of up to 65 536 rows /
 * 48 MB (two staging slots, so the gather of a chunk overlaps the copy of the one before).
 * Returns when the rows are on the device (the host buffers may be reused at once).  After an error other
 * than those two the slots named in this call hold unspecified rows (put them again); all other slots are intact. */
int  nbg_store_put(nbg_store* st, const uint64_t* idx, uint64_t m, const nbg_sts_sparse_batch* sts,
                   const nbg_svc_gen_batch* svc, const nbg_vs_gen_batch* vs, const nbg_cull_batch* cull);
/* One step over slots [0,n): outputs as nbg_step, into HOST buffers (synchronous) ... */
int  nbg_store_step(nbg_store* st, uint64_t n, const nbg_vs_gen_consts* vs_consts, const nbg_cull_params* params,
                    const nbg_step_outputs* out);
/* ... or into DEVICE buffers, enqueued on `stream` like nbg_step_device. */
int  nbg_store_step_device(nbg_store* st, uint64_t n, const nbg_vs_gen_consts* vs_consts, const nbg_cull_params* params,
                           const nbg_step_outputs* out, void* stream);
typedef struct nbg_store_stats {
  uint64_t capacity;        /* slots                                              */
  uint64_t fixed_bytes;     /* per-slot arrays                                    */
  uint64_t heap_bytes;      /* allocated for the heaps                            */
  uint64_t live_bytes;      /* of which belong to a current run                   */
  uint64_t garbage_bytes;   /* appended and since replaced                        */
  uint64_t compactions;     /* heap compactions so far                            */
  uint64_t puts, rows_put;  /* calls and rows since creation                      */
} nbg_store_stats;
int  nbg_store_get_stats(const nbg_store* st, nbg_store_stats* out);
/* nbg_check_step_device (below) on the store's own view of slots [0,n): every span inside the used part of its
 * heap, sparse runs aligned and as long as their masks say.  A self-check for tests and debug builds. */
int  nbg_store_check(nbg_store* st, uint64_t n, unsigned* what /*nullable*/);
/* CUDA ordinal of the context's device `device` (for callers that allocate next to it). */
int  nbg_device_id(const nbg_ctx* ctx, int device);

/* Metrics.scrape (components/notebook-controller/pkg/metrics/metrics.go:82-98) as a segmented
 * count: counts[ns] = number of StatefulSets j with ns_id[j] == ns whose pod-template label
 * "notebook-name" is present (flags[j] & 1) and equals the StatefulSet's name
 * (label_hash[j] == name_hash[j]).  ns_id are dense namespace ids < n_namespaces assigned by the
 * host (ids >= n_namespaces are ignored).  A namespace whose count stays 0 gets no gauge update
 * in the reference (stsCache has no entry for it). */
int nbg_count_running(nbg_ctx* ctx, const uint32_t* ns_id, const uint64_t* name_hash, const uint64_t* label_hash,
                      const uint8_t* flags, uint64_t m, uint32_t n_namespaces, uint32_t* counts);

/* ---- the ODH comparator family ------------------------------ */
/* Pairs (a = desired, b = found) of arbitrary objects reduced to two key/value
 * vectors each: "labels" and "spec".
 *  NBG_GEN_DEEPEQUAL  each vector is a FLATTENED value tree (nbh_flatten: one entry per
 *      leaf / container marker, key = hash of the path, sorted by key), so vectors are equal
 *      <=> reflect.DeepEqual of the trees.  This is CompareNotebookHTTPRoutes
 *      (odh notebook_route.go:134-138), CompareNotebookNetworkPolicies (notebook_network.go:
 *      125-129), CompareNotebookReferenceGrants (notebook_referencegrant.go:72-76) — labels
 *      and spec — and the RoleBinding subjects test (notebook_rbac.go:131).
 *  NBG_GEN_CONFIGMAP  the vectors are plain string maps (hash(key), hash(value)); the test is
 *      the loop of notebook_kube_rbac_auth.go:240-266: lengths differ, or some desired
 *      (k, v) has found[k] != v with a missing key reading as "" — on `spec` (= Data) and on
 *      `labels`.
 * reason bit0 = labels differ, bit1 = spec/data differ; bitmap bit = reason != 0 (the
 * reference's Compare* functions return the negation: true when equal). */
typedef struct nbg_generic_batch { nbg_kvmap labels_a, labels_b, spec_a, spec_b; } nbg_generic_batch;
#define NBG_GEN_DEEPEQUAL 0
#define NBG_GEN_CONFIGMAP 1
#define NBG_GEN_LABELS 0x01u
#define NBG_GEN_SPEC   0x02u
int nbg_diff_generic(nbg_ctx* ctx, const nbg_generic_batch* in, uint64_t n, int mode,
                     uint8_t* reason, uint64_t* bitmap);

/* The webhook's restart-block predicate, maybeRestartRunningNotebook (odh notebook_webhook.go:
 * 505-564), for a fleet.  Per Notebook three PodSpecs as sparse SEMANTIC tiles (the packer's
 * semantic flavour: nil == empty slices/maps, Quantity by value — equal tiles <=>
 * equality.Semantic.DeepEqual): the stored one, the one the user sent, and that one after
 * the webhook's injections. */
typedef struct nbg_restart_batch {
  const uint8_t* flags;                      /* n; NBG_RB_CREATE / _STOPPED / _RESTARTING */
  nbg_podsparse old_spec, updated_spec, mutated_spec;
} nbg_restart_batch;
#define NBG_RB_CREATE      0x01u /* req.Operation == Create (:515)                              */
#define NBG_RB_STOPPED     0x02u /* mutated notebook carries kubeflow-resource-stopped (:521)   */
#define NBG_RB_RESTARTING  0x04u /* ... carries notebooks.opendatahub.io/notebook-restart (:527) */
/* decision byte */
#define NBG_RB_BLOCKED         0x01u /* keep the updated PodSpec, report UpdatesPending (:559-563) */
#define NBG_RB_EXTERNAL_CHANGE 0x02u /* the user's update already changes the pod template (:548)  */
#define NBG_RB_NO_CHANGE       0x04u /* old == updated == mutated (:554)                           */
int nbg_restart_block(nbg_ctx* ctx, const nbg_restart_batch* in, uint64_t n, uint8_t* decision, uint64_t* blocked_bitmap);

/* ---- strings on the device (the host work right in front of the predicates) ---- */
/* A table of R strings as CSR: string r is bytes[off[r] .. off[r+1]).  `bytes` must be
 * 16-byte aligned and its allocation padded to a multiple of 16 bytes (the kernels copy
 * whole 16-byte vectors). */
typedef struct nbg_strings { const uint32_t* off; const uint8_t* bytes; } nbg_strings;
/* time.Parse(time.RFC3339, last_activity) for R kernel/terminal records (CULL:332,339) and,
 * when `state` is given, execution_state == "idle" (CULL:320): fills exactly the
 * (sec, nsf) words nbg_cull_batch wants (NBG_REC_PARSE_OK, NBG_REC_IDLE, nsec), plus the
 * zone offset in seconds of every record (nullable) for formatting results. */
int nbg_parse_records(nbg_ctx* ctx, const nbg_strings* last_activity, const nbg_strings* state /*nullable*/,
                      uint64_t R, int64_t* sec, uint32_t* nsf, int32_t* zone /*nullable*/);
/* The culler decision straight from the strings the reference handles (CULL:86-241,315-428):
 * kernel/terminal last_activity + execution_state texts and the two annotation values per
 * notebook; the device parses them (as nbg_parse_records) and decides (as nbg_cull) in one call,
 * so the host never parses a timestamp.  ann_text / chk_text hold one string per notebook ("" when
 * the key is absent — absence itself is in nb_flags / NBG_NBT_* below).  Outputs as nbg_cull, plus
 * the zone offset (seconds) of the record that supplies the new last-activity (0 otherwise), which
 * is all the host needs to format it as Format(time.RFC3339) does. */
typedef struct nbg_cull_text_batch {
  const uint32_t* k_off;        /* n+1: kernels of notebook i are records k_off[i]..k_off[i+1)      */
  nbg_strings k_last_activity;  /* K strings                                                        */
  nbg_strings k_state;          /* K strings                                                        */
  const uint32_t* t_off;        /* n+1: terminals                                                   */
  nbg_strings t_last_activity;  /* T strings                                                        */
  nbg_strings ann_text;         /* n strings: notebooks.kubeflow.org/last-activity                  */
  nbg_strings chk_text;         /* n strings: notebooks.kubeflow.org/last_activity_check_timestamp  */
  const uint8_t* nb_flags;      /* n: NBG_NB_* | NBG_NBT_*                                           */
} nbg_cull_text_batch;
#define NBG_NBT_STOP_SET    0x10u /* kubeflow-resource-stopped present                               */
#define NBG_NBT_CHK_PRESENT 0x20u /* check-timestamp key present                                      */
int nbg_cull_text(nbg_ctx* ctx, const nbg_cull_text_batch* in, uint64_t n, const nbg_cull_params* params,
                  uint8_t* action, uint32_t* argmax, int32_t* argmax_zone, uint64_t* cull_bitmap, uint64_t* period_bitmap);
/* nbg_hash64 of every string of the table (label keys/values, env values, ...). */
int nbg_hash_strings(nbg_ctx* ctx, const nbg_strings* strings, uint64_t R, uint64_t* out);

/* ---- device-pointer entry points (asynchronous, no copies) ------------- */
/* Same kernels on buffers already resident in HBM.  Every pointer inside the
 * batch and every output is a DEVICE pointer on `device` (an index into the
 * list given to nbg_init); `stream` is a cudaStream_t, passed through as is
 * (NULL = the CUDA default stream; nbg_ctx_stream() gives the context's own
 * non-blocking stream).  The call only enqueues work.  CSR offsets are NOT validated
 * here (the host-buffer calls validate them before upload). */
int nbg_diff_sts_device(nbg_ctx* ctx, int device, const nbg_sts_batch* in, uint64_t n,
                        uint8_t* reason, uint64_t* bitmap, void* stream);
int nbg_diff_sts_sparse_device(nbg_ctx* ctx, int device, const nbg_sts_sparse_batch* in, uint64_t n,
                               uint8_t* reason, uint64_t* bitmap, void* stream);
int nbg_diff_svc_device(nbg_ctx* ctx, int device, const nbg_svc_batch* in, uint64_t n,
                        uint8_t* reason, uint64_t* bitmap, void* stream);
int nbg_diff_vs_device(nbg_ctx* ctx, int device, const nbg_vs_batch* in, uint64_t n,
                       uint8_t* reason, uint64_t* bitmap, void* stream);
int nbg_cull_device(nbg_ctx* ctx, int device, const nbg_cull_batch* in, uint64_t n,
                    const nbg_cull_params* params,
                    uint8_t* action, uint32_t* argmax,
                    uint64_t* cull_bitmap, uint64_t* period_bitmap, void* stream);
/* Device form of nbg_owner_join.  `scratch` is device memory for the join's hash
 * table: nbg_owner_join_scratch_bytes(m) bytes, 8-byte aligned, contents
 * irrelevant on entry (the call clears it), clobbered on return. */
size_t nbg_owner_join_scratch_bytes(uint64_t m);
int nbg_owner_join_device(nbg_ctx* ctx, int device, const uint64_t* nb_key, uint64_t n,
                          const uint64_t* sts_key, uint64_t m, uint32_t* match,
                          void* scratch, size_t scratch_bytes, void* stream);
int nbg_diff_generic_device(nbg_ctx* ctx, int device, const nbg_generic_batch* in, uint64_t n, int mode,
                            uint8_t* reason, uint64_t* bitmap, void* stream);
int nbg_diff_svc_gen_device(nbg_ctx* ctx, int device, const nbg_svc_gen_batch* in, uint64_t n, uint8_t* reason,
                            uint64_t* bitmap, void* stream);
int nbg_diff_vs_gen_device(nbg_ctx* ctx, int device, const nbg_vs_gen_batch* in, const nbg_vs_gen_consts* consts,
                           uint64_t n, uint8_t* reason, uint64_t* bitmap, void* stream);
int nbg_count_running_device(nbg_ctx* ctx, int device, const uint32_t* ns_id, const uint64_t* name_hash,
                             const uint64_t* label_hash, const uint8_t* flags, uint64_t m, uint32_t n_namespaces,
                             uint32_t* counts, void* stream);
int nbg_restart_block_device(nbg_ctx* ctx, int device, const nbg_restart_batch* in, uint64_t n, uint8_t* decision,
                             uint64_t* blocked_bitmap, void* stream);
int nbg_parse_records_device(nbg_ctx* ctx, int device, const nbg_strings* last_activity, const nbg_strings* state,
                             uint64_t R, int64_t* sec, uint32_t* nsf, int32_t* zone, void* stream);
int nbg_hash_strings_device(nbg_ctx* ctx, int device, const nbg_strings* strings, uint64_t R, uint64_t* out,
                            void* stream);
/* The checks the host-buffer entry points make on their arguments, for batches that are already on the device
 * (the *_device entry points above only enqueue and therefore trust what they are given: a run that ends past its
 * payload array makes a kernel read outside it).  One thread per object looks at every (start, end) pair of its
 * rows: start <= end, end <= the payload's length as stated in `ext`, sparse runs even-aligned and as long as their
 * mask says.  Synchronous (it waits for its own kernel); returns NBG_E_RANGE when anything is wrong, with the kinds
 * of fault OR-ed into *what (nullable): 1 start > end, 2 past the payload, 4 misaligned, 8 length != mask.
 * Any batch may be NULL.  No counterpart in the reference (Go slices carry their bounds).  Meant for the moments
 * where a host-buffer call would have checked for free: after uploading a batch from a new packer, in tests, in a
 * debug build of the caller. */
typedef struct nbg_step_extents {          /* lengths, in elements, of the payload arrays the batches point into */
  uint64_t sts_val_from, sts_val_to;                                    /* u64 values   */
  uint64_t sts_labels_from, sts_labels_to, sts_ann_from, sts_ann_to;    /* nbg_kv       */
  uint64_t svc_labels_to, svc_ann_to, svc_sel_to;                       /* nbg_kv       */
  uint64_t vs_hdr_from, vs_hdr_to;                                      /* nbg_kv       */
  uint64_t cull_kernels, cull_terminals;                                /* records      */
} nbg_step_extents;
int nbg_check_step_device(nbg_ctx* ctx, int device, const nbg_sts_sparse_batch* sts, const nbg_svc_gen_batch* svc,
                          const nbg_vs_gen_batch* vs, const nbg_cull_batch* cull, uint64_t n,
                          const nbg_step_extents* ext, void* stream, unsigned* what);

/* Owner join that allocates nothing — it has no counterpart in the reference.
 * For callers that may not call the CUDA allocator while they run: code on a
 * reconcile path with a latency bound (cudaMalloc/cudaFree synchronise the whole
 * device), or a controller sharing its GPU with the notebooks it manages, where
 * an allocation made on demand can fail because the workloads got the memory
 * first.  Same inputs, outputs and result as nbg_owner_join_device (all pointers
 * are DEVICE pointers on the CURRENT CUDA device; no context is needed), but the
 * hash table is a fixed array inside the library, so:
 *   - at most NBG_JOIN_STATIC_MAX_STS StatefulSets per call; more is rejected
 *     with NBG_E_TOOBIG and nothing is launched;
 *   - every GPU the entry point has been called on holds the table (3 MiB) until exit;
 *   - the FIRST call on each device is the set-up: it looks the table's address up
 *     (cudaGetSymbolAddress), and with CUDA's lazy module loading that lookup is what makes
 *     the driver load this part of the library and claim the 3 MiB on that device.  It can
 *     therefore block, or fail with NBG_E_NOMEM when the device is already full.  Make it a
 *     warm-up call at start-up (n = 1 is enough; or run with CUDA_MODULE_LOADING=EAGER),
 *     before memory gets tight and before any stream capture.  Later calls touch neither the
 *     loader nor the allocator;
 *   - all calls on a device share the one table (each call clears it before
 *     use): two calls whose GPU work overlaps — different streams, or several
 *     threads — race.  Order them on one stream or with events.
 * The call itself performs no cudaMalloc/cudaFree and only enqueues work. */
#define NBG_JOIN_STATIC_MAX_STS 131072u
int nbg_owner_join_static(const uint64_t* nb_key, uint64_t n, const uint64_t* sts_key, uint64_t m,
                          uint32_t* match, void* stream);
/* The context's own (non-blocking) run stream on `device`, as a cudaStream_t. */
void* nbg_ctx_stream(nbg_ctx* ctx, int device);
/* Block until everything enqueued on the context's own stream of `device` is done. */
int nbg_sync(nbg_ctx* ctx, int device);
/* Number of kernels this library has launched since nbg_init (all devices). */
uint64_t nbg_launch_count(const nbg_ctx* ctx);

#ifdef __cplusplus
}
#endif
#endif /* NBGRAFT_H */
