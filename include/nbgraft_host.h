/*
 * nbgraft_host.h — C ABI of libnbgraft_host.so: the HOST side of the graft.
 *
 * In the reference this work is done by ordinary Go code before the predicates
 * run: decoding objects from the API server, building the desired objects
 * (generateStatefulSet/Service/VirtualService, NBCTL:433-658) and, in this
 * build, flattening them into the structure-of-arrays batches that
 * libnbgraft.so consumes.  No Go toolchain exists in the build image, so the
 * host side is C++ behind this C ABI (a Go packer would produce byte-identical
 * batches from the same objects; INTEGRATION.md shows the cgo binding).
 * Nothing here touches CUDA; the library loads on a machine without a GPU.
 *
 * Handles are opaque and owned by the caller (free with the matching *_free).
 * Functions returning int return 0 on success, -1 on failure with a message
 * in `err` (if err != NULL and errlen > 0).
 */
#ifndef NBGRAFT_HOST_H
#define NBGRAFT_HOST_H

#include <stddef.h>
#include <stdint.h>

#include "nbgraft.h"

#ifdef __cplusplus
extern "C" {
#endif

typedef struct nbh_obj nbh_obj;     /* one decoded object (a Go value tree)          */
typedef struct nbh_sts nbh_sts;     /* packed StatefulSet pairs  -> nbg_sts_batch    */
typedef struct nbh_svc nbh_svc;     /* packed Service pairs      -> nbg_svc_batch    */
typedef struct nbh_vs nbh_vs;       /* packed VirtualService pairs -> nbg_vs_batch   */
typedef struct nbh_cull nbh_cull;   /* packed culler inputs      -> nbg_cull_batch   */
typedef struct nbh_gen nbh_gen;     /* ODH comparator pairs      -> nbg_generic_batch */
typedef struct nbh_rb nbh_rb;       /* webhook restart-block     -> nbg_restart_batch */
typedef struct nbh_fleet nbh_fleet; /* seeded synthetic Notebook fleet (diff side)   */
typedef struct nbh_cullfleet nbh_cullfleet; /* seeded synthetic culler fleet         */

/* ---- objects ----------------------------------------------------------- */
/* type: "Notebook", "StatefulSet", "Service", "PodSpec", ... or "Unstructured"
 * (untyped JSON, what CopyVirtualService receives). */
int nbh_obj_from_json(const char* type, const char* json, nbh_obj** out, char* err, size_t errlen);
void nbh_obj_free(nbh_obj* o);
/* Debug rendering of the value tree as JSON (Ptr -> its pointee, Null -> null);
 * caller frees with nbh_str_free. */
char* nbh_obj_to_json(const nbh_obj* o);
void nbh_str_free(char* s);
/* Copy of the sub-object at a dotted field path ("spec.template.spec"); -1 if absent. */
int nbh_obj_get(const nbh_obj* o, const char* dotted_path, nbh_obj** out);
/* Edit in place: remove the field / set a string field at a dotted path; list elements are
 * addressed by number ("metadata.ownerReferences.0.uid"). */
int nbh_obj_erase(nbh_obj* o, const char* dotted_path);
int nbh_obj_set_str(nbh_obj* o, const char* dotted_path, const char* value);
/* canonical digest of the whole value: equal digests <=> reflect.DeepEqual (mod 2^-64) */
uint64_t nbh_obj_digest(const nbh_obj* o);

/* ---- desired-object builders ------------------------------------------- */
typedef struct nbh_env {
  int add_fsgroup;             /* ADD_FSGROUP unset or "true" (NBCTL:514)      */
  const char* cluster_domain;  /* CLUSTER_DOMAIN; NULL = "cluster.local"       */
  const char* istio_host;      /* ISTIO_HOST; NULL or "" = "*"                 */
  const char* istio_gateway;   /* ISTIO_GATEWAY; NULL or "" = kubeflow/kubeflow-gateway */
} nbh_env;
/* generateStatefulSet (NBCTL:433-523); env may be NULL for the defaults */
int nbh_generate_statefulset(const nbh_obj* notebook, const nbh_env* env, nbh_obj** out, char* err, size_t errlen);
/* generateService (NBCTL:525-552) */
int nbh_generate_service(const nbh_obj* notebook, nbh_obj** out, char* err, size_t errlen);
/* generateVirtualService (NBCTL:558-658) */
int nbh_generate_virtualservice(const nbh_obj* notebook, const nbh_env* env, nbh_obj** out, char* err, size_t errlen);
/* ODH mutating-webhook injections (ODHWH:105-114,177-326,666-781; notebook_runtime.go:216-285;
 * notebook_dspa_secret.go:375-450; notebook_feast_config.go:34-147) applied in place. flags: 1 lock, 2 CA bundle,
 * 4 runtime images, 8 elyra secret, 16 kube-rbac-proxy, 32 the label-gated Feast block of Handle (ODHWH:431-444:
 * label opendatahub.io/feast-integration == "true" mounts <name>-feast-config, otherwise a mounted one is removed). */
int nbh_odh_mutate(nbh_obj* notebook, int flags);
/* The same, followed by InjectProxyConfigEnvVars (ODHWH:566-616) with the cluster Proxy's three status fields:
 * HTTP_PROXY / HTTPS_PROXY / NO_PROXY are appended to the image container's env unless already present (a present
 * one keeps its value: the reference's update at :583-585 writes to the range copy). */
int nbh_odh_mutate_proxy(nbh_obj* notebook, int flags, const char* http_proxy, const char* https_proxy, const char* no_proxy);

/* Desired objects of the ODH controller, the `from` side of its comparators: what = "HTTPRoute"
 * (NewNotebookHTTPRoute, notebook_route.go:50-131), "NetworkPolicy" / "KubeRbacProxyNetworkPolicy"
 * (notebook_network.go:132-211), "ReferenceGrant" (notebook_referencegrant.go:39-69) or "RoleBinding"
 * (NewRoleBinding, notebook_rbac.go:36-59).  central_namespace = the controller's namespace; gateway_* = the
 * NOTEBOOK_GATEWAY_NAME / NOTEBOOK_GATEWAY_NAMESPACE overrides (NULL or "" = data-science-gateway / openshift-ingress);
 * the three role strings are read for "RoleBinding" only. */
int nbh_odh_build(const nbh_obj* notebook, const char* what, const char* central_namespace, const char* gateway_name,
                  const char* gateway_namespace, const char* rolebinding_name, const char* role_kind,
                  const char* role_name, nbh_obj** out);

/* The COPY half of CopyStatefulSetFields / CopyServiceFields / CopyVirtualService (COMMON:114-131,173-192,
 * 208-217): what the Go functions write into `to`, given the reason bits the compare kernel returned for the pair
 * (nbg_diff_sts / nbg_diff_svc / nbg_diff_vs).  caller_prestep != 0 also copies the template labels when the
 * kernel set NBG_STS_TMPL_LABELS, as Reconcile does before the call (NBCTL:190-194). */
int nbh_apply_sts_fields(const nbh_obj* from, nbh_obj* to, unsigned reason, int caller_prestep);
int nbh_apply_svc_fields(const nbh_obj* from, nbh_obj* to);
int nbh_apply_vs_fields(const nbh_obj* from, nbh_obj* to, unsigned reason);

/* ---- packed batches ---------------------------------------------------- */
nbh_sts* nbh_sts_new(void);
void nbh_sts_free(nbh_sts* b);
int nbh_sts_add(nbh_sts* b, const nbh_obj* from, const nbh_obj* to);
uint64_t nbh_sts_view(const nbh_sts* b, nbg_sts_batch* out);   /* returns n; pointers live as long as b is unchanged */

/* generateStatefulSet on the device (nbg_diff_sts_gen): append a pair whose `from` side is taken from the NOTEBOOK
 * (its own PodSpec tile, its labels, the digest of the desired template labels) instead of from a generated
 * StatefulSet; nbh_sts_view gives the nbg_sts_batch, nbh_sts_view_gen the per-notebook generator inputs (0 when
 * the batch was not filled with nbh_sts_add_gen alone), nbh_sts_gen_consts the constants.  -1 when
 * generateStatefulSet would fail or leave its common path (no containers; name longer than 52 characters). */
int nbh_sts_add_gen(nbh_sts* b, const nbh_obj* notebook, const nbh_obj* to, const nbh_env* env /*NULL = defaults*/);
uint64_t nbh_sts_view_gen(const nbh_sts* b, nbg_sts_gen_inputs* out);
int nbh_sts_gen_consts(const nbh_env* env /*NULL = defaults*/, nbg_sts_gen_consts* out);
/* For the sparse step (nbg_diff_sts_sparse, nbg_step): the pair nbh_sts_add(generateStatefulSet(notebook), to) would
 * append, array for array, WITHOUT building the StatefulSet - the Notebook's PodSpec is packed once and the desired
 * tile made from it on the host with the same few slot edits the kernel above makes (notebook_controller.go:495-521).
 * Works on lean batches; not to be mixed with nbh_sts_add_gen in one batch.  -1 as nbh_sts_add_gen. */
int nbh_sts_add_notebook(nbh_sts* b, const nbh_obj* notebook, const nbh_obj* to, const nbh_env* env /*NULL = defaults*/);

/* the same pairs with sparsely stored PodSpec tiles (nbg_diff_sts_sparse) */
uint64_t nbh_sts_view_sparse(const nbh_sts* b, nbg_sts_sparse_batch* out);

nbh_svc* nbh_svc_new(void);
void nbh_svc_free(nbh_svc* b);
int nbh_svc_add(nbh_svc* b, const nbh_obj* from, const nbh_obj* to);
uint64_t nbh_svc_view(const nbh_svc* b, nbg_svc_batch* out);

/* views for the device-generated desired side (nbg_diff_svc_gen / nbg_diff_vs_gen): valid when
 * every `from` object was produced by generate_service / generate_virtualservice */
uint64_t nbh_svc_view_gen(const nbh_svc* b, nbg_svc_gen_batch* out);
uint64_t nbh_vs_view_gen(const nbh_vs* b, nbg_vs_gen_batch* out);
int nbh_vs_gen_consts(const nbh_env* env /*NULL = defaults*/, nbg_vs_gen_consts* out);

nbh_vs* nbh_vs_new(void);
void nbh_vs_free(nbh_vs* b);
int nbh_vs_add(nbh_vs* b, const nbh_obj* from, const nbh_obj* to);
uint64_t nbh_vs_view(const nbh_vs* b, nbg_vs_batch* out);

/* ODH comparator pairs (desired a, found b): the sub-objects at labels_path / spec_path of both,
 * flattened (mode NBG_GEN_DEEPEQUAL) or as plain string maps (mode NBG_GEN_CONFIGMAP). */
nbh_gen* nbh_gen_new(void);
void nbh_gen_free(nbh_gen* b);
int nbh_gen_add(nbh_gen* b, const nbh_obj* a, const nbh_obj* found, const char* labels_path, const char* spec_path, int mode);
uint64_t nbh_gen_view(const nbh_gen* b, nbg_generic_batch* out);

/* Webhook restart-block inputs (ODHWH:505-564): the stored Notebook, the one the user sent and the
 * one after the webhook's mutations (spec.template.spec of each, as semantic sparse tiles). */
nbh_rb* nbh_rb_new(void);
void nbh_rb_free(nbh_rb* b);
int nbh_rb_add(nbh_rb* b, const nbh_obj* old_nb, const nbh_obj* updated_nb, const nbh_obj* mutated_nb, int create);
uint64_t nbh_rb_view(const nbh_rb* b, nbg_restart_batch* out);
nbh_obj* nbh_obj_clone(const nbh_obj* o);

/* One notebook as the culler sees it, as JSON:
 *   {"annotations": {..}|null, "pod_missing": bool,
 *    "kernels": [{"last_activity": "...", "execution_state": "..."}],
 *    "terminals": [{"last_activity": "..."}]}
 * Timestamps are parsed as Go's time.Parse(time.RFC3339, ...) would (CULL:212,229,332,339). */
nbh_cull* nbh_cull_new(void);
void nbh_cull_free(nbh_cull* b);
int nbh_cull_add_json(nbh_cull* b, const char* json, char* err, size_t errlen);
uint64_t nbh_cull_view(const nbh_cull* b, nbg_cull_batch* out);
/* number of kernel / terminal records */
uint64_t nbh_cull_records(const nbh_cull* b, uint64_t* kernels, uint64_t* terminals);
/* The new last-activity annotation value for notebook i given the kernel's
 * outputs (what CULL:386,399,425 would have written); "" when unchanged.
 * Writes a NUL-terminated string into buf; returns its length or -1. */
int nbh_cull_new_last_activity(const nbh_cull* b, uint64_t i, uint8_t action, uint32_t argmax,
                               int64_t now_sec, int32_t now_offset_sec, char* buf, size_t buflen);

/* time.Parse(time.RFC3339, s): 1 if it parses, else 0 */
int nbh_parse_rfc3339(const char* s, int64_t* sec, int32_t* nsec, int32_t* offset_sec);
/* Host-side counterpart of nbg_parse_records on all host threads (same CSR inputs and outputs). */
int nbh_parse_records(const uint32_t* off, const uint8_t* bytes, const uint32_t* s_off, const uint8_t* s_bytes,
                      uint64_t R, int64_t* sec, uint32_t* nsf, int32_t* zone);
/* t.Format(time.RFC3339) */
int nbh_format_rfc3339(int64_t sec, int32_t offset_sec, char* buf, size_t buflen);

/* ---- owner join keys (NBCTL:157-170) -------------------------------------- */
/* key of an owner (Notebook): nbg_join_key(hash(namespace), hash(metadata.uid)) */
uint64_t nbh_join_key_owner(const nbh_obj* owner);
/* key of a dependent (StatefulSet): its namespace + the UID of its controller owner reference
 * (metav1.GetControllerOf: first reference with controller == true), 0 if there is none */
uint64_t nbh_join_key_controlled(const nbh_obj* obj);

/* ---- synthetic fleets ------------------------------------- */
/* kind: 2 = Kubeflow loadtest-shaped notebooks (configs #2/#5), 4 = ODH-webhook-shaped (#4) */
nbh_fleet* nbh_fleet_new(int kind, uint64_t seed, uint64_t n, uint32_t n_namespaces);
void nbh_fleet_free(nbh_fleet* f);
/* which: 0 notebook, 1 sts desired, 2 sts live, 3 svc desired, 4 svc live, 5 vs desired, 6 vs live */
int nbh_fleet_get(const nbh_fleet* f, uint64_t i, int which, nbh_obj** out);
int nbh_fleet_drift_class(const nbh_fleet* f, uint64_t i);
uint64_t nbh_fleet_ns_hash(const nbh_fleet* f, uint64_t i);   /* nbg_hash64(namespace name) */
/* Pack objects idx[0..m) (idx == NULL: the range [i0, i0+m)) with all host threads. */
int nbh_fleet_pack(const nbh_fleet* f, const uint64_t* idx, uint64_t i0, uint64_t m,
                   nbh_sts** sts /*nullable*/, nbh_svc** svc /*nullable*/, nbh_vs** vs /*nullable*/);
/* The same, keeping only what the reconcile step reads of the StatefulSet pairs: the SPARSE PodSpec tiles
 * (nbh_sts_view_sparse); the dense tiles are not stored and nbh_sts_view returns NULL for pod_from / pod_to.
 * 0.6 KB instead of 3.7 KB per pair: what lets the 10 M-notebook fleet of BASELINE config #5 sit in host memory. */
int nbh_fleet_pack_lean(const nbh_fleet* f, const uint64_t* idx, uint64_t i0, uint64_t m,
                        nbh_sts** sts /*nullable*/, nbh_svc** svc /*nullable*/, nbh_vs** vs /*nullable*/);
int nbh_sts_is_lean(const nbh_sts* b);
/* Pack m notebooks whose objects the caller holds: objs[k*7 + which] as nbh_fleet_objects lays them out (which 0,
 * the Notebook, is not read).  This is the packing a controller does when its informer delivers objects. */
int nbh_pack_objects(const nbh_obj* const* objs, uint64_t m, int lean,
                     nbh_sts** sts /*nullable*/, nbh_svc** svc /*nullable*/, nbh_vs** vs /*nullable*/);
/* CPU seconds, summed over threads, that the LAST nbh_fleet_pack spent (a) synthesising the objects
 * (Notebook -> generateStatefulSet/Service/VirtualService -> drifted live copies) and (b) packing
 * them into the batch arrays.  Only (b) is work a real caller has; the bench reports both. */
void nbh_fleet_pack_timing(double* make_cpu_s, double* pack_cpu_s);

/* Indices i in [0,n) whose namespace maps to `shard` of `n_shards` (nbg_shard_of), ascending.
 * Returns the count; writes at most cap of them (call with out=NULL to size). */
uint64_t nbh_fleet_shard_indices(const nbh_fleet* f, uint32_t n_shards, uint32_t shard, uint64_t* out, uint64_t cap);
/* The fleet's namespaces: ns_hash[k] = nbg_hash64(name of namespace k), ns_size[k] = notebooks in it (either may
 * be NULL).  Returns the number of namespaces.  Input of nbg_shard_of / nbg_balance_shards. */
uint32_t nbh_fleet_ns_table(const nbh_fleet* f, uint64_t* ns_hash, uint64_t* ns_size);
/* As nbh_fleet_shard_indices, for an explicit namespace -> shard table (shard_of_ns[k] for namespace k). */
uint64_t nbh_fleet_shard_indices_by_table(const nbh_fleet* f, const uint32_t* shard_of_ns, uint32_t shard,
                                          uint64_t* out, uint64_t cap);
/* Materialise objects of m notebooks at once: out[k*7 + which] (see nbh_fleet_get). */
int nbh_fleet_objects(const nbh_fleet* f, const uint64_t* idx, uint64_t i0, uint64_t m, nbh_obj** out);

/* Synthetic StatefulSet list for notebooks [0,n) in seeded random order (see hostapi.cpp);
 * returns its length m (call with NULL outputs to size; at most cap entries are written). */
uint64_t nbh_fleet_join_lists(const nbh_fleet* f, uint64_t n, uint64_t* nb_keys, uint64_t* sts_keys,
                              int64_t* sts_owner, uint64_t cap);

nbh_cullfleet* nbh_cullfleet_new(uint64_t seed, uint64_t n, int64_t now_sec, int32_t now_nsec, double mean_kernels);
void nbh_cullfleet_free(nbh_cullfleet* f);
char* nbh_cullfleet_json(const nbh_cullfleet* f, uint64_t i);  /* free with nbh_str_free */
int nbh_cullfleet_pack(const nbh_cullfleet* f, const uint64_t* idx, uint64_t i0, uint64_t m, nbh_cull** out);

int nbh_threads(void);  /* host threads the packers use */

#ifdef __cplusplus
}
#endif
#endif /* NBGRAFT_HOST_H */
