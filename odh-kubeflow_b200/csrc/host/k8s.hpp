// k8s.hpp — typed decoding of Kubernetes JSON into GVal, the desired-object
// builders of the notebook controller, and the packer that flattens objects
// into the structure-of-arrays schema of include/nbgraft.h.
#pragma once
#include <cstdint>
#include <string>
#include <vector>

#include "../../../include/nbgraft.h"
#include "gval.hpp"

namespace nbh {

// ---- typed decoding (k8s.cpp) ---------------------------------------------
// Convert the generic JSON tree `j` (json.hpp) into the GVal of Go type `type`:
// "Notebook", "StatefulSet", "Service", "PodSpec", "ObjectMeta", ... as
// encoding/json would fill k8s.io/api v0.33.7 structs.  Fields this table does
// not know are kept as untyped JSON under "?<name>" so that they still take
// part in comparisons (never silently dropped).
bool decode_typed(const GVal& j, const std::string& type, GVal* out, std::string* err);
bool decode_typed_json(const std::string& json, const std::string& type, GVal* out, std::string* err);

// ---- builders (builders.cpp) ----------------------------------------------
struct BuilderEnv {            // process environment read by the reference builders
  bool add_fsgroup = true;     // ADD_FSGROUP unset or "true"  (NBCTL:514)
  std::string cluster_domain = "cluster.local";  // CLUSTER_DOMAIN (NBCTL:576)
  std::string istio_host = "*";                  // ISTIO_HOST (NBCTL:587-590)
  std::string istio_gateway = "kubeflow/kubeflow-gateway";  // ISTIO_GATEWAY (NBCTL:596-599)
};
// generateStatefulSet (NBCTL:433-523).  false + err where the Go code panics.
bool generate_statefulset(const GVal& notebook, const BuilderEnv& env, GVal* out, std::string* err);
// generateService (NBCTL:525-552)
bool generate_service(const GVal& notebook, GVal* out, std::string* err);
// generateVirtualService (NBCTL:558-658): an unstructured (generic) object
bool generate_virtualservice(const GVal& notebook, const BuilderEnv& env, GVal* out, std::string* err);
// ODH mutating-webhook injections that shape config #4 (ODHWH:177-326,666-781 and
// notebook_runtime.go:216-285, notebook_dspa_secret.go:375-450): applied to a Notebook.
struct OdhInject {
  bool lock = false, ca_bundle = true, runtime_images = true, elyra_secret = true, rbac_proxy = true;
  // the label-gated Feast block of Handle (ODHWH:431-444): label opendatahub.io/feast-integration == "true"
  // mounts <name>-feast-config (notebook_feast_config.go:53-117), any other value unmounts a mounted one (:120-147)
  bool feast = true;
  // InjectProxyConfigEnvVars (ODHWH:566-616): runs when INJECT_CLUSTER_PROXY_ENV parses as true and the cluster
  // Proxy has all three status fields; the map it ranges over holds exactly these three keys (ODHWH:340-342)
  bool proxy_env = false;
  std::string http_proxy, https_proxy, no_proxy;
};
// false when InjectKubeRbacProxy refuses the auth-sidecar resource annotations (the other injections
// are still applied; no proxy container, volumes or service account are added — ODHWH:177-183)
bool odh_webhook_mutate(GVal* notebook, const OdhInject& what);

// Desired objects of the ODH controller (the `from` side of its comparators): `what` is "HTTPRoute",
// "NetworkPolicy", "KubeRbacProxyNetworkPolicy", "ReferenceGrant" or "RoleBinding".
struct OdhBuildEnv {
  std::string central_namespace;                 // the controller's own namespace (routes live there)
  std::string gateway_name, gateway_namespace;   // NOTEBOOK_GATEWAY_NAME / _NAMESPACE; "" = the defaults
  std::string rolebinding_name, role_kind, role_name;
};
bool odh_build(const GVal& notebook, const std::string& what, const OdhBuildEnv& env, GVal* out);

// ---- the copy half of Copy*Fields (apply.cpp) ------------------------------
// `reason` = the bits the compare kernel returned for this pair (include/nbgraft.h).
void apply_statefulset_fields(const GVal& from, GVal* to, unsigned reason, bool caller_prestep);
void apply_service_fields(const GVal& from, GVal* to);
void apply_virtualservice(const GVal& from, GVal* to, unsigned reason);

// ---- packing (pack.cpp) ---------------------------------------------------
uint64_t hash_str(const std::string& s);
uint64_t digest(const GVal& v);                 // canonical 64-bit digest of a value tree
uint64_t digest_string_map(const GVal& m);      // nil != empty
void pack_kvmap(const GVal& m, std::vector<nbg_kv>* out);   // sorted by key hash; Null/absent => none
// Flatten a whole value tree into (hash(path), leaf-or-marker) entries sorted by key, such that
// two flattened vectors are equal  <=>  reflect.DeepEqual of the trees (NBG_GEN_DEEPEQUAL).
void flatten_tree(const GVal& v, std::vector<nbg_kv>* out);
void pack_podspec(const GVal& podspec, uint64_t* slots /*NBG_POD_SLOTS*/);
// the Notebook's PodSpec tile -> the tile of the StatefulSet generateStatefulSet makes from it (NBCTL:495-521)
bool patch_statefulset_tile(uint64_t* slots, uint64_t prefix_hash, bool add_fsgroup);
// The tree as apimachinery's equality.Semantic.DeepEqual sees it: nil and empty slices/maps are
// the same, a Quantity is its number.  pack_podspec(semantic_canon(x)) tiles are equal
// <=> equality.Semantic.DeepEqual(x, y)  (ODHWH:547-563).
GVal semantic_canon(const GVal& v);
void pack_service_ports(const GVal& ports, uint64_t* slots /*NBG_SVC_SLOTS*/);
// returns false when obj.spec is missing or not a map (NestedMap not found / error)
bool pack_vs_spec(const GVal& vs_obj, uint64_t* slots /*NBG_VS_SLOTS*/, std::vector<nbg_kv>* headers,
                  bool* headers_nil);

}  // namespace nbh
