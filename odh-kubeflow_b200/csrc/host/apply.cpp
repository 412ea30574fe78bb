// apply.cpp — the COPY half of the reference's Copy*Fields
// (components/common/reconcilehelper/util.go = COMMON).  The COMPARE half runs on the GPU and
// returns reason bits; what the Go functions then do to `to`, they do whatever the comparison
// said (the copies are unconditional, except Replicas), so it is plain host work on the two
// objects:
//   CopyStatefulSetFields  COMMON:114 to.Labels = from.Labels          :121 to.Annotations = from.Annotations
//                          COMMON:124 *to.Spec.Replicas = *from...     (only when they differ)
//                          COMMON:131 to.Spec.Template.Spec = from.Spec.Template.Spec
//                          (+ caller pre-step NBCTL:190-194: template labels, iff replicas differ and labels differ)
//   CopyServiceFields      COMMON:173,180 labels, annotations          :187 Selector   :192 Ports
//   CopyVirtualService     COMMON:200-206 from.spec missing -> nothing :208-217 SetNestedMap(to, fromSpec, "spec")
#include "../../../include/nbgraft.h"
#include "k8s.hpp"

namespace nbh {
namespace {

// struct field `name` of `s`, created as an empty struct when absent (set() would drop a zero struct)
GVal& child(GVal& s, const char* name) {
  if (!s.find(name)) {
    auto it = std::lower_bound(s.kv.begin(), s.kv.end(), std::string(name),
                               [](const std::pair<std::string, GVal>& e, const std::string& n) { return e.first < n; });
    s.kv.insert(it, std::make_pair(std::string(name), GVal::strct()));
  }
  return *s.find(name);
}
const GVal* walk(const GVal& s, std::initializer_list<const char*> names) {
  const GVal* cur = &s;
  for (const char* n : names) {
    cur = cur->find(n);
    if (!cur) return nullptr;
  }
  return cur;
}
// to.<name> = from.<name>  (a Go assignment: nil stays nil, the value is copied)
void assign(GVal& to_parent, const GVal* from_parent, const char* name) {
  const GVal* v = from_parent ? from_parent->find(name) : nullptr;
  if (v) to_parent.set(name, *v);
  else to_parent.erase(name);
}

}  // namespace

void apply_statefulset_fields(const GVal& from, GVal* to, unsigned reason, bool caller_prestep) {
  if (caller_prestep && (reason & NBG_STS_TMPL_LABELS))                                  // NBCTL:190-194
    assign(child(child(child(*to, "spec"), "template"), "metadata"), walk(from, {"spec", "template", "metadata"}), "labels");
  GVal& meta = child(*to, "metadata");
  assign(meta, from.find("metadata"), "labels");                                          // COMMON:114
  assign(meta, from.find("metadata"), "annotations");                                     // COMMON:121
  GVal& spec = child(*to, "spec");
  if (reason & NBG_STS_REPLICAS) assign(spec, from.find("spec"), "replicas");             // COMMON:123-126
  assign(child(spec, "template"), walk(from, {"spec", "template"}), "spec");              // COMMON:131
  to->canon();
}

void apply_service_fields(const GVal& from, GVal* to) {
  GVal& meta = child(*to, "metadata");
  assign(meta, from.find("metadata"), "labels");                                          // COMMON:173
  assign(meta, from.find("metadata"), "annotations");                                     // COMMON:180
  GVal& spec = child(*to, "spec");
  assign(spec, from.find("spec"), "selector");                                            // COMMON:187 (ClusterIP etc. untouched, :182)
  assign(spec, from.find("spec"), "ports");                                               // COMMON:192
  to->canon();
}

void apply_virtualservice(const GVal& from, GVal* to, unsigned reason) {
  if (reason & NBG_VS_FROM_MISSING) return;                                               // COMMON:200-206: returns false, no copy
  if (!(reason & (NBG_VS_SPEC_DIFFERS | NBG_VS_TO_MISSING))) return;                      // equal specs: nothing is written (:214)
  const GVal* spec = from.find("spec");
  if (spec) to->set("spec", *spec);                                                       // SetNestedMap deep-copies (:210,:216)
}

}  // namespace nbh
