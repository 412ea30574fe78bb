// fleet.cpp — packed batches and the seeded synthetic fleets of SURVEY.md §8d.
//
// Desired objects always come from the restated builders (builders.cpp) applied
// to a synthetic Notebook; live objects are the desired ones mutated by a drift
// class, so every quirk of COMMON:107-219 is exercised at a known rate.
#include "fleet.hpp"

#include <cmath>
#include <cstdio>

#include "../hash.h"
#include "quantity.hpp"
#include "rfc3339.hpp"

namespace nbh {
namespace {

const GVal kZero;
const GVal& field(const GVal& s, const char* name) {
  const GVal* p = (s.kind == GVal::Struct || s.kind == GVal::Map) ? s.find(name) : nullptr;
  return p ? *p : kZero;
}
const GVal& path(const GVal& s, std::initializer_list<const char*> names) {
  const GVal* cur = &s;
  for (const char* n : names) cur = &field(*cur, n);
  return *cur;
}
GVal* mpath(GVal& s, std::initializer_list<const char*> names) {
  GVal* cur = &s;
  for (const char* n : names) {
    if (!cur || (cur->kind != GVal::Struct && cur->kind != GVal::Map)) return nullptr;
    cur = cur->find(n);
  }
  return cur;
}
template <class T>
void extend(std::vector<T>& a, const std::vector<T>& b) { a.insert(a.end(), b.begin(), b.end()); }

}  // namespace

// ------------------------------------------------------------------ batches
void sparsify_tile(const uint64_t* tile, std::vector<uint64_t>& mask, std::vector<uint64_t>& val, std::vector<uint32_t>& off) {
  uint64_t m[3] = {0, 0, 0};
  for (unsigned s = 0; s < NBG_POD_SLOTS; ++s)
    if (tile[s]) { m[s >> 6] |= 1ull << (s & 63); val.push_back(tile[s]); }
  if (val.size() & 1) val.push_back(0);   // runs are padded to an even count: 16-byte aligned starts
  mask.insert(mask.end(), m, m + 3);
  off.push_back((uint32_t)val.size());
}

void RestartBatch::add(const GVal& old_nb, const GVal& updated_nb, const GVal& mutated_nb, bool create) {
  const GVal* nbs[3] = {&old_nb, &updated_nb, &mutated_nb};
  for (int k = 0; k < 3; ++k) {
    uint64_t tile[NBG_POD_SLOTS];
    pack_podspec(semantic_canon(path(*nbs[k], {"spec", "template", "spec"})), tile);
    sparsify_tile(tile, mask[k], val[k], off[k]);
  }
  const GVal& ann = path(mutated_nb, {"metadata", "annotations"});   // metav1.HasAnnotation(mutatedNotebook...), ODHWH:521,527
  uint8_t f = create ? NBG_RB_CREATE : 0;
  if (ann.kind == GVal::Map && ann.has("kubeflow-resource-stopped")) f |= NBG_RB_STOPPED;
  if (ann.kind == GVal::Map && ann.has("notebooks.opendatahub.io/notebook-restart")) f |= NBG_RB_RESTARTING;
  flags.push_back(f);
  ++n;
}
nbg_restart_batch RestartBatch::view() const {
  nbg_restart_batch b{};
  b.flags = flags.data();
  nbg_podsparse* dst[3] = {&b.old_spec, &b.updated_spec, &b.mutated_spec};
  for (int k = 0; k < 3; ++k) { dst[k]->mask = mask[k].data(); dst[k]->off = off[k].data(); dst[k]->val = val[k].data(); }
  return b;
}

void StsBatch::add(const GVal& from, const GVal& to) {
  auto replicas = [](const GVal& o) -> int32_t {
    const GVal& r = path(o, {"spec", "replicas"});
    return (r.kind == GVal::Ptr && !r.items.empty()) ? (int32_t)r.items[0].i : NBG_REPLICAS_NIL;
  };
  replicas_from.push_back(replicas(from));
  replicas_to.push_back(replicas(to));
  tmpl_from.push_back(digest_string_map(path(from, {"spec", "template", "metadata", "labels"})));
  tmpl_to.push_back(digest_string_map(path(to, {"spec", "template", "metadata", "labels"})));
  uint64_t tile_from[NBG_POD_SLOTS], tile_to[NBG_POD_SLOTS];
  pack_podspec(path(from, {"spec", "template", "spec"}), tile_from);
  pack_podspec(path(to, {"spec", "template", "spec"}), tile_to);
  if (!lean) {
    pod_from.insert(pod_from.end(), tile_from, tile_from + NBG_POD_SLOTS);
    pod_to.insert(pod_to.end(), tile_to, tile_to + NBG_POD_SLOTS);
  }
  labels_from.add(path(from, {"metadata", "labels"}));
  labels_to.add(path(to, {"metadata", "labels"}));
  ann_from.add(path(from, {"metadata", "annotations"}));
  ann_to.add(path(to, {"metadata", "annotations"}));
  sparsify_tile(tile_from, mask_from, val_from, voff_from);
  sparsify_tile(tile_to, mask_to, val_to, voff_to);
  ++n;
}
// The part of generateStatefulSet that is not the PodSpec (NBCTL:434-437,442,462-483), from the Notebook alone.
namespace {
struct NotebookSide {
  bool ok = false, stopped = false;
  const GVal* pod = nullptr;
  const GVal* env0 = nullptr;
  std::string prefix;
  GVal labels = GVal::map(), tmpl = GVal::map();
};
NotebookSide notebook_side(const GVal& nb) {
  NotebookSide r;
  const GVal& meta = field(nb, "metadata");
  const std::string name = meta.get_str("name"), ns = meta.get_str("namespace");
  r.pod = &path(nb, {"spec", "template", "spec"});
  const GVal& conts = field(*r.pod, "containers");
  if (name.size() > 52 || conts.kind != GVal::List || conts.items.empty()) return r;
  r.env0 = &field(conts.items[0], "env");
  const GVal& nb_ann = field(meta, "annotations");
  r.stopped = nb_ann.kind == GVal::Map && nb_ann.has("kubeflow-resource-stopped");   // NBCTL:434-437
  r.tmpl.set("statefulset", GVal::str(name));                                        // NBCTL:462-466,480-483
  r.tmpl.set("notebook-name", GVal::str(name));
  r.tmpl.set("opendatahub.io/workbenches", GVal::str("true"));
  const GVal& nb_labels = field(meta, "labels");                                     // NBCTL:442,474-477: a copy, non-nil even when empty
  if (nb_labels.kind == GVal::Map)
    for (const auto& e : nb_labels.kv) { r.tmpl.set(e.first, e.second); r.labels.set(e.first, e.second); }
  r.prefix = "/notebook/" + ns + "/" + name;
  r.ok = true;
  return r;
}
}  // namespace

bool StsBatch::add_gen(const GVal& nb, const GVal& to, const BuilderEnv& env) {
  if (lean) return false;                                  // the generator works on dense tiles
  const NotebookSide side = notebook_side(nb);
  if (!side.ok) return false;
  const bool too_long = side.env0->kind == GVal::List && side.env0->items.size() >= 14;   // the tile names 14 variables at most
  uint64_t tile_from[NBG_POD_SLOTS], tile_to[NBG_POD_SLOTS];
  if (too_long) {
    GVal ss;
    std::string err;
    if (!generate_statefulset(nb, env, &ss, &err)) return false;
    pack_podspec(path(ss, {"spec", "template", "spec"}), tile_from);
  } else {
    pack_podspec(*side.pod, tile_from);
  }
  pack_podspec(path(to, {"spec", "template", "spec"}), tile_to);
  const GVal& rt = path(to, {"spec", "replicas"});
  replicas_from.push_back(side.stopped ? 0 : 1);
  replicas_to.push_back((rt.kind == GVal::Ptr && !rt.items.empty()) ? (int32_t)rt.items[0].i : NBG_REPLICAS_NIL);
  tmpl_from.push_back(digest_string_map(side.tmpl));
  tmpl_to.push_back(digest_string_map(path(to, {"spec", "template", "metadata", "labels"})));
  pod_from.insert(pod_from.end(), tile_from, tile_from + NBG_POD_SLOTS);
  pod_to.insert(pod_to.end(), tile_to, tile_to + NBG_POD_SLOTS);
  labels_from.add(side.labels);
  labels_to.add(path(to, {"metadata", "labels"}));
  ann_from.add(GVal());                                    // the desired StatefulSet has no annotations (NBCTL:439-443)
  ann_to.add(path(to, {"metadata", "annotations"}));
  sparsify_tile(tile_from, mask_from, val_from, voff_from);   // kept consistent; the sparse form of a gen batch is the Notebook's tile too
  sparsify_tile(tile_to, mask_to, val_to, voff_to);
  prefix_hash.push_back(hash_str(side.prefix));
  gen_flags.push_back((uint8_t)((side.stopped ? NBG_GEN_STOPPED : 0) | (too_long ? NBG_GEN_PREPATCHED : 0)));
  ++n;
  return true;
}
bool StsBatch::add_notebook(const GVal& nb, const GVal& to, const BuilderEnv& env) {
  if (!prefix_hash.empty()) return false;                  // not into a batch add_gen has filled
  const NotebookSide side = notebook_side(nb);
  if (!side.ok) return false;
  uint64_t tile_from[NBG_POD_SLOTS], tile_to[NBG_POD_SLOTS];
  pack_podspec(*side.pod, tile_from);
  if (!patch_statefulset_tile(tile_from, hash_str(side.prefix), env.add_fsgroup)) {
    GVal ss;                                               // an env the tile cannot decide: build the object after all
    std::string err;
    if (!generate_statefulset(nb, env, &ss, &err)) return false;
    pack_podspec(path(ss, {"spec", "template", "spec"}), tile_from);
  }
  pack_podspec(path(to, {"spec", "template", "spec"}), tile_to);
  const GVal& rt = path(to, {"spec", "replicas"});
  replicas_from.push_back(side.stopped ? 0 : 1);
  replicas_to.push_back((rt.kind == GVal::Ptr && !rt.items.empty()) ? (int32_t)rt.items[0].i : NBG_REPLICAS_NIL);
  tmpl_from.push_back(digest_string_map(side.tmpl));
  tmpl_to.push_back(digest_string_map(path(to, {"spec", "template", "metadata", "labels"})));
  if (!lean) {
    pod_from.insert(pod_from.end(), tile_from, tile_from + NBG_POD_SLOTS);
    pod_to.insert(pod_to.end(), tile_to, tile_to + NBG_POD_SLOTS);
  }
  labels_from.add(side.labels);
  labels_to.add(path(to, {"metadata", "labels"}));
  ann_from.add(GVal());
  ann_to.add(path(to, {"metadata", "annotations"}));
  sparsify_tile(tile_from, mask_from, val_from, voff_from);
  sparsify_tile(tile_to, mask_to, val_to, voff_to);
  ++n;
  return true;
}
void StsBatch::append(const StsBatch& o) {
  {
    const uint32_t bf = (uint32_t)val_from.size(), bt = (uint32_t)val_to.size();
    for (size_t k = 1; k < o.voff_from.size(); ++k) voff_from.push_back(bf + o.voff_from[k]);
    for (size_t k = 1; k < o.voff_to.size(); ++k) voff_to.push_back(bt + o.voff_to[k]);
    extend(mask_from, o.mask_from); extend(mask_to, o.mask_to);
    extend(val_from, o.val_from); extend(val_to, o.val_to);
  }
  extend(replicas_from, o.replicas_from); extend(replicas_to, o.replicas_to);
  extend(tmpl_from, o.tmpl_from); extend(tmpl_to, o.tmpl_to);
  if (!lean) { extend(pod_from, o.pod_from); extend(pod_to, o.pod_to); }
  labels_from.append(o.labels_from); labels_to.append(o.labels_to);
  ann_from.append(o.ann_from); ann_to.append(o.ann_to);
  n += o.n;
}
nbg_sts_batch StsBatch::view() const {
  nbg_sts_batch b{};
  b.replicas_from = replicas_from.data(); b.replicas_to = replicas_to.data();
  b.tmpl_from = tmpl_from.data(); b.tmpl_to = tmpl_to.data();
  b.pod_from = lean ? nullptr : pod_from.data(); b.pod_to = lean ? nullptr : pod_to.data();
  b.labels_from = labels_from.view(); b.labels_to = labels_to.view();
  b.ann_from = ann_from.view(); b.ann_to = ann_to.view();
  return b;
}

nbg_sts_sparse_batch StsBatch::view_sparse() const {
  nbg_sts_sparse_batch b{};
  b.replicas_from = replicas_from.data(); b.replicas_to = replicas_to.data();
  b.tmpl_from = tmpl_from.data(); b.tmpl_to = tmpl_to.data();
  b.pod_from.mask = mask_from.data(); b.pod_from.off = voff_from.data(); b.pod_from.val = val_from.data();
  b.pod_to.mask = mask_to.data(); b.pod_to.off = voff_to.data(); b.pod_to.val = val_to.data();
  b.labels_from = labels_from.view(); b.labels_to = labels_to.view();
  b.ann_from = ann_from.view(); b.ann_to = ann_to.view();
  return b;
}

void SvcBatch::add(const GVal& from, const GVal& to) {
  const GVal &sf = path(from, {"spec", "selector"}), &st = path(to, {"spec", "selector"});
  flags.push_back((uint8_t)((sf.kind != GVal::Map ? NBG_F_FROM_SEL_NIL : 0) | (st.kind != GVal::Map ? NBG_F_TO_SEL_NIL : 0)));
  ports_from.resize(ports_from.size() + NBG_SVC_SLOTS);
  ports_to.resize(ports_to.size() + NBG_SVC_SLOTS);
  pack_service_ports(path(from, {"spec", "ports"}), &ports_from[n * NBG_SVC_SLOTS]);
  pack_service_ports(path(to, {"spec", "ports"}), &ports_to[n * NBG_SVC_SLOTS]);
  labels_from.add(path(from, {"metadata", "labels"})); labels_to.add(path(to, {"metadata", "labels"}));
  ann_from.add(path(from, {"metadata", "annotations"})); ann_to.add(path(to, {"metadata", "annotations"}));
  sel_from.add(sf); sel_to.add(st);
  // generateService's inputs, read back from the desired object it built
  name_hash.push_back(hash_str(path(from, {"metadata", "name"}).s));
  {
    const GVal& fp = path(from, {"spec", "ports"});
    int32_t tp = 8888;
    if (fp.kind == GVal::List && !fp.items.empty()) tp = (int32_t)path(fp.items[0], {"targetPort", "IntVal"}).i;
    target_port.push_back(tp);
  }
  ++n;
}
void SvcBatch::append(const SvcBatch& o) {
  extend(flags, o.flags); extend(ports_from, o.ports_from); extend(ports_to, o.ports_to);
  labels_from.append(o.labels_from); labels_to.append(o.labels_to);
  ann_from.append(o.ann_from); ann_to.append(o.ann_to);
  sel_from.append(o.sel_from); sel_to.append(o.sel_to);
  extend(name_hash, o.name_hash); extend(target_port, o.target_port);
  n += o.n;
}
nbg_svc_batch SvcBatch::view() const {
  nbg_svc_batch b{};
  b.flags = flags.data(); b.ports_from = ports_from.data(); b.ports_to = ports_to.data();
  b.labels_from = labels_from.view(); b.labels_to = labels_to.view();
  b.ann_from = ann_from.view(); b.ann_to = ann_to.view();
  b.sel_from = sel_from.view(); b.sel_to = sel_to.view();
  return b;
}

nbg_svc_gen_batch SvcBatch::view_gen() const {
  nbg_svc_gen_batch b{};
  b.name_hash = name_hash.data(); b.target_port = target_port.data();
  b.flags = flags.data(); b.ports_to = ports_to.data();
  b.labels_to = labels_to.view(); b.ann_to = ann_to.view(); b.sel_to = sel_to.view();
  return b;
}

void VsBatch::add(const GVal& from, const GVal& to) {
  spec_from.resize(spec_from.size() + NBG_VS_SLOTS);
  spec_to.resize(spec_to.size() + NBG_VS_SLOTS);
  bool fnil = true, tnil = true;
  std::vector<nbg_kv> hf, ht;
  const bool fok = pack_vs_spec(from, &spec_from[n * NBG_VS_SLOTS], &hf, &fnil);
  const bool tok = pack_vs_spec(to, &spec_to[n * NBG_VS_SLOTS], &ht, &tnil);
  flags.push_back((uint8_t)((fok ? 0 : NBG_F_FROM_SPEC_MISS) | (tok ? 0 : NBG_F_TO_SPEC_MISS) |
                            (fnil ? NBG_F_FROM_HDR_NIL : 0) | (tnil ? NBG_F_TO_HDR_NIL : 0)));
  extend(hdr_from.ent, hf); hdr_from.off.push_back((uint32_t)hdr_from.ent.size());
  extend(hdr_to.ent, ht); hdr_to.off.push_back((uint32_t)hdr_to.ent.size());
  prefix_hash.push_back(spec_from[n * NBG_VS_SLOTS + 2]);
  rewrite_hash.push_back(spec_from[n * NBG_VS_SLOTS + 3]);
  host_hash.push_back(spec_from[n * NBG_VS_SLOTS + 4]);
  ++n;
}
void VsBatch::append(const VsBatch& o) {
  extend(flags, o.flags); extend(spec_from, o.spec_from); extend(spec_to, o.spec_to);
  hdr_from.append(o.hdr_from); hdr_to.append(o.hdr_to);
  extend(prefix_hash, o.prefix_hash); extend(rewrite_hash, o.rewrite_hash); extend(host_hash, o.host_hash);
  n += o.n;
}
nbg_vs_batch VsBatch::view() const {
  nbg_vs_batch b{};
  b.flags = flags.data(); b.spec_from = spec_from.data(); b.spec_to = spec_to.data();
  b.hdr_from = hdr_from.view(); b.hdr_to = hdr_to.view();
  return b;
}

nbg_vs_gen_batch VsBatch::view_gen() const {
  nbg_vs_gen_batch b{};
  b.prefix_hash = prefix_hash.data(); b.rewrite_hash = rewrite_hash.data(); b.host_hash = host_hash.data();
  b.hdr_from = hdr_from.view(); b.flags = flags.data(); b.spec_to = spec_to.data(); b.hdr_to = hdr_to.view();
  return b;
}

void GenBatch::add(const GVal& la, const GVal& lb, const GVal& sa, const GVal& sb, bool flatten) {
  auto put = [&](KvCsr& c, const GVal& v) {
    if (flatten) flatten_tree(v, &c.ent); else pack_kvmap(v, &c.ent);
    c.off.push_back((uint32_t)c.ent.size());
  };
  put(labels_a, la); put(labels_b, lb); put(spec_a, sa); put(spec_b, sb);
  ++n;
}
nbg_generic_batch GenBatch::view() const {
  nbg_generic_batch b{};
  b.labels_a = labels_a.view(); b.labels_b = labels_b.view();
  b.spec_a = spec_a.view(); b.spec_b = spec_b.view();
  return b;
}

void CullBatch::add(const CullInput& in) {
  auto push = [](const std::vector<CullInput::Rec>& recs, bool kernels, std::vector<int64_t>& sec,
                 std::vector<uint32_t>& nsf, std::vector<int32_t>& zone, std::vector<uint32_t>& off) {
    for (const auto& r : recs) {
      Instant t;
      uint32_t w = 0;
      if (parse_rfc3339(r.last_activity, &t)) w = (uint32_t)t.nsec | NBG_REC_PARSE_OK; else t = Instant();
      if (kernels && r.state == "idle") w |= NBG_REC_IDLE;   // KERNEL_EXECUTION_STATE_IDLE, CULL:56,320
      sec.push_back(t.sec); nsf.push_back(w); zone.push_back(t.offset);
    }
    off.push_back((uint32_t)sec.size());
  };
  push(in.kernels, true, k_sec, k_nsf, k_zone, k_off);
  push(in.terminals, false, t_sec, t_nsf, t_zone, t_off);
  Instant a, c;
  uint32_t aw = 0, cw = 0;
  const bool have = in.has_annotations;
  if (have && in.la_present && parse_rfc3339(in.last_activity, &a)) aw = (uint32_t)a.nsec | NBG_ANN_PARSE_OK; else a = Instant();
  if (have && in.stop_set) aw |= NBG_ANN_STOP_SET;
  if (have && in.chk_present) {
    cw |= NBG_CHK_PRESENT;
    if (parse_rfc3339(in.check_ts, &c)) cw |= (uint32_t)c.nsec | NBG_CHK_PARSE_OK; else c = Instant();
  }
  ann_sec.push_back(a.sec); ann_nsf.push_back(aw);
  chk_sec.push_back(c.sec); chk_nsf.push_back(cw);
  nb_flags.push_back((uint8_t)((have && in.la_present ? NBG_NB_LA_PRESENT : 0) | (in.pod_missing ? NBG_NB_POD_MISSING : 0)));
  ++n;
}
void CullBatch::append(const CullBatch& o) {
  const uint32_t kb = (uint32_t)k_sec.size(), tb = (uint32_t)t_sec.size();
  for (size_t k = 1; k < o.k_off.size(); ++k) k_off.push_back(kb + o.k_off[k]);
  for (size_t k = 1; k < o.t_off.size(); ++k) t_off.push_back(tb + o.t_off[k]);
  extend(k_sec, o.k_sec); extend(k_nsf, o.k_nsf); extend(k_zone, o.k_zone);
  extend(t_sec, o.t_sec); extend(t_nsf, o.t_nsf); extend(t_zone, o.t_zone);
  extend(ann_sec, o.ann_sec); extend(ann_nsf, o.ann_nsf); extend(chk_sec, o.chk_sec); extend(chk_nsf, o.chk_nsf);
  extend(nb_flags, o.nb_flags);
  n += o.n;
}
nbg_cull_batch CullBatch::view() const {
  nbg_cull_batch b{};
  b.k_off = k_off.data(); b.k_sec = k_sec.data(); b.k_nsf = k_nsf.data();
  b.t_off = t_off.data(); b.t_sec = t_sec.data(); b.t_nsf = t_nsf.data();
  b.ann_sec = ann_sec.data(); b.ann_nsf = ann_nsf.data();
  b.chk_sec = chk_sec.data(); b.chk_nsf = chk_nsf.data();
  b.nb_flags = nb_flags.data();
  return b;
}

// ------------------------------------------------------------------- fleets
void Fleet::init() {
  if (n_namespaces == 0) n_namespaces = 1;
  ns_cdf.resize(n_namespaces);
  double acc = 0;
  for (uint32_t k = 0; k < n_namespaces; ++k) { acc += 1.0 / std::pow((double)(k + 1), 1.1); ns_cdf[k] = acc; }
  for (auto& v : ns_cdf) v /= acc;
}
uint32_t Fleet::namespace_of(uint64_t i) const {
  Rng r(seed ^ nbg_mix64(i * 2 + 1));
  const double u = r.unit();
  const auto it = std::lower_bound(ns_cdf.begin(), ns_cdf.end(), u);
  const size_t k = it - ns_cdf.begin();
  return (uint32_t)(k < ns_cdf.size() ? k : ns_cdf.size() - 1);
}
std::string Fleet::namespace_name(uint64_t i) const {
  char buf[32];
  snprintf(buf, sizeof buf, "ns-%06u", namespace_of(i));
  return buf;
}

namespace {

const char* kImages[] = {
    "quay.io/opendatahub/workbench-images:jupyter-minimal-ubi9-python-3.11",
    "quay.io/opendatahub/workbench-images:jupyter-datascience-ubi9-python-3.11",
    "quay.io/opendatahub/workbench-images:jupyter-pytorch-ubi9-python-3.11",
    "quay.io/opendatahub/workbench-images:jupyter-tensorflow-ubi9-python-3.11",
    "gcr.io/kubeflow-images-public/tensorflow-1.14.0-notebook-cpu:v1.0.0",
    "quay.io/thoth-station/s2i-minimal-notebook:v0.2.2",
    "kubeflownotebookswg/jupyter-scipy:v1.9.0",
    "kubeflownotebookswg/codeserver-python:v1.9.0"};
const char* kCpu[] = {"500m", "1", "2", "0.5", "250m", "4"};
const char* kMem[] = {"1Gi", "2Gi", "512Mi", "1024Mi", "4Gi", "8Gi"};
const char* kLabelKeys[] = {"app", "team", "env", "opendatahub.io/dashboard", "opendatahub.io/odh-managed",
                            "owner", "cost-center", "tier"};
const char* kTeams[] = {"ml-platform", "fraud", "search", "ads", "risk", "vision", "nlp", "infra"};

GVal quantity(const char* text) { GVal q; parse_quantity(text, &q); return q; }

GVal make_notebook(const Fleet& f, uint64_t i, Rng& r) {
  char name[32];
  snprintf(name, sizeof name, "nb-%07llu", (unsigned long long)i);
  const std::string ns = f.namespace_name(i);
  GVal meta = GVal::strct();
  meta.set("name", GVal::str(name));
  meta.set("namespace", GVal::str(ns));
  {
    char uid[40];
    snprintf(uid, sizeof uid, "%016llx-nb", (unsigned long long)nbg_mix64(f.seed ^ (i * 0x9E3779B97F4A7C15ull)));
    meta.set("uid", GVal::str(uid));
  }
  GVal labels = GVal::map();
  const int nl = 3 + (int)r.below(4);
  for (int k = 0; k < nl; ++k) {
    const char* key = kLabelKeys[(r.below(8) + k) % 8];
    std::string val;
    switch (r.below(4)) {
      case 0: val = "true"; break;
      case 1: val = kTeams[r.below(8)]; break;
      case 2: val = "user-" + std::to_string(r.below(50000)); break;
      default: val = "v" + std::to_string(r.below(12)); break;
    }
    labels.set(key, GVal::str(val));
  }
  meta.set("labels", std::move(labels));
  GVal ann = GVal::map();
  if (r.chance(0.6)) ann.set("openshift.io/display-name", GVal::str(std::string("Workbench ") + name));
  if (r.chance(0.5)) ann.set("notebooks.opendatahub.io/last-image-selection", GVal::str("jupyter-datascience:2024.1"));
  if (r.chance(0.3)) ann.set("kubectl.kubernetes.io/last-applied-configuration", GVal::str("{\"kind\":\"Notebook\"}"));
  if (r.chance(0.7)) ann.set("notebooks.kubeflow.org/last-activity", GVal::str("2025-12-31T23:10:00Z"));
  if (r.chance(0.1)) ann.set("opendatahub.io/username", GVal::str("user-" + std::to_string(r.below(50000))));
  if (!ann.kv.empty() || r.chance(0.5)) meta.set("annotations", std::move(ann));

  GVal c = GVal::strct();
  c.set("name", GVal::str(name));
  c.set("image", GVal::str(kImages[r.below(8)]));
  GVal req = GVal::map();
  req.set("cpu", quantity(kCpu[r.below(6)]));
  req.set("memory", quantity(kMem[r.below(6)]));
  GVal res = GVal::strct();
  res.set("requests", std::move(req));
  if (r.chance(0.5)) {
    GVal lim = GVal::map();
    lim.set("cpu", quantity(kCpu[r.below(6)]));
    lim.set("memory", quantity(kMem[r.below(6)]));
    if (r.chance(0.1)) lim.set("nvidia.com/gpu", quantity("1"));
    res.set("limits", std::move(lim));
  }
  c.set("resources", std::move(res));
  auto mount = [](const char* n, const char* p) {
    GVal m = GVal::strct();
    m.set("name", GVal::str(n));
    m.set("mountPath", GVal::str(p));
    return m;
  };
  GVal mounts = GVal::list();
  mounts.items.push_back(mount("workspace", "/home/jovyan"));
  mounts.items.push_back(mount("dshm", "/dev/shm"));
  c.set("volumeMounts", std::move(mounts));
  const int ne = (int)r.below(5);
  if (ne) {
    GVal env = GVal::list();
    for (int k = 0; k < ne; ++k) {
      GVal e = GVal::strct();
      e.set("name", GVal::str("VAR_" + std::to_string(k)));
      e.set("value", GVal::str("value-" + std::to_string(r.below(1000))));
      env.items.push_back(std::move(e));
    }
    c.set("env", std::move(env));
  }
  if (r.chance(0.15)) c.set("workingDir", GVal::str("/opt/app-root/src"));
  if (r.chance(0.2)) c.set("imagePullPolicy", GVal::str("Always"));

  GVal pvc = GVal::strct();
  pvc.set("claimName", GVal::str(std::string("workspace-") + name));
  GVal v0 = GVal::strct();
  v0.set("name", GVal::str("workspace"));
  v0.set("persistentVolumeClaim", GVal::ptr(std::move(pvc)));
  GVal ed = GVal::strct();
  ed.set("medium", GVal::str("Memory"));
  GVal v1 = GVal::strct();
  v1.set("name", GVal::str("dshm"));
  v1.set("emptyDir", GVal::ptr(std::move(ed)));
  GVal vols = GVal::list();
  vols.items.push_back(std::move(v0));
  vols.items.push_back(std::move(v1));

  GVal pod = GVal::strct();
  pod.set("serviceAccountName", GVal::str("default-editor"));
  pod.set("volumes", std::move(vols));
  GVal conts = GVal::list();
  conts.items.push_back(std::move(c));
  pod.set("containers", std::move(conts));
  if (r.chance(0.1)) {
    GVal t = GVal::strct();
    t.set("key", GVal::str("nvidia.com/gpu"));
    t.set("operator", GVal::str("Exists"));
    t.set("effect", GVal::str("NoSchedule"));
    GVal tl = GVal::list();
    tl.items.push_back(std::move(t));
    pod.set("tolerations", std::move(tl));
  }
  GVal tmpl = GVal::strct();
  tmpl.set("spec", std::move(pod));
  GVal spec = GVal::strct();
  spec.set("template", std::move(tmpl));
  GVal nb = GVal::strct();
  nb.set("metadata", std::move(meta));
  nb.set("spec", std::move(spec));
  return nb;
}

// What the API server's defaulting adds to a stored StatefulSet's PodSpec.
void apply_api_defaults(GVal* pod) {
  pod->set("restartPolicy", GVal::str("Always"));
  pod->set("terminationGracePeriodSeconds", GVal::ptr(GVal::integer(30)));
  pod->set("dnsPolicy", GVal::str("ClusterFirst"));
  pod->set("schedulerName", GVal::str("default-scheduler"));
  pod->set("serviceAccount", GVal::str(pod->get_str("serviceAccountName")));
  if (GVal* cs = pod->find("containers"))
    for (auto& c : cs->items) {
      c.set("terminationMessagePath", GVal::str("/dev/termination-log"));
      c.set("terminationMessagePolicy", GVal::str("File"));
      if (c.get_str("imagePullPolicy").empty()) c.set("imagePullPolicy", GVal::str("IfNotPresent"));
    }
}

}  // namespace

void fleet_make(const Fleet& f, uint64_t i, Pair* out) {
  Rng r(f.seed ^ nbg_mix64(i * 2));
  out->notebook = make_notebook(f, i, r);
  // a second stream for the Feast / cluster-proxy decisions, so that adding them left every other draw as it was
  Rng r2(f.seed ^ nbg_mix64(i * 2 + 1) ^ 0x6665617374ull);
  bool feast_on = false, proxy_on = false;
  if (f.kind == FLEET_ODH) {
    OdhInject inj;
    inj.lock = r.chance(0.2);
    inj.elyra_secret = r.chance(0.7);
    inj.runtime_images = r.chance(0.8);
    // Feast (ODHWH:431-444): 12 % of the notebooks carry the label and get the 6th volume; 3 % had it once —
    // the label now says "false" and the user's spec still holds the volume and mount, which the webhook removes
    const double uf = r2.unit();
    GVal* labels = mpath(out->notebook, {"metadata", "labels"});
    if (uf < 0.12) {
      feast_on = true;
      labels->set("opendatahub.io/feast-integration", GVal::str("true"));
    } else if (uf < 0.15) {
      labels->set("opendatahub.io/feast-integration", GVal::str("false"));
      OdhInject once;
      once.ca_bundle = once.runtime_images = once.elyra_secret = once.rbac_proxy = false;
      GVal was = out->notebook;
      mpath(was, {"metadata", "labels"})->set("opendatahub.io/feast-integration", GVal::str("true"));
      odh_webhook_mutate(&was, once);
      *mpath(out->notebook, {"spec", "template", "spec"}) = *mpath(was, {"spec", "template", "spec"});
    }
    // cluster-wide proxy (ODHWH:457-473): a quarter of the notebooks were admitted while it was switched on;
    // one in ten of those already had an HTTP_PROXY of their own, which the injector leaves alone (:583-585)
    if (r2.chance(0.25)) {
      proxy_on = true;
      inj.proxy_env = true;
      inj.http_proxy = "http://proxy.corp.example:3128";
      inj.https_proxy = "http://proxy.corp.example:3129";
      inj.no_proxy = ".cluster.local,.svc,10.0.0.0/8,localhost";
      if (r2.chance(0.1)) {
        GVal* cs = mpath(out->notebook, {"spec", "template", "spec", "containers"});
        if (cs && cs->kind == GVal::List && !cs->items.empty()) {
          GVal e = GVal::strct();
          e.set("name", GVal::str("HTTP_PROXY"));
          e.set("value", GVal::str("http://my-own-proxy:8080"));
          GVal* env = cs->items[0].find("env");
          if (!env || env->kind != GVal::List) { cs->items[0].set("env", GVal::list()); env = cs->items[0].find("env"); }
          env->items.push_back(std::move(e));
        }
      }
    }
    odh_webhook_mutate(&out->notebook, inj);
    GVal* ann = mpath(out->notebook, {"metadata", "annotations"});
    if (ann && ann->kind == GVal::Map) {
      const int nh = (int)r.below(4);
      if (nh) {
        std::string js = "{";
        for (int k = 0; k < nh; ++k)
          js += std::string(k ? "," : "") + "\"X-Hdr-" + std::to_string(k) + "\":\"v" + std::to_string(r.below(100)) + "\"";
        ann->set("notebooks.kubeflow.org/http-headers-request-set", GVal::str(js + "}"));
      }
      if (r.chance(0.1)) ann->set("notebooks.kubeflow.org/http-rewrite-uri", GVal::str("/custom/"));
    }
  }
  // a stopped notebook (culled, or carrying the ODH lock) is desired at 0 replicas
  const bool stopped = r.chance(0.05);
  if (stopped) {
    GVal* meta = out->notebook.find("metadata");
    if (!meta->find("annotations")) meta->set("annotations", GVal::map());
    meta->find("annotations")->set("kubeflow-resource-stopped", GVal::str("2025-12-31T22:00:00Z"));
  }
  std::string err;
  generate_statefulset(out->notebook, f.env, &out->sts_from, &err);
  generate_service(out->notebook, &out->svc_from, &err);
  generate_virtualservice(out->notebook, f.env, &out->vs_from, &err);
  {  // ctrl.SetControllerReference(instance, ss, r.Scheme) — NBCTL:153,208
    const GVal* nbm = out->notebook.find("metadata");
    GVal ref = GVal::strct();
    ref.set("apiVersion", GVal::str("kubeflow.org/v1beta1"));
    ref.set("kind", GVal::str("Notebook"));
    ref.set("name", GVal::str(nbm->get_str("name")));
    ref.set("uid", GVal::str(nbm->get_str("uid")));
    ref.set("controller", GVal::ptr(GVal::boolean(true)));
    ref.set("blockOwnerDeletion", GVal::ptr(GVal::boolean(true)));
    GVal refs = GVal::list();
    refs.items.push_back(std::move(ref));
    out->sts_from.find("metadata")->set("ownerReferences", refs);
    out->svc_from.find("metadata")->set("ownerReferences", refs);
  }
  out->sts_to = out->sts_from;
  out->svc_to = out->svc_from;
  out->vs_to = out->vs_from;

  // ---- StatefulSet drift classes
  const double u = r.unit();
  GVal* to_meta = out->sts_to.find("metadata");
  GVal* to_labels = to_meta->find("labels");
  GVal* to_pod = mpath(out->sts_to, {"spec", "template", "spec"});
  GVal* to_c0 = &to_pod->find("containers")->items[0];
  int cls = 0;
  if (u < 0.40) cls = 0;                                           // identical
  else if (u < 0.60) { cls = 1; apply_api_defaults(to_pod); }      // API-defaulted PodSpec => P
  else if (u < 0.70) {                                             // live label changed / extra => L
    cls = 2;
    if (to_labels && !to_labels->kv.empty() && r.chance(0.5)) to_labels->kv[r.below(to_labels->kv.size())].second = GVal::str("changed");
    else if (to_labels) to_labels->set("controller-revision-hash", GVal::str("nb-" + std::to_string(r.below(1000))));
  } else if (u < 0.80) {                                           // desired-only label => not detected
    cls = 3;
    if (to_labels && !to_labels->kv.empty()) to_labels->kv.erase(to_labels->kv.begin() + r.below(to_labels->kv.size()));
  } else if (u < 0.85) {                                           // live (k,"") absent from desired => equal
    cls = 4;
    if (to_labels) to_labels->set("empty-valued", GVal::str(""));
  } else if (u < 0.90) {                                           // replicas 0 <-> 1 => R (+ D1')
    cls = 5;
    GVal* rep = mpath(out->sts_to, {"spec", "replicas"});
    rep->items[0].i = rep->items[0].i ? 0 : 1;
    if (r.chance(0.5)) mpath(out->sts_to, {"spec", "template", "metadata", "labels"})->set("pod-template-hash", GVal::str("x"));
  } else if (u < 0.95) {                                           // live annotation => A
    cls = 6;
    if (!to_meta->find("annotations")) to_meta->set("annotations", GVal::map());
    to_meta->find("annotations")->set("deployment.kubernetes.io/revision", GVal::str("3"));
  } else if (u < 0.98) {                                           // nil <-> empty list => P
    cls = 7;
    if (to_c0->find("command")) to_c0->erase("command"); else to_c0->set("command", GVal::list());
  } else if (u < 0.99) {                                           // Quantity: same number, other scale => P
    cls = 8;
    GVal* req = mpath(*to_c0, {"resources", "requests"});
    const GVal cur = req->get("cpu");
    req->set("cpu", digest(cur) == digest(quantity("500m")) ? quantity("0.5") : quantity("500m"));
    if (digest(cur) == digest(req->get("cpu"))) req->set("cpu", quantity("0.5"));
  } else {                                                         // Quantity: only the cached string differs => P
    cls = 9;
    GVal* req = mpath(*to_c0, {"resources", "requests"});
    // "1Gi" keeps s="1Gi"; "1024Mi" parses to the same (value, scale, format) with s="".
    // (the one class that edits the desired side after the builder ran: the
    // Notebook would have asked for 1024Mi)
    req->set("memory", quantity("1Gi"));
    mpath(out->sts_from, {"spec", "template", "spec"})->find("containers")->items[0].find("resources")->find("requests")->set("memory", quantity("1024Mi"));
  }
  // ODH only, on top of the class above: StatefulSets created before the webhook learnt a mutation.  Both make
  // the PodSpecs differ (P), so they are applied only where the class already expects P or nothing else is checked.
  if (f.kind == FLEET_ODH && (cls == 1 || cls == 7)) {
    if (feast_on && r2.chance(0.5)) {                                // live pod still without the Feast volume and mount
      GVal* vols = to_pod->find("volumes");
      if (vols && vols->kind == GVal::List)
        for (size_t k = 0; k < vols->items.size(); ++k)
          if (vols->items[k].get_str("name") == "odh-feast-config") { vols->items.erase(vols->items.begin() + k); break; }
      GVal* ms = to_c0->find("volumeMounts");
      if (ms && ms->kind == GVal::List)
        for (size_t k = 0; k < ms->items.size(); ++k)
          if (ms->items[k].get_str("name") == "odh-feast-config") { ms->items.erase(ms->items.begin() + k); break; }
    }
    if (proxy_on && r2.chance(0.5)) {                                // live pod with the proxy of an earlier cluster config
      GVal* env = to_c0->find("env");
      if (env && env->kind == GVal::List)
        for (auto& e : env->items)
          if (e.get_str("name") == "NO_PROXY") e.set("value", GVal::str(".cluster.local,.svc,localhost"));
    }
  }
  out->drift_class = cls;

  // ---- Service drift
  const double us = r.unit();
  GVal* svc_spec = out->svc_to.find("spec");
  svc_spec->set("clusterIP", GVal::str("10.0." + std::to_string(r.below(255)) + "." + std::to_string(r.below(255))));  // never compared (COMMON:182)
  if (us < 0.80) {
  } else if (us < 0.84) svc_spec->find("selector")->set("statefulset", GVal::str("other"));
  else if (us < 0.87) svc_spec->find("ports")->items[0].set("targetPort", [] { GVal g = GVal::strct(); g.set("Type", GVal::integer(1)); g.set("StrVal", GVal::str("8888")); return g; }());
  else if (us < 0.90) svc_spec->find("ports")->items[0].set("nodePort", GVal::integer(30000 + (int64_t)r.below(2000)));
  else if (us < 0.94) { GVal l = GVal::map(); l.set("app.kubernetes.io/managed-by", GVal::str("helm")); out->svc_to.find("metadata")->set("labels", std::move(l)); }
  else if (us < 0.97) { GVal a = GVal::map(); a.set("service.beta.openshift.io/serving-cert-secret-name", GVal::str("tls")); out->svc_to.find("metadata")->set("annotations", std::move(a)); }
  else if (us < 0.98) svc_spec->erase("selector");                 // nil selector
  else if (us < 0.99) { GVal l = GVal::map(); l.set("blank", GVal::str("")); out->svc_to.find("metadata")->set("labels", std::move(l)); }  // equal
  else svc_spec->find("ports")->items.push_back(svc_spec->find("ports")->items[0]);

  // ---- VirtualService drift
  const double uv = r.unit();
  GVal* vspec = out->vs_to.find("spec");
  GVal* http0 = &vspec->find("http")->items[0];
  if (uv < 0.85) {
  } else if (uv < 0.90) {                                          // int64(80) vs float64(80)
    *mpath(http0->find("route")->items[0], {"destination", "port", "number"}) = GVal::real(80.0);
  } else if (uv < 0.92) out->vs_to.erase("spec");                  // to.spec missing => copy, true
  else if (uv < 0.95) mpath(*http0, {"headers", "request", "set"})->set("X-Extra", GVal::str("1"));
  else if (uv < 0.97) vspec->find("hosts")->items[0] = GVal::str("notebooks.example.com");
  else if (uv < 0.98) vspec->set("exportTo", [] { GVal l = GVal::list(); l.items.push_back(GVal::str(".")); return l; }());
  else if (uv < 0.99) http0->set("timeout", GVal::str("300s"));
  else mpath(*http0, {"rewrite"})->set("uri", GVal::str("/"));
}

// ------------------------------------------------------------- cull fleets
namespace {
std::string fmt_time(int64_t sec, int32_t nsec, int32_t offset, int frac_digits) {
  std::string s = format_rfc3339(sec, offset);
  if (frac_digits > 0) {
    char buf[16];
    snprintf(buf, sizeof buf, "%09d", nsec);
    const size_t tz = offset == 0 ? s.size() - 1 : s.size() - 6;
    s.insert(tz, "." + std::string(buf, frac_digits));
  }
  return s;
}
}  // namespace

void cullfleet_make(const CullFleet& f, uint64_t i, CullInput* out) {
  Rng r(f.seed ^ nbg_mix64(i * 2 + 0x63756C6Cull));
  *out = CullInput();
  const int64_t three_days = 3 * 86400;
  auto geometric = [&](double mean, int cap) {
    if (mean <= 0) return 0;
    const double p = 1.0 / (mean + 1.0);
    int k = (int)std::floor(std::log(1.0 - r.unit()) / std::log(1.0 - p));
    return k > cap ? cap : k;
  };
  auto activity = [&]() {
    const int64_t sec = f.now_sec - (int64_t)r.below(three_days);
    const int32_t nsec = (int32_t)r.below(1000000000);
    const double u = r.unit();
    if (u < 0.005) {  // malformed
      static const char* bad[] = {"", "should-fail", "2025-13-01T00:00:00Z", "2025-12-31 23:00:00Z", "2025-12-31T23:00:00",
                                  "2025-02-30T10:00:00Z", "2025-12-31t23:00:00Z", "2025-12-31T23:00:60Z"};
      return std::string(bad[r.below(8)]);
    }
    if (u < 0.10) return fmt_time(sec, 0, 0, 0);                                   // whole seconds
    if (u < 0.15) return fmt_time(sec, nsec, (int32_t)(r.below(27) - 13) * 3600, 6);  // zoned
    return fmt_time(sec, nsec, 0, 6);                                              // Jupyter: microseconds, Z
  };
  int nk = r.chance(0.10) ? 0 : 1 + geometric(f.mean_kernels / 0.9 - 1.0, 31);  // overall mean = mean_kernels
  if (nk > 32) nk = 32;
  for (int k = 0; k < nk; ++k) {
    CullInput::Rec rec;
    rec.last_activity = activity();
    const double u = r.unit();
    rec.state = u < 0.85 ? "idle" : u < 0.95 ? "busy" : u < 0.99 ? "starting" : "dead";
    out->kernels.push_back(std::move(rec));
  }
  const int nt = geometric(0.5, 8);
  for (int k = 0; k < nt; ++k) {
    CullInput::Rec rec;
    rec.last_activity = activity();
    out->terminals.push_back(std::move(rec));
  }
  out->has_annotations = !r.chance(0.002);
  out->la_present = !r.chance(0.01);
  const int64_t asec = f.now_sec - (int64_t)r.below(three_days);
  out->last_activity = r.chance(0.01) ? fmt_time(asec, (int32_t)r.below(1000000000), 0, 6)
                       : r.chance(0.005) ? std::string("should-fail") : fmt_time(asec, 0, 0, 0);
  out->stop_set = r.chance(0.01);
  out->chk_present = !r.chance(0.005);
  out->check_ts = r.chance(0.005) ? std::string("not-a-time") : fmt_time(f.now_sec - (int64_t)r.below(121), 0, 0, 0);
  out->pod_missing = r.chance(0.003);
}

static void json_escape(const std::string& s, std::string* o) {
  o->push_back('"');
  for (char c : s) {
    if (c == '"' || c == '\\') { o->push_back('\\'); o->push_back(c); }
    else if ((unsigned char)c < 0x20) { char b[8]; snprintf(b, sizeof b, "\\u%04x", c); *o += b; }
    else o->push_back(c);
  }
  o->push_back('"');
}

std::string cull_input_json(const CullInput& in) {
  std::string o = "{\"annotations\":";
  if (!in.has_annotations) o += "null";
  else {
    o += "{";
    bool first = true;
    auto kv = [&](const char* k, const std::string& v) {
      if (!first) o += ",";
      first = false;
      json_escape(k, &o); o += ":"; json_escape(v, &o);
    };
    if (in.stop_set) kv("kubeflow-resource-stopped", "2025-12-31T22:00:00Z");
    if (in.la_present) kv("notebooks.kubeflow.org/last-activity", in.last_activity);
    if (in.chk_present) kv("notebooks.kubeflow.org/last_activity_check_timestamp", in.check_ts);
    o += "}";
  }
  o += ",\"pod_missing\":";
  o += in.pod_missing ? "true" : "false";
  o += ",\"kernels\":[";
  for (size_t k = 0; k < in.kernels.size(); ++k) {
    if (k) o += ",";
    o += "{\"last_activity\":"; json_escape(in.kernels[k].last_activity, &o);
    o += ",\"execution_state\":"; json_escape(in.kernels[k].state, &o); o += "}";
  }
  o += "],\"terminals\":[";
  for (size_t k = 0; k < in.terminals.size(); ++k) {
    if (k) o += ",";
    o += "{\"last_activity\":"; json_escape(in.terminals[k].last_activity, &o); o += "}";
  }
  o += "]}";
  return o;
}

}  // namespace nbh
