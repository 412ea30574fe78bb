// fleet.hpp — seeded synthetic Notebook fleets and the packed
// batches built from them.  Object i of a fleet is a pure function of
// (kind, seed, i): fleets are never stored as objects, they are regenerated on
// demand, so a 10 M-notebook fleet costs no memory until it is packed.
#pragma once
#include <cstdint>
#include <string>
#include <vector>

#include "../../../include/nbgraft.h"
#include "gval.hpp"
#include "k8s.hpp"

namespace nbh {

struct Rng {  // splitmix64
  uint64_t s;
  explicit Rng(uint64_t seed) : s(seed) {}
  uint64_t next() {
    uint64_t z = (s += 0x9E3779B97F4A7C15ull);
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
  }
  uint64_t below(uint64_t n) { return n ? next() % n : 0; }
  double unit() { return (next() >> 11) * (1.0 / 9007199254740992.0); }
  bool chance(double p) { return unit() < p; }
};

// ---- packed batches (own their arrays) --------------------------------------
struct KvCsr {
  std::vector<uint32_t> off{0};
  std::vector<nbg_kv> ent;
  void add(const GVal& m) { pack_kvmap(m, &ent); off.push_back((uint32_t)ent.size()); }
  void append(const KvCsr& o) {
    const uint32_t base = (uint32_t)ent.size();
    ent.insert(ent.end(), o.ent.begin(), o.ent.end());
    for (size_t k = 1; k < o.off.size(); ++k) off.push_back(base + o.off[k]);
  }
  nbg_kvmap view() const { nbg_kvmap m{}; m.off = off.data(); m.ent = ent.data(); return m; }
};

struct StsBatch {
  std::vector<int32_t> replicas_from, replicas_to;
  std::vector<uint64_t> tmpl_from, tmpl_to, pod_from, pod_to;
  KvCsr labels_from, labels_to, ann_from, ann_to;
  // the same tiles stored sparsely (nbg_podsparse): presence mask + non-zero slots
  std::vector<uint64_t> mask_from, mask_to, val_from, val_to;
  std::vector<uint32_t> voff_from{0}, voff_to{0};
  uint64_t n = 0;
  // lean: keep only the sparse form of the PodSpec tiles (what the step reads).  A 10 M-notebook batch then
  // costs 0.59 KB per pair on the host instead of 3.7 KB; view() of a lean batch has no dense tiles.
  bool lean = false;
  // generateStatefulSet on the device (nbg_diff_sts_gen): per-notebook inputs of the generator; filled by add_gen only
  std::vector<uint64_t> prefix_hash;
  std::vector<uint8_t> gen_flags;
  void add(const GVal& from, const GVal& to);
  // the `from` side taken from the NOTEBOOK (its own PodSpec tile, its labels); false when generateStatefulSet would
  // fail or take its generateName branch (no containers, name longer than 52 characters)
  bool add_gen(const GVal& notebook, const GVal& to, const BuilderEnv& env);
  // the same pair as add(generate_statefulset(notebook), to), array for array, without building the StatefulSet:
  // the desired tile is the Notebook's tile patched on the host (patch_statefulset_tile).  Works on lean batches.
  bool add_notebook(const GVal& notebook, const GVal& to, const BuilderEnv& env);
  void append(const StsBatch& o);
  nbg_sts_batch view() const;
  nbg_sts_sparse_batch view_sparse() const;
};
struct SvcBatch {
  std::vector<uint8_t> flags;
  std::vector<uint64_t> ports_from, ports_to;
  KvCsr labels_from, labels_to, ann_from, ann_to, sel_from, sel_to;
  // the few Notebook fields generateService depends on (nbg_svc_gen_batch)
  std::vector<uint64_t> name_hash;
  std::vector<int32_t> target_port;
  uint64_t n = 0;
  void add(const GVal& from, const GVal& to);
  void append(const SvcBatch& o);
  nbg_svc_batch view() const;
  nbg_svc_gen_batch view_gen() const;
};
struct VsBatch {
  std::vector<uint8_t> flags;
  std::vector<uint64_t> spec_from, spec_to;
  KvCsr hdr_from, hdr_to;
  // the per-notebook words of generateVirtualService (nbg_vs_gen_batch)
  std::vector<uint64_t> prefix_hash, rewrite_hash, host_hash;
  uint64_t n = 0;
  void add(const GVal& from, const GVal& to);
  void append(const VsBatch& o);
  nbg_vs_batch view() const;
  nbg_vs_gen_batch view_gen() const;
};

// Pairs for the ODH comparator family (nbg_generic_batch).
struct GenBatch {
  KvCsr labels_a, labels_b, spec_a, spec_b;
  uint64_t n = 0;
  // flattened trees (NBG_GEN_DEEPEQUAL) or plain string maps (NBG_GEN_CONFIGMAP)
  void add(const GVal& la, const GVal& lb, const GVal& sa, const GVal& sb, bool flatten);
  nbg_generic_batch view() const;
};

// Append one dense 192-slot tile in sparse form (mask + even-padded run of non-zero slots).
void sparsify_tile(const uint64_t* tile, std::vector<uint64_t>& mask, std::vector<uint64_t>& val, std::vector<uint32_t>& off);

// Webhook restart-block inputs (nbg_restart_batch): old / updated / mutated Notebook PodSpecs as
// semantic sparse tiles.
struct RestartBatch {
  std::vector<uint8_t> flags;
  std::vector<uint64_t> mask[3], val[3];
  std::vector<uint32_t> off[3] = {{0}, {0}, {0}};
  uint64_t n = 0;
  void add(const GVal& old_nb, const GVal& updated_nb, const GVal& mutated_nb, bool create);
  nbg_restart_batch view() const;
};

// One notebook as the culler sees it: annotations + what the two Jupyter
// endpoints returned (CULL:243-313), all still strings.
struct CullInput {
  bool has_annotations = true;           // metadata.annotations != nil
  bool stop_set = false;                 // kubeflow-resource-stopped present
  bool la_present = false, chk_present = false;
  std::string last_activity, check_ts;   // annotation values
  bool pod_missing = false;
  struct Rec { std::string last_activity, state; };
  std::vector<Rec> kernels, terminals;   // empty == nil for the reference (CULL:373,409)
};
struct CullBatch {
  std::vector<uint32_t> k_off{0}, t_off{0};
  std::vector<int64_t> k_sec, t_sec, ann_sec, chk_sec;
  std::vector<uint32_t> k_nsf, t_nsf, ann_nsf, chk_nsf;
  std::vector<uint8_t> nb_flags;
  std::vector<int32_t> k_zone, t_zone;   // host-only: zone offsets, to format results
  uint64_t n = 0;
  void add(const CullInput& in);
  void append(const CullBatch& o);
  nbg_cull_batch view() const;
};

// ---- fleets -------------------------------------------------------------------
enum FleetKind { FLEET_KUBEFLOW = 2, FLEET_ODH = 4 };  // numbered after BASELINE configs
struct Fleet {
  int kind = FLEET_KUBEFLOW;
  uint64_t seed = 0, n = 0;
  uint32_t n_namespaces = 1;
  std::vector<double> ns_cdf;            // Zipf(s = 1.1) over namespaces
  BuilderEnv env;
  void init();
  uint32_t namespace_of(uint64_t i) const;
  std::string namespace_name(uint64_t i) const;
};
struct Pair {
  GVal notebook, sts_from, sts_to, svc_from, svc_to, vs_from, vs_to;
  int drift_class = 0;
};
void fleet_make(const Fleet& f, uint64_t i, Pair* out);

struct CullFleet {
  uint64_t seed = 0, n = 0;
  int64_t now_sec = 1767225600;          // 2026-01-01T00:00:00Z
  int32_t now_nsec = 123456789;
  double mean_kernels = 4.0;
};
void cullfleet_make(const CullFleet& f, uint64_t i, CullInput* out);
std::string cull_input_json(const CullInput& in);

}  // namespace nbh
