// hostapi.cpp — C ABI of libnbgraft_host.so (include/nbgraft_host.h).
#include <atomic>
#include <chrono>
#include <cstdio>
#include <cstring>
#include <sstream>
#include <thread>

#include "../../../include/nbgraft_host.h"
#include "../hash.h"
#include "fleet.hpp"
#include "json.hpp"
#include "k8s.hpp"
#include "rfc3339.hpp"

#define NBH_API extern "C" __attribute__((visibility("default")))

using namespace nbh;

struct nbh_obj { GVal v; };
struct nbh_sts { StsBatch b; };
struct nbh_svc { SvcBatch b; };
struct nbh_vs { VsBatch b; };
struct nbh_cull { CullBatch b; };
struct nbh_gen { GenBatch b; };
struct nbh_rb { RestartBatch b; };
struct nbh_fleet { Fleet f; };
struct nbh_cullfleet { CullFleet f; };

namespace {

// dynamic parallel-for over blocks with plain std::thread (no OpenMP runtime needed)
int host_threads() {
  const unsigned hc = std::thread::hardware_concurrency();
  const char* e = getenv("NBH_THREADS");
  int n = e ? atoi(e) : (int)hc;
  if (n < 1) n = 1;
  if (n > 256) n = 256;
  return n;
}
template <class F>
void parallel_blocks(uint64_t nblocks, F&& body) {
  const int nt = (int)std::min<uint64_t>((uint64_t)host_threads(), nblocks);
  if (nt <= 1) { for (uint64_t b = 0; b < nblocks; ++b) body(b); return; }
  std::atomic<uint64_t> next{0};
  std::vector<std::thread> pool;
  for (int t = 0; t < nt; ++t)
    pool.emplace_back([&] { for (uint64_t b; (b = next.fetch_add(1)) < nblocks;) body(b); });
  for (auto& th : pool) th.join();
}

int fail(char* err, size_t errlen, const std::string& msg) {
  if (err && errlen) { snprintf(err, errlen, "%s", msg.c_str()); }
  return -1;
}

void dump_json(const GVal& v, std::string* o) {
  auto esc = [&](const std::string& s) {
    o->push_back('"');
    for (char c : s) {
      if (c == '"' || c == '\\') { o->push_back('\\'); o->push_back(c); }
      else if ((unsigned char)c < 0x20) { char b[8]; snprintf(b, sizeof b, "\\u%04x", c); *o += b; }
      else o->push_back(c);
    }
    o->push_back('"');
  };
  switch (v.kind) {
    case GVal::Null: *o += "null"; break;
    case GVal::Bool: *o += v.i ? "true" : "false"; break;
    case GVal::Int: *o += std::to_string(v.i); break;
    case GVal::Float: { std::ostringstream os; os.precision(17); os << v.f; std::string t = os.str();
                        if (t.find_first_of(".eEn") == std::string::npos) { t += ".0"; }
                        *o += t; break; }
    case GVal::Str: esc(v.s); break;
    case GVal::List:
      o->push_back('[');
      for (size_t k = 0; k < v.items.size(); ++k) { if (k) o->push_back(','); dump_json(v.items[k], o); }
      o->push_back(']');
      break;
    case GVal::Map: case GVal::Struct: {
      o->push_back('{');
      bool first = true;
      for (const auto& e : v.kv) {
        if (!first) o->push_back(',');
        first = false;
        esc(e.first); o->push_back(':'); dump_json(e.second, o);
      }
      o->push_back('}');
      break;
    }
    case GVal::Ptr:
      if (v.items.empty()) *o += "\"<extracted>\""; else dump_json(v.items[0], o);
      break;
  }
}

BuilderEnv to_env(const nbh_env* e) {
  BuilderEnv b;
  if (!e) return b;
  b.add_fsgroup = e->add_fsgroup != 0;
  if (e->cluster_domain) b.cluster_domain = e->cluster_domain;
  if (e->istio_host && *e->istio_host) b.istio_host = e->istio_host;
  if (e->istio_gateway && *e->istio_gateway) b.istio_gateway = e->istio_gateway;
  return b;
}

bool cull_input_from_json(const std::string& text, CullInput* in, std::string* err) {
  GVal j;
  if (!parse_json(text, &j, err)) return false;
  if (j.kind != GVal::Map) { *err = "culler input must be a JSON object"; return false; }
  *in = CullInput();
  const GVal* ann = j.find("annotations");
  in->has_annotations = ann && ann->kind == GVal::Map;
  if (in->has_annotations) {
    if (ann->has("kubeflow-resource-stopped")) in->stop_set = true;
    if (const GVal* la = ann->find("notebooks.kubeflow.org/last-activity")) { in->la_present = true; in->last_activity = la->s; }
    if (const GVal* ck = ann->find("notebooks.kubeflow.org/last_activity_check_timestamp")) { in->chk_present = true; in->check_ts = ck->s; }
  }
  if (const GVal* pm = j.find("pod_missing")) in->pod_missing = pm->kind == GVal::Bool && pm->i;
  auto recs = [&](const char* key, std::vector<CullInput::Rec>* out) {
    const GVal* l = j.find(key);
    if (!l || l->kind != GVal::List) return;
    for (const auto& e : l->items) {
      CullInput::Rec r;
      if (e.kind == GVal::Map) { r.last_activity = e.get_str("last_activity"); r.state = e.get_str("execution_state"); }
      out->push_back(std::move(r));
    }
  };
  recs("kernels", &in->kernels);
  recs("terminals", &in->terminals);
  return true;
}

}  // namespace

// ------------------------------------------------------------------ objects
NBH_API int nbh_obj_from_json(const char* type, const char* json, nbh_obj** out, char* err, size_t errlen) {
  if (!type || !json || !out) return fail(err, errlen, "null argument");
  nbh_obj* o = new nbh_obj();
  std::string e;
  if (!decode_typed_json(json, type, &o->v, &e)) { delete o; return fail(err, errlen, e); }
  *out = o;
  return 0;
}
NBH_API void nbh_obj_free(nbh_obj* o) { delete o; }
NBH_API char* nbh_obj_to_json(const nbh_obj* o) {
  std::string s;
  if (o) dump_json(o->v, &s);
  char* r = (char*)malloc(s.size() + 1);
  if (r) memcpy(r, s.c_str(), s.size() + 1);
  return r;
}
NBH_API void nbh_str_free(char* s) { free(s); }
NBH_API int nbh_obj_get(const nbh_obj* o, const char* dotted_path, nbh_obj** out) {
  if (!o || !dotted_path || !out) return -1;
  const GVal* cur = &o->v;
  std::string path(dotted_path);
  size_t at = 0;
  while (!path.empty()) {
    const size_t dot = path.find('.', at);
    const std::string key = path.substr(at, dot == std::string::npos ? std::string::npos : dot - at);
    while (cur->kind == GVal::Ptr && !cur->items.empty()) cur = &cur->items[0];
    if (cur->kind != GVal::Struct && cur->kind != GVal::Map) return -1;
    cur = cur->find(key);
    if (!cur) return -1;
    if (dot == std::string::npos) break;
    at = dot + 1;
  }
  nbh_obj* r = new nbh_obj();
  r->v = *cur;
  *out = r;
  return 0;
}
NBH_API uint64_t nbh_obj_digest(const nbh_obj* o) { return o ? digest(o->v) : 0; }

namespace {
// walk "a.b.0.c" (numbers index lists, pointers are looked through); returns the parent of the last element
GVal* walk_parent(GVal* cur, const std::string& path, std::string* last) {
  size_t at = 0;
  while (true) {
    const size_t dot = path.find('.', at);
    const std::string key = path.substr(at, dot == std::string::npos ? std::string::npos : dot - at);
    while (cur->kind == GVal::Ptr && !cur->items.empty()) cur = &cur->items[0];
    if (dot == std::string::npos) { *last = key; return cur; }
    if (cur->kind == GVal::List) {
      const size_t k = (size_t)atoll(key.c_str());
      if (k >= cur->items.size()) return nullptr;
      cur = &cur->items[k];
    } else if (cur->kind == GVal::Struct || cur->kind == GVal::Map) {
      cur = cur->find(key);
      if (!cur) return nullptr;
    } else {
      return nullptr;
    }
    at = dot + 1;
  }
}
}  // namespace

NBH_API int nbh_obj_erase(nbh_obj* o, const char* dotted_path) {
  if (!o || !dotted_path) return -1;
  std::string last;
  GVal* parent = walk_parent(&o->v, dotted_path, &last);
  if (!parent || (parent->kind != GVal::Struct && parent->kind != GVal::Map)) return -1;
  parent->erase(last);
  return 0;
}
NBH_API int nbh_obj_set_str(nbh_obj* o, const char* dotted_path, const char* value) {
  if (!o || !dotted_path || !value) return -1;
  std::string last;
  GVal* parent = walk_parent(&o->v, dotted_path, &last);
  if (!parent || (parent->kind != GVal::Struct && parent->kind != GVal::Map)) return -1;
  parent->set(last, GVal::str(value));
  return 0;
}

// ----------------------------------------------------------------- builders
NBH_API int nbh_generate_statefulset(const nbh_obj* nb, const nbh_env* env, nbh_obj** out, char* err, size_t errlen) {
  if (!nb || !out) return fail(err, errlen, "null argument");
  nbh_obj* o = new nbh_obj();
  std::string e;
  if (!generate_statefulset(nb->v, to_env(env), &o->v, &e)) { delete o; return fail(err, errlen, e); }
  *out = o;
  return 0;
}
NBH_API int nbh_generate_service(const nbh_obj* nb, nbh_obj** out, char* err, size_t errlen) {
  if (!nb || !out) return fail(err, errlen, "null argument");
  nbh_obj* o = new nbh_obj();
  std::string e;
  if (!generate_service(nb->v, &o->v, &e)) { delete o; return fail(err, errlen, e); }
  *out = o;
  return 0;
}
NBH_API int nbh_generate_virtualservice(const nbh_obj* nb, const nbh_env* env, nbh_obj** out, char* err, size_t errlen) {
  if (!nb || !out) return fail(err, errlen, "null argument");
  nbh_obj* o = new nbh_obj();
  std::string e;
  if (!generate_virtualservice(nb->v, to_env(env), &o->v, &e)) { delete o; return fail(err, errlen, e); }
  *out = o;
  return 0;
}
NBH_API int nbh_odh_mutate(nbh_obj* nb, int flags) {
  if (!nb) return -1;
  OdhInject inj;
  inj.lock = flags & 1; inj.ca_bundle = flags & 2; inj.runtime_images = flags & 4;
  inj.elyra_secret = flags & 8; inj.rbac_proxy = flags & 16; inj.feast = flags & 32;
  return odh_webhook_mutate(&nb->v, inj) ? 0 : 1;     // 1: kube-rbac-proxy injection refused (bad resource annotations)
}
NBH_API int nbh_odh_build(const nbh_obj* nb, const char* what, const char* central_namespace, const char* gateway_name,
                          const char* gateway_namespace, const char* rolebinding_name, const char* role_kind,
                          const char* role_name, nbh_obj** out) {
  if (!nb || !what || !out) return -1;
  OdhBuildEnv env;
  auto s = [](const char* p) { return std::string(p ? p : ""); };
  env.central_namespace = s(central_namespace); env.gateway_name = s(gateway_name); env.gateway_namespace = s(gateway_namespace);
  env.rolebinding_name = s(rolebinding_name); env.role_kind = s(role_kind); env.role_name = s(role_name);
  nbh_obj* o = new nbh_obj();
  if (!odh_build(nb->v, what, env, &o->v)) { delete o; return -1; }
  *out = o;
  return 0;
}
NBH_API int nbh_odh_mutate_proxy(nbh_obj* nb, int flags, const char* http_proxy, const char* https_proxy, const char* no_proxy) {
  if (!nb || !http_proxy || !https_proxy || !no_proxy) return -1;
  OdhInject inj;
  inj.lock = flags & 1; inj.ca_bundle = flags & 2; inj.runtime_images = flags & 4;
  inj.elyra_secret = flags & 8; inj.rbac_proxy = flags & 16; inj.feast = flags & 32;
  inj.proxy_env = true; inj.http_proxy = http_proxy; inj.https_proxy = https_proxy; inj.no_proxy = no_proxy;
  return odh_webhook_mutate(&nb->v, inj) ? 0 : 1;
}

// ---- what Copy*Fields does to `to` once the kernel has said how the pair compares (apply.cpp)
NBH_API int nbh_apply_sts_fields(const nbh_obj* from, nbh_obj* to, unsigned reason, int caller_prestep) {
  if (!from || !to) return -1;
  apply_statefulset_fields(from->v, &to->v, reason, caller_prestep != 0);
  return 0;
}
NBH_API int nbh_apply_svc_fields(const nbh_obj* from, nbh_obj* to) {
  if (!from || !to) return -1;
  apply_service_fields(from->v, &to->v);
  return 0;
}
NBH_API int nbh_apply_vs_fields(const nbh_obj* from, nbh_obj* to, unsigned reason) {
  if (!from || !to) return -1;
  apply_virtualservice(from->v, &to->v, reason);
  return 0;
}

// ------------------------------------------------------------------ batches
NBH_API nbh_sts* nbh_sts_new(void) { return new nbh_sts(); }
NBH_API void nbh_sts_free(nbh_sts* b) { delete b; }
NBH_API int nbh_sts_add(nbh_sts* b, const nbh_obj* from, const nbh_obj* to) {
  if (!b || !from || !to) return -1;
  b->b.add(from->v, to->v);
  return 0;
}
NBH_API uint64_t nbh_sts_view(const nbh_sts* b, nbg_sts_batch* out) {
  if (!b || !out) return 0;
  *out = b->b.view();
  return b->b.n;
}
NBH_API uint64_t nbh_sts_view_sparse(const nbh_sts* b, nbg_sts_sparse_batch* out) {
  if (!b || !out) return 0;
  *out = b->b.view_sparse();
  return b->b.n;
}
NBH_API nbh_svc* nbh_svc_new(void) { return new nbh_svc(); }
NBH_API void nbh_svc_free(nbh_svc* b) { delete b; }
NBH_API int nbh_svc_add(nbh_svc* b, const nbh_obj* from, const nbh_obj* to) {
  if (!b || !from || !to) return -1;
  b->b.add(from->v, to->v);
  return 0;
}
NBH_API uint64_t nbh_svc_view(const nbh_svc* b, nbg_svc_batch* out) {
  if (!b || !out) return 0;
  *out = b->b.view();
  return b->b.n;
}
NBH_API uint64_t nbh_svc_view_gen(const nbh_svc* b, nbg_svc_gen_batch* out) {
  if (!b || !out) return 0;
  *out = b->b.view_gen();
  return b->b.n;
}
NBH_API uint64_t nbh_vs_view_gen(const nbh_vs* b, nbg_vs_gen_batch* out) {
  if (!b || !out) return 0;
  *out = b->b.view_gen();
  return b->b.n;
}
// The notebook-independent words of a generated VirtualService for this environment.
NBH_API int nbh_vs_gen_consts(const nbh_env* env, nbg_vs_gen_consts* out) {
  if (!out) return -1;
  GVal nb = GVal::strct(), meta = GVal::strct();
  meta.set("name", GVal::str("x"));
  meta.set("namespace", GVal::str("y"));
  nb.set("metadata", std::move(meta));
  GVal vs;
  std::string err;
  if (!generate_virtualservice(nb, to_env(env), &vs, &err)) return -1;
  uint64_t slots[NBG_VS_SLOTS];
  std::vector<nbg_kv> hdr;
  bool hdr_nil;
  if (!pack_vs_spec(vs, slots, &hdr, &hdr_nil)) return -1;
  out->istio_host_hash = slots[0];
  out->gateway_hash = slots[1];
  out->residual = slots[7];
  return 0;
}
NBH_API nbh_vs* nbh_vs_new(void) { return new nbh_vs(); }
NBH_API void nbh_vs_free(nbh_vs* b) { delete b; }
NBH_API int nbh_vs_add(nbh_vs* b, const nbh_obj* from, const nbh_obj* to) {
  if (!b || !from || !to) return -1;
  b->b.add(from->v, to->v);
  return 0;
}
NBH_API uint64_t nbh_vs_view(const nbh_vs* b, nbg_vs_batch* out) {
  if (!b || !out) return 0;
  *out = b->b.view();
  return b->b.n;
}

// ODH comparator pairs: `labels_path` / `spec_path` are dotted paths inside the two objects
// ("metadata.labels", "spec", "data", "subjects"); a missing path is the nil / zero value.
NBH_API nbh_gen* nbh_gen_new(void) { return new nbh_gen(); }
NBH_API void nbh_gen_free(nbh_gen* b) { delete b; }
NBH_API int nbh_gen_add(nbh_gen* b, const nbh_obj* a, const nbh_obj* o, const char* labels_path, const char* spec_path,
                        int mode) {
  if (!b || !a || !o || !labels_path || !spec_path) return -1;
  auto at = [](const GVal& root, const std::string& path) -> GVal {
    const GVal* cur = &root;
    size_t pos = 0;
    while (!path.empty()) {
      const size_t dot = path.find('.', pos);
      const std::string key = path.substr(pos, dot == std::string::npos ? std::string::npos : dot - pos);
      while (cur->kind == GVal::Ptr && !cur->items.empty()) cur = &cur->items[0];
      if (cur->kind != GVal::Struct && cur->kind != GVal::Map) return GVal();
      cur = cur->find(key);
      if (!cur) return GVal();
      if (dot == std::string::npos) break;
      pos = dot + 1;
    }
    return *cur;
  };
  b->b.add(at(a->v, labels_path), at(o->v, labels_path), at(a->v, spec_path), at(o->v, spec_path), mode == NBG_GEN_DEEPEQUAL);
  return 0;
}
NBH_API uint64_t nbh_gen_view(const nbh_gen* b, nbg_generic_batch* out) {
  if (!b || !out) return 0;
  *out = b->b.view();
  return b->b.n;
}

NBH_API nbh_rb* nbh_rb_new(void) { return new nbh_rb(); }
NBH_API void nbh_rb_free(nbh_rb* b) { delete b; }
NBH_API int nbh_rb_add(nbh_rb* b, const nbh_obj* old_nb, const nbh_obj* updated_nb, const nbh_obj* mutated_nb, int create) {
  if (!b || !old_nb || !updated_nb || !mutated_nb) return -1;
  b->b.add(old_nb->v, updated_nb->v, mutated_nb->v, create != 0);
  return 0;
}
NBH_API uint64_t nbh_rb_view(const nbh_rb* b, nbg_restart_batch* out) {
  if (!b || !out) return 0;
  *out = b->b.view();
  return b->b.n;
}
NBH_API nbh_obj* nbh_obj_clone(const nbh_obj* o) {
  if (!o) return nullptr;
  nbh_obj* r = new nbh_obj();
  r->v = o->v;
  return r;
}

NBH_API nbh_cull* nbh_cull_new(void) { return new nbh_cull(); }
NBH_API void nbh_cull_free(nbh_cull* b) { delete b; }
NBH_API int nbh_cull_add_json(nbh_cull* b, const char* json, char* err, size_t errlen) {
  if (!b || !json) return fail(err, errlen, "null argument");
  CullInput in;
  std::string e;
  if (!cull_input_from_json(json, &in, &e)) return fail(err, errlen, e);
  b->b.add(in);
  return 0;
}
NBH_API uint64_t nbh_cull_view(const nbh_cull* b, nbg_cull_batch* out) {
  if (!b || !out) return 0;
  *out = b->b.view();
  return b->b.n;
}
NBH_API uint64_t nbh_cull_records(const nbh_cull* b, uint64_t* kernels, uint64_t* terminals) {
  if (!b) return 0;
  if (kernels) *kernels = b->b.k_sec.size();
  if (terminals) *terminals = b->b.t_sec.size();
  return b->b.k_sec.size() + b->b.t_sec.size();
}
NBH_API int nbh_cull_new_last_activity(const nbh_cull* b, uint64_t i, uint8_t action, uint32_t argmax,
                                       int64_t now_sec, int32_t now_offset_sec, char* buf, size_t buflen) {
  if (!b || !buf || !buflen || i >= b->b.n) return -1;
  std::string s;
  switch (action & NBG_ACT_SRC_MASK) {
    case 0: break;
    case 1: s = format_rfc3339(now_sec, now_offset_sec); break;                  // createTimestamp(), CULL:512-515
    case 2: {
      const uint64_t r = (uint64_t)b->b.k_off[i] + argmax;
      if (r >= b->b.k_off[i + 1]) return -1;
      s = format_rfc3339(b->b.k_sec[r], b->b.k_zone[r]);                          // CULL:348
      break;
    }
    case 3: {
      const uint64_t r = (uint64_t)b->b.t_off[i] + argmax;
      if (r >= b->b.t_off[i + 1]) return -1;
      s = format_rfc3339(b->b.t_sec[r], b->b.t_zone[r]);
      break;
    }
  }
  if (s.size() + 1 > buflen) return -1;
  memcpy(buf, s.c_str(), s.size() + 1);
  return (int)s.size();
}

NBH_API int nbh_parse_rfc3339(const char* s, int64_t* sec, int32_t* nsec, int32_t* offset_sec) {
  Instant t;
  if (!s || !parse_rfc3339(s, strlen(s), &t)) return 0;
  if (sec) *sec = t.sec;
  if (nsec) *nsec = t.nsec;
  if (offset_sec) *offset_sec = t.offset;
  return 1;
}
// Host-side counterpart of nbg_parse_records (same inputs, same outputs) on all host threads:
// what a packer without the device parser has to do, and the CPU baseline of that kernel.
NBH_API int nbh_parse_records(const uint32_t* off, const uint8_t* bytes, const uint32_t* s_off, const uint8_t* s_bytes,
                              uint64_t R, int64_t* sec, uint32_t* nsf, int32_t* zone) {
  if (!off || !sec || !nsf) return -1;
  const uint64_t kBlock = 16384, nblocks = (R + kBlock - 1) / kBlock;
  parallel_blocks(nblocks, [&](uint64_t blk) {
    const uint64_t lo = blk * kBlock, hi = lo + kBlock < R ? lo + kBlock : R;
    for (uint64_t r = lo; r < hi; ++r) {
      Instant t;
      uint32_t w = 0;
      if (parse_rfc3339((const char*)bytes + off[r], off[r + 1] - off[r], &t)) w = (uint32_t)t.nsec | NBG_REC_PARSE_OK;
      else t = Instant();
      if (s_off && s_off[r + 1] - s_off[r] == 4 && memcmp(s_bytes + s_off[r], "idle", 4) == 0) w |= NBG_REC_IDLE;
      sec[r] = t.sec; nsf[r] = w;
      if (zone) zone[r] = t.offset;
    }
  });
  return 0;
}
NBH_API int nbh_format_rfc3339(int64_t sec, int32_t offset_sec, char* buf, size_t buflen) {
  const std::string s = format_rfc3339(sec, offset_sec);
  if (!buf || s.size() + 1 > buflen) return -1;
  memcpy(buf, s.c_str(), s.size() + 1);
  return (int)s.size();
}

// --------------------------------------------------------------- owner join
namespace {
uint64_t join_key(uint64_t ns_hash, uint64_t uid_hash) {   // == nbg_join_key (capi.cu)
  const uint64_t k = nbg_combine(nbg_combine(0x6A6F696E6B657921ull, ns_hash), uid_hash);
  return k ? k : 1;
}
const GVal* meta_of(const GVal& o) { return (o.kind == GVal::Struct || o.kind == GVal::Map) ? o.find("metadata") : nullptr; }
}  // namespace

NBH_API uint64_t nbh_join_key_owner(const nbh_obj* owner) {
  const GVal* m = owner ? meta_of(owner->v) : nullptr;
  if (!m) return 0;
  return join_key(hash_str(m->get_str("namespace")), hash_str(m->get_str("uid")));
}
// metav1.GetControllerOf: the first owner reference with controller != nil && *controller
NBH_API uint64_t nbh_join_key_controlled(const nbh_obj* obj) {
  const GVal* m = obj ? meta_of(obj->v) : nullptr;
  if (!m) return 0;
  const GVal* refs = m->find("ownerReferences");
  if (!refs || refs->kind != GVal::List) return 0;
  for (const auto& r : refs->items) {
    const GVal* c = (r.kind == GVal::Struct || r.kind == GVal::Map) ? r.find("controller") : nullptr;
    const bool is_ctrl = c && ((c->kind == GVal::Ptr && !c->items.empty() && c->items[0].i) || (c->kind == GVal::Bool && c->i));
    if (is_ctrl) return join_key(hash_str(m->get_str("namespace")), hash_str(r.get_str("uid")));
  }
  return 0;
}

// ------------------------------------------------------------------- fleets
NBH_API nbh_fleet* nbh_fleet_new(int kind, uint64_t seed, uint64_t n, uint32_t n_namespaces) {
  if (kind != FLEET_KUBEFLOW && kind != FLEET_ODH) return nullptr;
  nbh_fleet* f = new nbh_fleet();
  f->f.kind = kind; f->f.seed = seed; f->f.n = n; f->f.n_namespaces = n_namespaces;
  f->f.init();
  return f;
}
NBH_API void nbh_fleet_free(nbh_fleet* f) { delete f; }
NBH_API int nbh_fleet_get(const nbh_fleet* f, uint64_t i, int which, nbh_obj** out) {
  if (!f || !out || which < 0 || which > 6) return -1;
  Pair p;
  fleet_make(f->f, i, &p);
  nbh_obj* o = new nbh_obj();
  GVal* src[] = {&p.notebook, &p.sts_from, &p.sts_to, &p.svc_from, &p.svc_to, &p.vs_from, &p.vs_to};
  o->v = std::move(*src[which]);
  *out = o;
  return 0;
}
NBH_API int nbh_fleet_drift_class(const nbh_fleet* f, uint64_t i) {
  if (!f) return -1;
  Pair p;
  fleet_make(f->f, i, &p);
  return p.drift_class;
}
NBH_API uint64_t nbh_fleet_ns_hash(const nbh_fleet* f, uint64_t i) {
  if (!f) return 0;
  return hash_str(f->f.namespace_name(i));
}

NBH_API int nbh_threads(void) { return host_threads(); }

NBH_API uint64_t nbh_fleet_shard_indices(const nbh_fleet* f, uint32_t n_shards, uint32_t shard, uint64_t* out,
                                         uint64_t cap) {
  if (!f || n_shards == 0) return 0;
  const uint64_t n = f->f.n, kBlock = 65536, nblocks = (n + kBlock - 1) / kBlock;
  // the namespace -> shard map is tiny: compute it once per namespace
  std::vector<uint8_t> mine(f->f.n_namespaces);
  for (uint32_t k = 0; k < f->f.n_namespaces; ++k) {
    char buf[32];
    snprintf(buf, sizeof buf, "ns-%06u", k);
    mine[k] = nbg_shard_of_inline(nbg_hash_bytes(buf, strlen(buf)), n_shards) == shard;
  }
  std::vector<std::vector<uint64_t>> parts(nblocks);
  parallel_blocks(nblocks, [&](uint64_t blk) {
    const uint64_t lo = blk * kBlock, hi = lo + kBlock < n ? lo + kBlock : n;
    for (uint64_t i = lo; i < hi; ++i)
      if (mine[f->f.namespace_of(i)]) parts[blk].push_back(i);
  });
  uint64_t count = 0;
  for (const auto& v : parts)
    for (uint64_t i : v) { if (out && count < cap) out[count] = i; ++count; }
  return count;
}

NBH_API int nbh_fleet_objects(const nbh_fleet* f, const uint64_t* idx, uint64_t i0, uint64_t m, nbh_obj** out) {
  if (!f || !out) return -1;
  const uint64_t kBlock = 256, nblocks = (m + kBlock - 1) / kBlock;
  parallel_blocks(nblocks, [&](uint64_t blk) {
    const uint64_t lo = blk * kBlock, hi = lo + kBlock < m ? lo + kBlock : m;
    Pair p;
    for (uint64_t k = lo; k < hi; ++k) {
      fleet_make(f->f, idx ? idx[k] : i0 + k, &p);
      GVal* src[] = {&p.notebook, &p.sts_from, &p.sts_to, &p.svc_from, &p.svc_to, &p.vs_from, &p.vs_to};
      for (int w = 0; w < 7; ++w) { nbh_obj* o = new nbh_obj(); o->v = std::move(*src[w]); out[k * 7 + w] = o; }
    }
  });
  return 0;
}

// CPU seconds (summed over threads) the last nbh_fleet_pack spent synthesising objects and packing them.
static std::atomic<uint64_t> g_make_ns{0}, g_pack_ns{0};
NBH_API void nbh_fleet_pack_timing(double* make_cpu_s, double* pack_cpu_s) {
  if (make_cpu_s) *make_cpu_s = g_make_ns.load() * 1e-9;
  if (pack_cpu_s) *pack_cpu_s = g_pack_ns.load() * 1e-9;
}

namespace {
int fleet_pack_impl(const nbh_fleet* f, const uint64_t* idx, uint64_t i0, uint64_t m,
                    nbh_sts** sts, nbh_svc** svc, nbh_vs** vs, bool lean) {
  if (!f) return -1;
  g_make_ns = 0; g_pack_ns = 0;
  typedef std::chrono::steady_clock clk;
  const uint64_t kBlock = 2048;
  const uint64_t nblocks = (m + kBlock - 1) / kBlock;
  std::vector<StsBatch> ps(sts ? nblocks : 0);
  std::vector<SvcBatch> pv(svc ? nblocks : 0);
  std::vector<VsBatch> pw(vs ? nblocks : 0);
  for (auto& b : ps) b.lean = lean;
  parallel_blocks(nblocks, [&](uint64_t blk) {
    const uint64_t lo = blk * kBlock, hi = lo + kBlock < m ? lo + kBlock : m;
    Pair p;
    uint64_t make_ns = 0, pack_ns = 0;
    for (uint64_t k = lo; k < hi; ++k) {
      const auto t0 = clk::now();
      fleet_make(f->f, idx ? idx[k] : i0 + k, &p);
      const auto t1 = clk::now();
      if (sts) ps[blk].add(p.sts_from, p.sts_to);
      if (svc) pv[blk].add(p.svc_from, p.svc_to);
      if (vs) pw[blk].add(p.vs_from, p.vs_to);
      const auto t2 = clk::now();
      make_ns += std::chrono::duration_cast<std::chrono::nanoseconds>(t1 - t0).count();
      pack_ns += std::chrono::duration_cast<std::chrono::nanoseconds>(t2 - t1).count();
    }
    g_make_ns += make_ns; g_pack_ns += pack_ns;
  });
  if (sts) {
    nbh_sts* o = new nbh_sts();
    o->b.lean = lean;
    if (!lean) { o->b.pod_from.reserve(m * NBG_POD_SLOTS); o->b.pod_to.reserve(m * NBG_POD_SLOTS); }
    // the value runs are the bulk of a sparse batch: size them once instead of letting the vectors double
    size_t vf = 0, vt = 0;
    for (const auto& b : ps) { vf += b.val_from.size(); vt += b.val_to.size(); }
    if (vf >= (1ull << 32) || vt >= (1ull << 32)) { delete o; return -1; }   // run offsets are 32-bit (nbgraft.h)
    o->b.val_from.reserve(vf); o->b.val_to.reserve(vt);
    o->b.mask_from.reserve(3 * m); o->b.mask_to.reserve(3 * m);
    for (auto& b : ps) { o->b.append(b); b = StsBatch(); }
    *sts = o;
  }
  if (svc) {
    nbh_svc* o = new nbh_svc();
    o->b.ports_from.reserve(m * NBG_SVC_SLOTS); o->b.ports_to.reserve(m * NBG_SVC_SLOTS);
    for (auto& b : pv) { o->b.append(b); b = SvcBatch(); }
    *svc = o;
  }
  if (vs) {
    nbh_vs* o = new nbh_vs();
    o->b.spec_from.reserve(m * NBG_VS_SLOTS); o->b.spec_to.reserve(m * NBG_VS_SLOTS);
    for (auto& b : pw) { o->b.append(b); b = VsBatch(); }
    *vs = o;
  }
  return 0;
}
}  // namespace

NBH_API int nbh_fleet_pack(const nbh_fleet* f, const uint64_t* idx, uint64_t i0, uint64_t m,
                           nbh_sts** sts, nbh_svc** svc, nbh_vs** vs) {
  return fleet_pack_impl(f, idx, i0, m, sts, svc, vs, false);
}
NBH_API int nbh_fleet_pack_lean(const nbh_fleet* f, const uint64_t* idx, uint64_t i0, uint64_t m,
                                nbh_sts** sts, nbh_svc** svc, nbh_vs** vs) {
  return fleet_pack_impl(f, idx, i0, m, sts, svc, vs, true);
}
// Pack objects the caller already holds (rows of 7 handles in nbh_fleet_objects order; the Notebook itself,
// row[0], is not read): the work a controller has when its informer hands it objects.  All host threads.
NBH_API int nbh_pack_objects(const nbh_obj* const* objs, uint64_t m, int lean, nbh_sts** sts, nbh_svc** svc, nbh_vs** vs) {
  if (!objs && m) return -1;
  for (uint64_t k = 0; k < m * 7; ++k) if (!objs[k]) return -1;
  const uint64_t kBlock = 2048, nblocks = (m + kBlock - 1) / kBlock;
  std::vector<StsBatch> ps(sts ? nblocks : 0);
  std::vector<SvcBatch> pv(svc ? nblocks : 0);
  std::vector<VsBatch> pw(vs ? nblocks : 0);
  for (auto& b : ps) b.lean = lean != 0;
  parallel_blocks(nblocks, [&](uint64_t blk) {
    const uint64_t lo = blk * kBlock, hi = lo + kBlock < m ? lo + kBlock : m;
    for (uint64_t k = lo; k < hi; ++k) {
      const nbh_obj* const* r = objs + k * 7;
      if (sts) ps[blk].add(r[1]->v, r[2]->v);
      if (svc) pv[blk].add(r[3]->v, r[4]->v);
      if (vs) pw[blk].add(r[5]->v, r[6]->v);
    }
  });
  if (sts) { nbh_sts* o = new nbh_sts(); o->b.lean = lean != 0; for (auto& b : ps) { o->b.append(b); b = StsBatch(); } *sts = o; }
  if (svc) { nbh_svc* o = new nbh_svc(); for (auto& b : pv) { o->b.append(b); b = SvcBatch(); } *svc = o; }
  if (vs) { nbh_vs* o = new nbh_vs(); for (auto& b : pw) { o->b.append(b); b = VsBatch(); } *vs = o; }
  return 0;
}
// ---- generateStatefulSet on the device: the batch's `from` side from the Notebook itself (nbg_diff_sts_gen)
NBH_API int nbh_sts_add_gen(nbh_sts* b, const nbh_obj* notebook, const nbh_obj* to, const nbh_env* env) {
  if (!b || !notebook || !to) return -1;
  BuilderEnv e;
  if (env) {
    e.add_fsgroup = env->add_fsgroup != 0;
    if (env->cluster_domain) e.cluster_domain = env->cluster_domain;
    if (env->istio_host && *env->istio_host) e.istio_host = env->istio_host;
    if (env->istio_gateway && *env->istio_gateway) e.istio_gateway = env->istio_gateway;
  }
  return b->b.add_gen(notebook->v, to->v, e) ? 0 : -1;
}
NBH_API int nbh_sts_add_notebook(nbh_sts* b, const nbh_obj* notebook, const nbh_obj* to, const nbh_env* env) {
  if (!b || !notebook || !to) return -1;
  BuilderEnv e;
  if (env) e.add_fsgroup = env->add_fsgroup != 0;
  return b->b.add_notebook(notebook->v, to->v, e) ? 0 : -1;
}
NBH_API uint64_t nbh_sts_view_gen(const nbh_sts* b, nbg_sts_gen_inputs* out) {
  if (!b || !out || b->b.prefix_hash.size() != b->b.n) return 0;
  out->prefix_hash = b->b.prefix_hash.data();
  out->flags = b->b.gen_flags.data();
  return b->b.n;
}
NBH_API int nbh_sts_gen_consts(const nbh_env* env, nbg_sts_gen_consts* out) {
  if (!out) return -1;
  // the default port exactly as generateStatefulSet builds it (NBCTL:499-505), through the packer itself
  GVal port = GVal::strct();
  port.set("containerPort", GVal::integer(8888));
  port.set("name", GVal::str("notebook-port"));
  port.set("protocol", GVal::str("TCP"));
  GVal ports = GVal::list();
  ports.items.push_back(std::move(port));
  GVal c = GVal::strct();
  c.set("ports", std::move(ports));
  GVal conts = GVal::list();
  conts.items.push_back(std::move(c));
  GVal pod = GVal::strct();
  pod.set("containers", std::move(conts));
  uint64_t tile[NBG_POD_SLOTS];
  pack_podspec(pod, tile);
  out->working_dir = nbg_hash_bytes("/home/jovyan", 12);
  out->port_name = tile[52];                               // containers[0].ports[0]: name, then the digest of the rest
  out->port_rest = tile[53];
  out->env_name = nbg_hash_bytes("NB_PREFIX", 9);
  out->add_fsgroup = (!env || env->add_fsgroup) ? 1u : 0u;
  out->pad = 0;
  return 0;
}
NBH_API int nbh_sts_is_lean(const nbh_sts* b) { return b && b->b.lean ? 1 : 0; }

// Namespace table of the fleet: ns_hash[k] = nbg_hash64 of namespace k's name, ns_size[k] = notebooks in it.
NBH_API uint32_t nbh_fleet_ns_table(const nbh_fleet* f, uint64_t* ns_hash, uint64_t* ns_size) {
  if (!f) return 0;
  const uint32_t K = f->f.n_namespaces;
  if (ns_hash)
    for (uint32_t k = 0; k < K; ++k) {
      char buf[32];
      snprintf(buf, sizeof buf, "ns-%06u", k);
      ns_hash[k] = nbg_hash_bytes(buf, strlen(buf));
    }
  if (ns_size) {
    const uint64_t n = f->f.n, kBlock = 65536, nblocks = (n + kBlock - 1) / kBlock;
    std::vector<std::atomic<uint64_t>> cnt(K);
    for (auto& c : cnt) c = 0;
    parallel_blocks(nblocks, [&](uint64_t blk) {
      const uint64_t lo = blk * kBlock, hi = lo + kBlock < n ? lo + kBlock : n;
      std::vector<uint32_t> local(K, 0);
      for (uint64_t i = lo; i < hi; ++i) ++local[f->f.namespace_of(i)];
      for (uint32_t k = 0; k < K; ++k) if (local[k]) cnt[k] += local[k];
    });
    for (uint32_t k = 0; k < K; ++k) ns_size[k] = cnt[k].load();
  }
  return K;
}

// Indices i in [0,n) whose namespace k has shard_of_ns[k] == shard, ascending (any namespace -> shard table:
// nbg_shard_of per namespace, or the size-aware one of nbg_balance_shards).
NBH_API uint64_t nbh_fleet_shard_indices_by_table(const nbh_fleet* f, const uint32_t* shard_of_ns, uint32_t shard,
                                                  uint64_t* out, uint64_t cap) {
  if (!f || !shard_of_ns) return 0;
  const uint64_t n = f->f.n, kBlock = 65536, nblocks = (n + kBlock - 1) / kBlock;
  std::vector<std::vector<uint64_t>> parts(nblocks);
  parallel_blocks(nblocks, [&](uint64_t blk) {
    const uint64_t lo = blk * kBlock, hi = lo + kBlock < n ? lo + kBlock : n;
    for (uint64_t i = lo; i < hi; ++i)
      if (shard_of_ns[f->f.namespace_of(i)] == shard) parts[blk].push_back(i);
  });
  uint64_t count = 0;
  for (const auto& v : parts)
    for (uint64_t i : v) { if (out && count < cap) out[count] = i; ++count; }
  return count;
}

// A synthetic "List(StatefulSets)" for notebooks [0,n): most notebooks have one StatefulSet,
// 3 % none yet, 2 % two (the first in list order must win), plus 2 % StatefulSets without a
// controller reference and 1 % controlled by a Notebook that no longer exists.  The list is in
// a seeded random order.  sts_owner[j] = owning notebook index, -1 (no controller), -2 (gone).
NBH_API uint64_t nbh_fleet_join_lists(const nbh_fleet* f, uint64_t n, uint64_t* nb_keys, uint64_t* sts_keys,
                                      int64_t* sts_owner, uint64_t cap) {
  if (!f) return 0;
  struct Ent { uint64_t order, key; int64_t owner; };
  std::vector<Ent> ents;
  ents.reserve(n + n / 8);
  for (uint64_t i = 0; i < n; ++i) {
    Rng r(f->f.seed ^ nbg_mix64(i * 2 + 0x6A6F696Eull));
    char uid[40];
    snprintf(uid, sizeof uid, "%016llx-nb", (unsigned long long)nbg_mix64(f->f.seed ^ (i * 0x9E3779B97F4A7C15ull)));
    const uint64_t nsh = hash_str(f->f.namespace_name(i));
    const uint64_t key = join_key(nsh, hash_str(uid));
    if (nb_keys) nb_keys[i] = key;
    const double u = r.unit();
    const int copies = u < 0.03 ? 0 : u < 0.05 ? 2 : 1;
    for (int c = 0; c < copies; ++c) ents.push_back({r.next(), key, (int64_t)i});
    if (r.chance(0.02)) ents.push_back({r.next(), 0, -1});
    if (r.chance(0.01)) ents.push_back({r.next(), join_key(nsh, hash_str(std::string(uid) + "-deleted")), -2});
  }
  std::sort(ents.begin(), ents.end(), [](const Ent& a, const Ent& b) { return a.order < b.order; });
  for (uint64_t j = 0; j < ents.size() && j < cap; ++j) {
    if (sts_keys) sts_keys[j] = ents[j].key;
    if (sts_owner) sts_owner[j] = ents[j].owner;
  }
  return ents.size();
}

NBH_API nbh_cullfleet* nbh_cullfleet_new(uint64_t seed, uint64_t n, int64_t now_sec, int32_t now_nsec, double mean_kernels) {
  nbh_cullfleet* f = new nbh_cullfleet();
  f->f.seed = seed; f->f.n = n; f->f.now_sec = now_sec; f->f.now_nsec = now_nsec;
  if (mean_kernels > 0) f->f.mean_kernels = mean_kernels;
  return f;
}
NBH_API void nbh_cullfleet_free(nbh_cullfleet* f) { delete f; }
NBH_API char* nbh_cullfleet_json(const nbh_cullfleet* f, uint64_t i) {
  if (!f) return nullptr;
  CullInput in;
  cullfleet_make(f->f, i, &in);
  const std::string s = cull_input_json(in);
  char* r = (char*)malloc(s.size() + 1);
  if (r) memcpy(r, s.c_str(), s.size() + 1);
  return r;
}
NBH_API int nbh_cullfleet_pack(const nbh_cullfleet* f, const uint64_t* idx, uint64_t i0, uint64_t m, nbh_cull** out) {
  if (!f || !out) return -1;
  const uint64_t kBlock = 8192;
  const uint64_t nblocks = (m + kBlock - 1) / kBlock;
  std::vector<CullBatch> parts(nblocks);
  parallel_blocks(nblocks, [&](uint64_t blk) {
    const uint64_t lo = blk * kBlock, hi = lo + kBlock < m ? lo + kBlock : m;
    CullInput in;
    for (uint64_t k = lo; k < hi; ++k) {
      cullfleet_make(f->f, idx ? idx[k] : i0 + k, &in);
      parts[blk].add(in);
    }
  });
  nbh_cull* o = new nbh_cull();
  for (auto& b : parts) { o->b.append(b); b = CullBatch(); }
  *out = o;
  return 0;
}
