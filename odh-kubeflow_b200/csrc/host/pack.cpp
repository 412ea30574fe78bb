// pack.cpp — flatten Go-shaped objects (GVal) into the structure-of-arrays
// schema of include/nbgraft.h.
//
// Soundness rule ("injective up to hash collision"): a tile must be a function
// of the whole value, so that   tiles equal  <=>  reflect.DeepEqual(values).
// Every tile therefore consists of
//   * named slots holding a leaf (raw integer, or the 64-bit hash of a string,
//     "" == 0) or the digest of a named sub-tree,
//   * count tags that tell nil / empty / n for every list whose elements are
//     slotted by position, and
//   * a residual digest of everything that has no slot of its own (including
//     list elements beyond the slotted capacity and fields outside the
//     schema).  Nothing is ever dropped.
// tests/test_pack.py checks the equivalence against the semantic oracle on
// generated and hand-built pairs.
#include <cstring>

#include "../hash.h"
#include "k8s.hpp"
#include "quantity.hpp"

namespace nbh {

uint64_t hash_str(const std::string& s) { return nbg_hash_bytes(s.data(), s.size()); }

namespace {

constexpr uint64_t K_NULL = 0x6E696C6E696C6E69ull, K_BOOL = 0xB001, K_INT = 0x1A7, K_FLOAT = 0xF10A7,
                   K_STR = 0x57A, K_LIST = 0x1157, K_MAP = 0x3A9, K_STRUCT = 0x57AC7, K_PTR = 0x97A,
                   K_MARK = 0x3A4B3A4B3A4B3A4Bull;

const GVal kZero;
const GVal& field(const GVal& s, const char* name) {
  const GVal* p = (s.kind == GVal::Struct || s.kind == GVal::Map) ? s.find(name) : nullptr;
  return p ? *p : kZero;
}
uint64_t str_of(const GVal& s, const char* name) {
  const GVal& f = field(s, name);
  return f.kind == GVal::Str ? hash_str(f.s) : 0;
}
// (a marker node is a Ptr without pointee, which no decoder can produce: digest K_PTR over K_MARK)

// count tag of a slice field: 0xFFFF nil, else min(len, 0xFFFE)
uint64_t count_tag(const GVal& l) {
  if (l.kind != GVal::List) return 0xFFFF;
  return l.items.size() > 0xFFFE ? 0xFFFE : l.items.size();
}
// digest of elements [from, end) of a list (0 when there are none)
uint64_t overflow_digest(const GVal& l, size_t from) {
  if (l.kind != GVal::List || l.items.size() <= from) return 0;
  uint64_t h = nbg_combine(K_LIST, l.items.size() - from);
  for (size_t k = from; k < l.items.size(); ++k) h = nbg_combine(h, digest(l.items[k]));
  return h;
}
// digest of a struct with some fields left out — same value as digest(copy with the fields
// erased), computed without copying the tree (0 when nothing is left)
uint64_t digest_without(const GVal& s, std::initializer_list<const char*> skip) {
  if (s.kind != GVal::Struct) return digest(s);
  uint64_t h = K_STRUCT;
  size_t n = 0;
  for (const auto& e : s.kv) {
    if (e.second.is_zero()) continue;
    bool skipped = false;
    for (const char* k : skip) if (e.first == k) { skipped = true; break; }
    if (skipped) continue;
    h = nbg_combine(nbg_combine(h, hash_str(e.first)), digest(e.second));
    ++n;
  }
  return n ? nbg_combine(h, n) : 0;
}

// A resource.Quantity: ONE word, the digest of the struct reflect.DeepEqual walks - int64Amount{value, scale}, the
// inf.Dec, the cached string s and the Format - so "1" != "1000m" and "1Gi" != "1024Mi" (COMMON:128).  The other two
// slots of the triple stay zero (the sparse form pays per non-zero slot).  An absent entry is 0; a present one is not,
// whatever its value: digest() of a struct never returns 0.
void pack_quantity(const GVal& q, bool present, uint64_t* s3) {
  s3[0] = s3[1] = s3[2] = 0;
  if (present) s3[0] = digest(q);
}

constexpr int kPodLevel = 16, kVolSlots = 3, kMaxVol = 8, kContBase = kPodLevel + kVolSlots * kMaxVol,
              kContSlots = 76, kMaxCont = 2;
constexpr int kMaxPorts = 2, kMaxEnv = 14, kMaxMounts = 10;
static_assert(kContBase + kContSlots * kMaxCont == (int)NBG_POD_SLOTS, "PodSpec tile layout");

// A list element that has a name (container port, volume mount) in its two slots: ONE word for the whole element,
// name and everything else folded together, the second slot zero - an element is what drifts, and the sparse form
// pays per non-zero slot.  (Env variables keep name and value apart: setPrefixEnvVar looks names up, NBCTL:417-431.)
void pack_named_element(const GVal& e, uint64_t* p) {
  p[0] = nbg_combine(str_of(e, "name"), digest_without(e, {"name"}));
  p[1] = 0;
}

void pack_container(const GVal& c, uint64_t* s) {
  s[0] = str_of(c, "name");
  s[1] = str_of(c, "image");
  s[2] = str_of(c, "workingDir");
  s[3] = str_of(c, "imagePullPolicy");
  const GVal &cmd = field(c, "command"), &args = field(c, "args"), &ports = field(c, "ports"),
             &env = field(c, "env"), &mounts = field(c, "volumeMounts"), &env_from = field(c, "envFrom");
  s[4] = cmd.is_null() ? 0 : digest(cmd);
  s[5] = args.is_null() ? 0 : digest(args);
  s[6] = count_tag(ports) | (count_tag(env) << 16) | (count_tag(mounts) << 32) | (count_tag(env_from) << 48);
  const GVal& sc = field(c, "securityContext");
  s[7] = sc.is_null() ? 0 : digest(sc);
  const GVal &lp = field(c, "livenessProbe"), &rp = field(c, "readinessProbe");
  s[8] = lp.is_null() ? 0 : digest(lp);
  s[9] = rp.is_null() ? 0 : digest(rp);
  s[10] = env_from.is_null() ? 0 : digest(env_from);

  uint64_t* p = s + 12;
  for (int k = 0; k < kMaxPorts; ++k, p += 2) {
    p[0] = p[1] = 0;
    if (ports.kind == GVal::List && (size_t)k < ports.items.size()) {
      pack_named_element(ports.items[k], p);
    }
  }
  for (int k = 0; k < kMaxEnv; ++k, p += 2) {
    p[0] = p[1] = 0;
    if (env.kind == GVal::List && (size_t)k < env.items.size()) {
      const GVal& e = env.items[k];
      p[0] = str_of(e, "name");
      bool plain = e.kind == GVal::Struct;
      if (plain) for (const auto& f : e.kv) if (f.first != "name" && f.first != "value") plain = false;
      p[1] = plain ? str_of(e, "value") : nbg_combine(K_MARK, digest_without(e, {"name"}));
    }
  }
  const GVal& res = field(c, "resources");
  const GVal &lim = field(res, "limits"), &req = field(res, "requests");
  static const char* rn[2] = {"cpu", "memory"};
  for (int side = 0; side < 2; ++side) {
    const GVal& m = side == 0 ? lim : req;
    for (int r = 0; r < 2; ++r, p += 3) {
      const GVal* q = m.kind == GVal::Map ? m.find(rn[r]) : nullptr;
      pack_quantity(q ? *q : kZero, q != nullptr, p);
    }
  }
  for (int k = 0; k < kMaxMounts; ++k, p += 2) {
    p[0] = p[1] = 0;
    if (mounts.kind == GVal::List && (size_t)k < mounts.items.size()) {
      pack_named_element(mounts.items[k], p);
    }
  }
  // residual: the container minus everything slotted above; nil-ness and emptiness of the
  // resource maps survive because only the four slotted entries are left out
  uint64_t h = K_STRUCT;
  size_t nrest = 0;
  for (const auto& e : c.kind == GVal::Struct ? c.kv : kZero.kv) {
    if (e.second.is_zero()) continue;
    static const char* slotted[] = {"name", "image", "workingDir", "imagePullPolicy", "command", "args", "securityContext",
                                    "livenessProbe", "readinessProbe", "envFrom", "ports", "env", "volumeMounts"};
    bool skip = false;
    for (const char* k : slotted) if (e.first == k) { skip = true; break; }
    if (skip) continue;
    uint64_t dv;
    if (e.first == "resources" && e.second.kind == GVal::Struct) {
      // digest(resources) with cpu/memory left out of limits and requests
      uint64_t hr = K_STRUCT;
      size_t nr = 0;
      for (const auto& f : e.second.kv) {
        if (f.second.is_zero()) continue;
        uint64_t df;
        if ((f.first == "limits" || f.first == "requests") && f.second.kind == GVal::Map) {
          uint64_t hm = K_MAP;
          size_t nm = 0;
          for (const auto& q : f.second.kv) {
            if (q.first == "cpu" || q.first == "memory") continue;
            hm = nbg_combine(nbg_combine(hm, hash_str(q.first)), digest(q.second));
            ++nm;
          }
          df = nbg_combine(hm, nm);
        } else {
          df = digest(f.second);
        }
        hr = nbg_combine(nbg_combine(hr, hash_str(f.first)), df);
        ++nr;
      }
      dv = nbg_combine(hr, nr);
    } else {
      dv = digest(e.second);
    }
    h = nbg_combine(nbg_combine(h, hash_str(e.first)), dv);
    ++nrest;
  }
  h = nrest ? nbg_combine(h, nrest) : 0;
  const uint64_t o1 = overflow_digest(ports, kMaxPorts), o2 = overflow_digest(env, kMaxEnv),
                 o3 = overflow_digest(mounts, kMaxMounts);
  if (o1 | o2 | o3) h = nbg_combine(nbg_combine(nbg_combine(h, o1), o2), o3);
  s[11] = h;
}

}  // namespace

uint64_t digest(const GVal& v) {
  switch (v.kind) {
    case GVal::Null: return K_NULL;
    case GVal::Bool: return nbg_combine(K_BOOL, (uint64_t)v.i);
    case GVal::Int: return nbg_combine(K_INT, (uint64_t)v.i);
    case GVal::Float: {
      double f = v.f == 0.0 ? 0.0 : v.f;
      uint64_t b; memcpy(&b, &f, 8);
      return nbg_combine(K_FLOAT, b);
    }
    case GVal::Str: return nbg_combine(K_STR, hash_str(v.s));
    case GVal::List: {
      uint64_t h = nbg_combine(K_LIST, v.items.size());
      for (const auto& e : v.items) h = nbg_combine(h, digest(e));
      return h;
    }
    case GVal::Map: case GVal::Struct: {
      uint64_t h = v.kind == GVal::Map ? K_MAP : K_STRUCT;
      size_t n = 0;
      for (const auto& e : v.kv) {
        if (v.kind == GVal::Struct && e.second.is_zero()) continue;
        h = nbg_combine(nbg_combine(h, hash_str(e.first)), digest(e.second));
        ++n;
      }
      return nbg_combine(h, n);
    }
    case GVal::Ptr: return nbg_combine(K_PTR, v.items.empty() ? K_MARK : digest(v.items[0]));
  }
  return 0;
}

uint64_t digest_string_map(const GVal& m) { return m.kind == GVal::Map ? digest(m) : K_NULL; }

void pack_kvmap(const GVal& m, std::vector<nbg_kv>* out) {
  if (m.kind != GVal::Map) return;
  const size_t at = out->size();
  for (const auto& e : m.kv) {
    nbg_kv kv;
    kv.key = hash_str(e.first);
    kv.val = e.second.kind == GVal::Str ? hash_str(e.second.s) : nbg_combine(K_MARK, digest(e.second));
    out->push_back(kv);
  }
  std::sort(out->begin() + at, out->end(), [](const nbg_kv& a, const nbg_kv& b) { return a.key < b.key; });
}

namespace {
constexpr uint64_t K_KEY = 0x6B65796B65796B65ull, K_FIELD = 0x6669656C64ull, K_INDEX = 0x696E646578ull;
void flatten_into(const GVal& v, uint64_t path, std::vector<nbg_kv>* out) {
  switch (v.kind) {
    case GVal::Struct:                              // a zero struct emits nothing: it IS the absent field
      for (const auto& e : v.kv)
        if (!e.second.is_zero()) flatten_into(e.second, nbg_combine(nbg_combine(path, K_FIELD), hash_str(e.first)), out);
      return;
    case GVal::Map:
      out->push_back({path, nbg_combine(K_MAP, v.kv.size())});          // non-nil, this many entries
      for (const auto& e : v.kv) flatten_into(e.second, nbg_combine(nbg_combine(path, K_KEY), hash_str(e.first)), out);
      return;
    case GVal::List:
      out->push_back({path, nbg_combine(K_LIST, v.items.size())});
      for (size_t k = 0; k < v.items.size(); ++k) {
        const uint64_t p = nbg_combine(nbg_combine(path, K_INDEX), k);
        // a zero struct element still occupies its position; the length marker covers it
        flatten_into(v.items[k], p, out);
      }
      return;
    case GVal::Ptr:
      out->push_back({path, K_PTR});                                     // non-nil pointer, whatever it points to
      if (!v.items.empty()) flatten_into(v.items[0], nbg_combine(path, K_PTR), out);
      return;
    default:
      out->push_back({path, digest(v)});                                 // typed leaf (Null included)
      return;
  }
}
}  // namespace

void flatten_tree(const GVal& v, std::vector<nbg_kv>* out) {
  const size_t at = out->size();
  // an absent root and a zero struct / zero leaf are the same Go value: neither emits anything
  if (v.is_zero()) return;
  flatten_into(v, 0x726F6F74726F6F74ull, out);
  std::sort(out->begin() + at, out->end(), [](const nbg_kv& a, const nbg_kv& b) { return a.key < b.key; });
}

GVal semantic_canon(const GVal& v) {
  if (v.tag == kTagQuantity) {
    std::string digits;
    int64_t exp;
    quantity_number(v, &digits, &exp);
    GVal q = GVal::strct();
    q.set("num", GVal::str(digits));
    q.kv.emplace_back("exp", GVal::integer(exp));      // kept even when 0: "num" alone identifies zero anyway
    std::sort(q.kv.begin(), q.kv.end(), [](const auto& a, const auto& b) { return a.first < b.first; });
    return q;
  }
  GVal o = v;
  for (auto& c : o.items) c = semantic_canon(c);
  for (auto& e : o.kv) e.second = semantic_canon(e.second);
  if ((o.kind == GVal::List && o.items.empty()) || (o.kind == GVal::Map && o.kv.empty())) return GVal();   // == nil
  if (o.kind == GVal::Struct) o.canon();
  return o;
}

void pack_podspec(const GVal& pod, uint64_t* s) {
  memset(s, 0, NBG_POD_SLOTS * sizeof(uint64_t));
  s[0] = str_of(pod, "serviceAccountName");
  s[1] = str_of(pod, "serviceAccount");
  s[2] = str_of(pod, "restartPolicy");
  s[3] = str_of(pod, "dnsPolicy");
  const GVal& tg = field(pod, "terminationGracePeriodSeconds");
  s[4] = tg.kind == GVal::Ptr && !tg.items.empty() ? (uint64_t)tg.items[0].i : 0;
  s[5] = str_of(pod, "schedulerName");
  s[6] = str_of(pod, "nodeName");
  const GVal& scp = field(pod, "securityContext");
  const GVal& sc = scp.kind == GVal::Ptr && !scp.items.empty() ? scp.items[0] : kZero;
  const GVal &fsg = field(sc, "fsGroup"), &rau = field(sc, "runAsUser");
  s[7] = fsg.kind == GVal::Ptr && !fsg.items.empty() ? (uint64_t)fsg.items[0].i : 0;
  s[8] = rau.kind == GVal::Ptr && !rau.items.empty() ? (uint64_t)rau.items[0].i : 0;
  s[9] = digest_without(sc, {"fsGroup", "runAsUser"});
  const GVal &nsel = field(pod, "nodeSelector"), &tol = field(pod, "tolerations"), &aff = field(pod, "affinity"),
             &ips = field(pod, "imagePullSecrets"), &vols = field(pod, "volumes"), &conts = field(pod, "containers");
  s[10] = nsel.is_null() ? 0 : digest(nsel);
  s[11] = tol.is_null() ? 0 : digest(tol);
  s[12] = aff.is_null() ? 0 : digest(aff);
  s[13] = ips.is_null() ? 0 : digest(ips);
  s[14] = (uint64_t)(tg.kind == GVal::Ptr) | ((uint64_t)(scp.kind == GVal::Ptr) << 1) |
          ((uint64_t)(fsg.kind == GVal::Ptr) << 2) | ((uint64_t)(rau.kind == GVal::Ptr) << 3) |
          (count_tag(vols) << 16) | (count_tag(conts) << 32);

  uint64_t* p = s + kPodLevel;
  for (int k = 0; k < kMaxVol; ++k, p += kVolSlots) {
    if (vols.kind != GVal::List || (size_t)k >= vols.items.size()) continue;
    const GVal& v = vols.items[k];
    p[0] = str_of(v, "name");
    // the usual sources: kind + the one string that names the backing object
    static const struct Known { const char* src; const char* ref; uint64_t src_hash; } known[] = {
        {"persistentVolumeClaim", "claimName", hash_str("persistentVolumeClaim")}, {"configMap", "name", hash_str("configMap")},
        {"secret", "secretName", hash_str("secret")}, {"emptyDir", "medium", hash_str("emptyDir")}};
    // p[1] != 0 says "this source, non-nil": when nothing else is set in it (the usual volume) the source leaves the
    // rest, so that slot 2 stays zero; otherwise the pointer stays in the rest, &T{...} with what is left of it.
    // The rest is digested in place - the value digest() gives for a copy of the volume with the name, the reference
    // and (when nothing else is in it) the source erased - without building that copy.
    const GVal* chosen = nullptr;
    const char* chosen_ref = nullptr;
    bool drop_source = false;
    if (v.kind == GVal::Struct)
      for (const auto& kn : known) {
        const GVal* src = v.find(kn.src);
        if (!src || src->kind != GVal::Ptr || src->items.empty() || src->items[0].kind != GVal::Struct) continue;
        p[1] = nbg_combine(kn.src_hash, str_of(src->items[0], kn.ref));
        chosen = src; chosen_ref = kn.ref;
        drop_source = true;
        for (const auto& f : src->items[0].kv)
          if (f.first != kn.ref && !f.second.is_zero()) { drop_source = false; break; }
        break;
      }
    size_t left = 0, counted = 0;
    uint64_t hr = K_STRUCT;
    if (v.kind == GVal::Struct)
      for (const auto& e : v.kv) {
        if (e.first == "name") continue;
        if (&e.second == chosen) {
          if (drop_source) continue;
          ++left; ++counted;
          hr = nbg_combine(nbg_combine(hr, hash_str(e.first)), nbg_combine(K_PTR, digest_without(chosen->items[0], {chosen_ref})));
          continue;
        }
        ++left;
        if (e.second.is_zero()) continue;
        hr = nbg_combine(nbg_combine(hr, hash_str(e.first)), digest(e.second));
        ++counted;
      }
    p[2] = left == 0 ? 0 : nbg_combine(hr, counted);
    // the usual volume - a name and one plain source - is ONE word; with anything else in it the three slots stay
    // apart and the third is non-zero, so the two spellings cannot meet
    if (p[2] == 0 && p[1] != 0) { p[0] = nbg_combine(p[0], p[1]); p[1] = 0; }
  }
  for (int k = 0; k < kMaxCont; ++k)
    if (conts.kind == GVal::List && (size_t)k < conts.items.size())
      pack_container(conts.items[k], s + kContBase + k * kContSlots);

  uint64_t h = digest_without(pod.kind == GVal::Struct ? pod : kZero,
                              {"serviceAccountName", "serviceAccount", "restartPolicy", "dnsPolicy",
                               "terminationGracePeriodSeconds", "schedulerName", "nodeName", "securityContext",
                               "nodeSelector", "tolerations", "affinity", "imagePullSecrets", "volumes", "containers"});
  if (pod.kind != GVal::Struct) h = 0;
  const uint64_t o1 = overflow_digest(vols, kMaxVol), o2 = overflow_digest(conts, kMaxCont);
  if (o1 | o2) h = nbg_combine(nbg_combine(h, o1), o2);
  s[15] = h;
}

// generateStatefulSet's edits of the PodSpec (NBCTL:495-521), applied to the Notebook's own tile: what the desired
// StatefulSet's tile is, without building the StatefulSet.  False when the tile cannot tell (env longer than the
// tile names, so an NB_PREFIX beyond it would go unseen) - the caller then builds and packs the object.
bool patch_statefulset_tile(uint64_t* s, uint64_t prefix_hash, bool add_fsgroup) {
  static const uint64_t kWorkingDir = hash_str("/home/jovyan"), kEnvName = hash_str("NB_PREFIX");
  static const struct DefaultPort { uint64_t name, rest; DefaultPort() {
    GVal port = GVal::strct();
    port.set("containerPort", GVal::integer(8888));
    port.set("name", GVal::str("notebook-port"));
    port.set("protocol", GVal::str("TCP"));
    uint64_t p[2];
    pack_named_element(port, p);
    name = p[0];
    rest = p[1];
  } } kPort;
  uint64_t* c = s + kContBase;                    // containers[0]
  uint64_t* ports = c + 12;
  uint64_t* env = ports + 2 * kMaxPorts;
  const uint64_t tags = c[6];
  uint64_t n_ports = tags & 0xFFFF, n_env = (tags >> 16) & 0xFFFF;
  const bool env_nil = n_env == 0xFFFF;
  const uint64_t have = env_nil ? 0 : n_env;
  if (have >= (uint64_t)kMaxEnv) return false;
  if (c[2] == 0) c[2] = kWorkingDir;
  if (n_ports == 0xFFFF) { ports[0] = kPort.name; ports[1] = kPort.rest; n_ports = 1; }
  bool there = false;
  for (uint64_t e = 0; e < have; ++e) there = there || env[2 * e] == kEnvName;
  if (!there) { env[2 * have] = kEnvName; env[2 * have + 1] = prefix_hash; n_env = have + 1; }
  c[6] = (tags & ~0xFFFFFFFFull) | n_ports | (n_env << 16);
  if (add_fsgroup && !(s[14] & 2)) {              // no securityContext: &PodSecurityContext{FSGroup: &100}
    s[7] = 100;
    s[9] = 0;                                     // nothing else in it
    s[14] |= 2 | 4;
  }
  return true;
}

void pack_service_ports(const GVal& ports, uint64_t* s) {
  // ports[0] field by field (generateService builds exactly one port, NBCTL:538-548, and that is what drifts);
  // every further port is folded into ONE digest: equal lists <=> equal tiles either way, in half the bytes.
  memset(s, 0, NBG_SVC_SLOTS * sizeof(uint64_t));
  if (ports.kind == GVal::List && !ports.items.empty()) {
    const GVal& sp = ports.items[0];
    s[0] = str_of(sp, "name");
    s[1] = str_of(sp, "protocol");
    const GVal& ap = field(sp, "appProtocol");
    s[2] = ap.kind == GVal::Ptr ? nbg_combine(K_PTR, ap.items.empty() ? 0 : hash_str(ap.items[0].s)) : 0;
    s[3] = (uint64_t)(uint32_t)field(sp, "port").i | ((uint64_t)(uint32_t)field(sp, "nodePort").i << 32);
    const GVal& tp = field(sp, "targetPort");
    s[4] = (uint64_t)(uint32_t)field(tp, "IntVal").i | ((uint64_t)field(tp, "Type").i << 32);
    // targetPort.StrVal and anything unexpected: the digest of the port without the fields slotted above, taken in
    // place (what digest() gives for a canonical copy with those fields erased; 0 when nothing is left)
    uint64_t hr = K_STRUCT;
    size_t counted = 0;
    if (sp.kind == GVal::Struct)
      for (const auto& e : sp.kv) {
        const std::string& f = e.first;
        if (f == "name" || f == "protocol" || f == "appProtocol" || f == "port" || f == "nodePort") continue;
        uint64_t d;
        if (f == "targetPort" && e.second.kind == GVal::Struct) {
          d = digest_without(e.second, {"IntVal", "Type"});
          if (d == 0) continue;                  // nothing but the slotted int and type
        } else {
          if (e.second.is_zero()) continue;
          d = digest(e.second);
        }
        hr = nbg_combine(nbg_combine(hr, hash_str(f)), d);
        ++counted;
      }
    s[5] = counted ? nbg_combine(hr, counted) : 0;
  }
  s[6] = ports.kind == GVal::List ? ports.items.size() : 0xFFFFFFFFull;
  s[7] = overflow_digest(ports, 1);
}

namespace {
// digest(v) with some nodes of the tree read as marker(): what digest() gives for a copy of v in which those nodes
// were replaced by a marker - the slotted leaves are taken OUT of the residual that way, the shape around them stays.
uint64_t digest_marked(const GVal& v, const GVal* const* marks, int n_marks) {
  for (int k = 0; k < n_marks; ++k)
    if (marks[k] == &v) return nbg_combine(K_PTR, K_MARK);       // == digest(marker())
  switch (v.kind) {
    case GVal::List: {
      uint64_t h = nbg_combine(K_LIST, v.items.size());
      for (const auto& e : v.items) h = nbg_combine(h, digest_marked(e, marks, n_marks));
      return h;
    }
    case GVal::Map: case GVal::Struct: {
      uint64_t h = v.kind == GVal::Map ? K_MAP : K_STRUCT;
      size_t n = 0;
      for (const auto& e : v.kv) {
        bool marked = false;
        for (int k = 0; k < n_marks; ++k) marked |= marks[k] == &e.second;
        if (v.kind == GVal::Struct && !marked && e.second.is_zero()) continue;
        h = nbg_combine(nbg_combine(h, hash_str(e.first)), digest_marked(e.second, marks, n_marks));
        ++n;
      }
      return nbg_combine(h, n);
    }
    case GVal::Ptr: return nbg_combine(K_PTR, v.items.empty() ? K_MARK : digest_marked(v.items[0], marks, n_marks));
    default: return digest(v);
  }
}
}  // namespace

bool pack_vs_spec(const GVal& obj, uint64_t* s, std::vector<nbg_kv>* headers, bool* headers_nil) {
  memset(s, 0, NBG_VS_SLOTS * sizeof(uint64_t));
  *headers_nil = true;
  // unstructured.NestedMap(obj, "spec"): found only if present, ok only if a map (COMMON:200,208)
  const GVal* specp = obj.kind == GVal::Map ? obj.find("spec") : nullptr;
  if (!specp || specp->kind != GVal::Map) return false;
  const GVal& spec = *specp;
  const GVal* marks[8];                  // leaves with a slot of their own: read as a marker by the residual digest
  int n_marks = 0;

  auto take_str = [&](const GVal* holder, uint64_t* slot) {
    if (holder && holder->kind == GVal::Str) { *slot = hash_str(holder->s); marks[n_marks++] = holder; }
  };
  auto first = [](const GVal* l) -> const GVal* { return (l && l->kind == GVal::List && !l->items.empty()) ? &l->items[0] : nullptr; };
  auto child = [](const GVal* m, const char* k) -> const GVal* { return (m && m->kind == GVal::Map) ? m->find(k) : nullptr; };

  take_str(first(child(&spec, "hosts")), &s[0]);
  take_str(first(child(&spec, "gateways")), &s[1]);
  const GVal* http0 = first(child(&spec, "http"));
  take_str(child(child(first(child(http0, "match")), "uri"), "prefix"), &s[2]);
  take_str(child(child(http0, "rewrite"), "uri"), &s[3]);
  const GVal* dest = child(first(child(http0, "route")), "destination");
  take_str(child(dest, "host"), &s[4]);
  if (const GVal* num = child(child(dest, "port"), "number")) {
    if (num->kind == GVal::Int) { s[5] = (uint64_t)num->i; s[6] = 1; marks[n_marks++] = num; }
    else if (num->kind == GVal::Float) { memcpy(&s[5], &num->f, 8); if (num->f == 0.0) s[5] = 0; s[6] = 2; marks[n_marks++] = num; }
  }
  if (const GVal* set = child(child(child(http0, "headers"), "request"), "set")) {
    bool all_str = set->kind == GVal::Map;
    if (all_str) for (const auto& e : set->kv) if (e.second.kind != GVal::Str) all_str = false;
    if (all_str) { pack_kvmap(*set, headers); *headers_nil = false; marks[n_marks++] = set; }
  }
  s[7] = digest_marked(spec, marks, n_marks);
  return true;
}

}  // namespace nbh
