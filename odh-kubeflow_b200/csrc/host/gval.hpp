// gval.hpp — a Go value, as far as reflect.DeepEqual can tell values apart.
//
// The reference's predicates receive fully materialised Go objects
// (appsv1.StatefulSet, corev1.Service, unstructured.Unstructured) and compare
// them with reflect.DeepEqual (COMMON:128,184,189,214).  DeepEqual's verdict
// depends only on: kind, nil-ness of pointers/slices/maps/interfaces, lengths,
// and leaf values.  GVal keeps exactly that and nothing else:
//
//   Null    nil pointer / nil slice / nil map / nil interface
//   Bool Int Float Str        leaves (Int = any Go integer type, Float = float64)
//   List    non-nil slice (possibly empty)
//   Map     non-nil map with string keys (possibly empty), entries sorted by key
//   Struct  struct value; fields sorted by name; a field holding its Go zero
//           value is NOT stored (so "absent" == "zero", as in Go)
//   Ptr     non-nil pointer; items[0] is the pointee
//
// Field names are the Kubernetes JSON names.  Inlined Go structs (VolumeSource
// in Volume, ProbeHandler in Probe) are flattened, which cannot change a
// DeepEqual verdict because both sides are flattened alike.
#pragma once
#include <algorithm>
#include <cstdint>
#include <cstring>
#include <string>
#include <string_view>
#include <utility>
#include <vector>

namespace nbh {

constexpr uint8_t kTagQuantity = 1;

struct GVal {
  enum Kind : uint8_t { Null, Bool, Int, Float, Str, List, Map, Struct, Ptr };
  Kind kind = Null;
  uint8_t tag = 0; // kTagQuantity marks a resource.Quantity struct (semantic equality compares it by value)
  int64_t i = 0;   // Bool, Int
  double f = 0.0;  // Float
  std::string s;   // Str
  std::vector<std::pair<std::string, GVal>> kv;  // Map entries / Struct fields, sorted by name
  std::vector<GVal> items;                       // List elements; Ptr pointee

  // ---- constructors
  static GVal null() { return GVal(); }
  static GVal boolean(bool b) { GVal g; g.kind = Bool; g.i = b ? 1 : 0; return g; }
  static GVal integer(int64_t v) { GVal g; g.kind = Int; g.i = v; return g; }
  static GVal real(double v) { GVal g; g.kind = Float; g.f = v; return g; }
  static GVal str(std::string v) { GVal g; g.kind = Str; g.s = std::move(v); return g; }
  static GVal list() { GVal g; g.kind = List; return g; }
  static GVal map() { GVal g; g.kind = Map; return g; }
  static GVal strct() { GVal g; g.kind = Struct; return g; }
  static GVal ptr(GVal pointee) { GVal g; g.kind = Ptr; g.items.push_back(std::move(pointee)); return g; }

  bool is_null() const { return kind == Null; }

  // Go zero value?  (non-nil empty slices/maps and non-nil pointers are NOT zero)
  bool is_zero() const {
    switch (kind) {
      case Null: return true;
      case Bool: case Int: return i == 0;
      case Float: return f == 0.0;
      case Str: return s.empty();
      case Struct:
        for (const auto& e : kv) if (!e.second.is_zero()) return false;
        return true;
      default: return false;
    }
  }

  // ---- Map / Struct access (sorted by name)
  // (string_view: looking a field up by a literal name builds no temporary std::string)
  const GVal* find(std::string_view name) const {
    if (kv.size() <= 32) {
      // API objects have a handful of fields: a scan that looks at lengths first beats a binary search whose every
      // step is a lexicographic compare (a quarter of the packer's time went into those)
      for (const auto& e : kv)
        if (e.first.size() == name.size() && memcmp(e.first.data(), name.data(), name.size()) == 0) return &e.second;
      return nullptr;
    }
    auto it = std::lower_bound(kv.begin(), kv.end(), name,
                               [](const std::pair<std::string, GVal>& e, std::string_view n) { return std::string_view(e.first) < n; });
    return (it != kv.end() && std::string_view(it->first) == name) ? &it->second : nullptr;
  }
  GVal* find(std::string_view name) { return const_cast<GVal*>(static_cast<const GVal*>(this)->find(name)); }
  const GVal* find(const std::string& name) const { return find(std::string_view(name)); }
  GVal* find(const std::string& name) { return find(std::string_view(name)); }
  const GVal* find(const char* name) const { return find(std::string_view(name)); }
  GVal* find(const char* name) { return find(std::string_view(name)); }
  bool has(const std::string& name) const { return find(name) != nullptr; }

  // Insert or overwrite.  Struct fields holding a zero value are erased instead.
  void set(const std::string& name, GVal v) {
    auto it = std::lower_bound(kv.begin(), kv.end(), name,
                               [](const std::pair<std::string, GVal>& e, const std::string& n) { return e.first < n; });
    const bool present = it != kv.end() && it->first == name;
    if (kind == Struct && v.is_zero()) {
      if (present) kv.erase(it);
      return;
    }
    if (present) it->second = std::move(v);
    else kv.insert(it, std::make_pair(name, std::move(v)));
  }
  void erase(const std::string& name) {
    auto it = std::lower_bound(kv.begin(), kv.end(), name,
                               [](const std::pair<std::string, GVal>& e, const std::string& n) { return e.first < n; });
    if (it != kv.end() && it->first == name) kv.erase(it);
  }
  // Field of a struct as a value (zero GVal if absent).
  GVal get(const std::string& name) const { const GVal* p = find(name); return p ? *p : GVal(); }
  // String leaf of a struct field / map entry ("" if absent).
  std::string get_str(const std::string& name) const {
    const GVal* p = find(name);
    return (p && p->kind == Str) ? p->s : std::string();
  }

  // Drop zero-valued struct fields everywhere (after in-place edits).
  void canon() {
    for (auto& c : items) c.canon();
    for (auto& e : kv) e.second.canon();
    if (kind == Struct)
      kv.erase(std::remove_if(kv.begin(), kv.end(),
                              [](const std::pair<std::string, GVal>& e) { return e.second.is_zero(); }),
               kv.end());
  }
};

// map[string]string helpers (labels, annotations, selector)
inline GVal string_map(const std::vector<std::pair<std::string, std::string>>& kvs) {
  GVal m = GVal::map();
  for (const auto& e : kvs) m.set(e.first, GVal::str(e.second));
  return m;
}

}  // namespace nbh
