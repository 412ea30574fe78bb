// oracle_semantic.cpp — L1 "semantic" oracle: a CPU restatement of the
// reference's Go predicates on whole objects and strings.
//
// TEST INFRASTRUCTURE ONLY.  Nothing under odh-kubeflow_b200/ may include, link
// or call this file; only tests/, __graft_entry__.smoke() and bench.py's
// cpu_baseline / --impl reference legs use it, and only as the checker.
//
// What it follows, line by line (paths under /root/reference/components):
//   common/reconcilehelper/util.go:105-134   CopyStatefulSetFields
//   common/reconcilehelper/util.go:165-195   CopyServiceFields
//   common/reconcilehelper/util.go:197-219   CopyVirtualService
//   notebook-controller/controllers/notebook_controller.go:188-194  template-label pre-step
//   notebook-controller/controllers/culling_controller.go:86-203    Reconcile decision flow
//   notebook-controller/controllers/culling_controller.go:207-241   cullingCheckPeriodHasPassed, notebookIsIdle
//   notebook-controller/controllers/culling_controller.go:315-438   allKernelsAreIdle ... updateLastCullingCheck...
//   notebook-controller/controllers/culling_controller.go:440-502   annotationsExist ... StopAnnotationIsSet
//
// Arithmetic that lives in dependencies ABSENT from /root/reference (restated
// from their published behaviour):
//   reflect.DeepEqual, time.Parse/Format/After/Before/Add — Go stdlib, toolchain
//     go 1.25.7 (notebook-controller/go.mod:3)
//   unstructured.NestedMap / SetNestedMap — k8s.io/apimachinery v0.33.7
//     (notebook-controller/go.mod:14), pkg/apis/meta/v1/unstructured/helpers.go
//
// PARITY PINNING.  The reference cannot be compiled or run here (Go, no
// toolchain, un-vendored modules), so there is no oracle/_ref.  This oracle is
// pinned against every known-answer case the reference's own tests hold for
// the path (tests/golden/reference_cases.json, taken from
// culling_controller_test.go:58-263 and notebook_controller_bdd_test.go:49-90).
// Those pin C1 (allKernelsAreIdle), C7 (notebookIsIdle), C9 and the label copy.
// For everything else — reflect.DeepEqual on PodSpec/Service/VS trees (no
// util_test.go exists), C2, C3, C8 — PARITY IS UNPINNED: expected values of the
// extra cases in tests/golden/derived_cases.json are derived by hand from the
// cited lines.
//
// Objects are nbh::GVal trees (odh-kubeflow_b200/csrc/host/gval.hpp: the data
// model only — Go kinds, nil-ness, lengths, leaves).  The algorithms here are
// this file's own.
#include <atomic>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <map>
#include <string>
#include <thread>
#include <vector>

#include "../odh-kubeflow_b200/csrc/host/gval.hpp"
#include "../odh-kubeflow_b200/csrc/host/json.hpp"

using nbh::GVal;

#define ORC_API extern "C" __attribute__((visibility("default")))

struct nbh_obj { GVal v; };  // same layout as libnbgraft_host.so's handle

namespace {

// ------------------------------------------------------- reflect.DeepEqual
// Go's deepValueEqual, by Kind (reflect/deepequal.go): kinds must match; nil
// only equals nil; slices/maps compare lengths then elements; structs field by
// field (a field absent from a GVal struct holds its zero value).
bool deep_equal(const GVal& a, const GVal& b);

bool struct_equal(const GVal& a, const GVal& b) {
  // walk the union of field names; an absent field is the zero value
  size_t i = 0, j = 0;
  static const GVal zero;
  while (i < a.kv.size() || j < b.kv.size()) {
    int c;
    if (i >= a.kv.size()) c = 1;
    else if (j >= b.kv.size()) c = -1;
    else c = a.kv[i].first.compare(b.kv[j].first);
    if (c == 0) { if (!deep_equal(a.kv[i].second, b.kv[j].second)) return false; ++i; ++j; }
    else if (c < 0) { if (!a.kv[i].second.is_zero()) return false; ++i; }
    else { if (!b.kv[j].second.is_zero()) return false; ++j; }
  }
  return true;
}

bool deep_equal(const GVal& a, const GVal& b) {
  // a zero struct and an absent (Null) struct field are the same Go value
  if (a.kind == GVal::Struct && b.kind == GVal::Null) return a.is_zero();
  if (a.kind == GVal::Null && b.kind == GVal::Struct) return b.is_zero();
  if (a.kind != b.kind) {
    // a zero leaf against an absent one (only arises for struct fields)
    if (a.kind == GVal::Null) return b.is_zero();
    if (b.kind == GVal::Null) return a.is_zero();
    return false;
  }
  switch (a.kind) {
    case GVal::Null: return true;
    case GVal::Bool: case GVal::Int: return a.i == b.i;
    case GVal::Float: return a.f == b.f;
    case GVal::Str: return a.s == b.s;
    case GVal::List:
      if (a.items.size() != b.items.size()) return false;
      for (size_t k = 0; k < a.items.size(); ++k)
        if (!deep_equal(a.items[k], b.items[k])) return false;
      return true;
    case GVal::Map:
      if (a.kv.size() != b.kv.size()) return false;
      for (size_t k = 0; k < a.kv.size(); ++k) {
        const GVal* o = b.find(a.kv[k].first);
        if (!o) return false;
        // map VALUES are full Go values: nil vs "" etc. stay distinct under interface{}
        if (a.kv[k].second.kind != o->kind) return false;
        if (!deep_equal(a.kv[k].second, *o)) return false;
      }
      return true;
    case GVal::Struct: return struct_equal(a, b);
    case GVal::Ptr:
      if (a.items.empty() || b.items.empty()) return a.items.empty() == b.items.empty();
      return deep_equal(a.items[0], b.items[0]);
  }
  return false;
}

// List elements under interface{} ([]interface{}) must also match in kind.
// deep_equal() above lets Null equal a zero leaf, which is right for struct
// fields; strict_equal is the entry point for untyped (unstructured) trees.
bool strict_equal(const GVal& a, const GVal& b) {
  if (a.kind != b.kind) return false;
  switch (a.kind) {
    case GVal::List:
      if (a.items.size() != b.items.size()) return false;
      for (size_t k = 0; k < a.items.size(); ++k)
        if (!strict_equal(a.items[k], b.items[k])) return false;
      return true;
    case GVal::Map:
      if (a.kv.size() != b.kv.size()) return false;
      for (size_t k = 0; k < a.kv.size(); ++k) {
        const GVal* o = b.find(a.kv[k].first);
        if (!o || !strict_equal(a.kv[k].second, *o)) return false;
      }
      return true;
    default: return deep_equal(a, b);
  }
}

const GVal kZero;
const GVal& field(const GVal& s, const char* name) {
  const GVal* p = (s.kind == GVal::Struct || s.kind == GVal::Map) ? s.find(name) : nullptr;
  return p ? *p : kZero;
}
GVal* mfield(GVal& s, const char* name) {
  return (s.kind == GVal::Struct || s.kind == GVal::Map) ? s.find(name) : nullptr;
}
// mutable struct field, created when absent (kept even while still zero)
GVal& ensure(GVal& s, const char* name, GVal init) {
  if (!s.find(name)) {
    auto it = std::lower_bound(s.kv.begin(), s.kv.end(), std::string(name),
                               [](const std::pair<std::string, GVal>& e, const std::string& n) { return e.first < n; });
    s.kv.insert(it, std::make_pair(std::string(name), std::move(init)));
  }
  return *s.find(name);
}

// `for k, v := range to { if from[k] != v {...} }` on map[string]string where
// either map may be nil (ranging over nil is empty; indexing nil yields "").
bool live_entry_differs(const GVal& from, const GVal& to) {
  bool differs = false;
  if (to.kind != GVal::Map) return false;
  for (const auto& e : to.kv) {
    const GVal* f = from.kind == GVal::Map ? from.find(e.first) : nullptr;
    const std::string fv = (f && f->kind == GVal::Str) ? f->s : std::string();
    if (fv != e.second.s) differs = true;
  }
  return differs;
}

enum { R_LABELS = 1, R_ANN = 2, R_THIRD = 4, R_FOURTH = 8, R_TMPL = 16, R_NILREPLICAS = 0x80 };

void assign_field(GVal& obj, const char* name, const GVal& value) {
  if (obj.kind != GVal::Struct) obj = GVal::strct();
  if (value.is_zero()) obj.erase(name); else obj.set(name, value);
}

// util.go:105-134 (+ notebook_controller.go:188-194 when prestep != 0)
int copy_statefulset_fields(const GVal& from, GVal& to, int prestep) {
  int reason = 0;
  // create every struct we will hold a reference to BEFORE taking references
  // (inserting a field may move its siblings)
  ensure(to, "metadata", GVal::strct());
  ensure(to, "spec", GVal::strct());
  const GVal& from_meta = field(from, "metadata");
  GVal& to_meta = *to.find("metadata");
  GVal& to_spec = *to.find("spec");
  ensure(to_spec, "template", GVal::strct());
  ensure(*to_spec.find("template"), "metadata", GVal::strct());
  const GVal& from_rep = field(field(from, "spec"), "replicas");
  GVal* to_rep = mfield(to_spec, "replicas");
  const bool from_nil = from_rep.kind != GVal::Ptr, to_nil = !to_rep || to_rep->kind != GVal::Ptr;

  if (prestep && !from_nil && !to_nil && from_rep.items[0].i != to_rep->items[0].i) {   // :190
    const GVal& fl = field(field(field(field(from, "spec"), "template"), "metadata"), "labels");
    GVal& tt = *to_spec.find("template")->find("metadata");
    const GVal tl = field(tt, "labels");
    if (!(fl.kind == tl.kind && deep_equal(fl, tl))) {                                 // :191
      assign_field(tt, "labels", fl);
      reason |= R_TMPL;
    }
  }

  if (live_entry_differs(field(from_meta, "labels"), field(to_meta, "labels"))) reason |= R_LABELS;  // :109-113
  assign_field(to_meta, "labels", field(from_meta, "labels"));                                        // :114
  if (live_entry_differs(field(from_meta, "annotations"), field(to_meta, "annotations"))) reason |= R_ANN;  // :116-120
  assign_field(to_meta, "annotations", field(from_meta, "annotations"));                              // :121

  if (from_nil || to_nil) {
    reason |= R_NILREPLICAS;                         // nil pointer dereference at :123
  } else if (from_rep.items[0].i != to_rep->items[0].i) {
    to_rep->items[0].i = from_rep.items[0].i;        // :124
    reason |= R_THIRD;
  }

  const GVal& from_pod = field(field(field(from, "spec"), "template"), "spec");
  GVal& to_tmpl = *to_spec.find("template");
  if (!deep_equal(field(to_tmpl, "spec"), from_pod)) reason |= R_FOURTH;   // :128
  assign_field(to_tmpl, "spec", from_pod);                                 // :131
  return reason;
}

// map[string]string / []ServicePort DeepEqual incl. nil-vs-empty
bool typed_equal(const GVal& a, const GVal& b) {
  if ((a.kind == GVal::Null) != (b.kind == GVal::Null)) return false;
  return deep_equal(a, b);
}

// util.go:165-195
int copy_service_fields(const GVal& from, GVal& to) {
  int reason = 0;
  ensure(to, "metadata", GVal::strct());
  ensure(to, "spec", GVal::strct());
  const GVal& from_meta = field(from, "metadata");
  GVal& to_meta = *to.find("metadata");
  if (live_entry_differs(field(from_meta, "labels"), field(to_meta, "labels"))) reason |= R_LABELS;
  assign_field(to_meta, "labels", field(from_meta, "labels"));
  if (live_entry_differs(field(from_meta, "annotations"), field(to_meta, "annotations"))) reason |= R_ANN;
  assign_field(to_meta, "annotations", field(from_meta, "annotations"));
  const GVal& from_spec = field(from, "spec");
  GVal& to_spec = *to.find("spec");
  if (!typed_equal(field(to_spec, "selector"), field(from_spec, "selector"))) reason |= R_THIRD;   // :184
  assign_field(to_spec, "selector", field(from_spec, "selector"));
  if (!typed_equal(field(to_spec, "ports"), field(from_spec, "ports"))) reason |= R_FOURTH;        // :189
  assign_field(to_spec, "ports", field(from_spec, "ports"));
  return reason;
}

// unstructured.NestedMap(obj, "spec"): (deep copy, found, err).  found=false if
// the key is absent; err != nil if the value is not a map[string]interface{}
// (a JSON null value is "found" but not a map => error).
void nested_map(const GVal& obj, const char* key, const GVal** out, bool* found, bool* err) {
  *out = nullptr; *found = false; *err = false;
  const GVal* v = obj.kind == GVal::Map ? obj.find(key) : nullptr;
  if (!v) return;
  *found = true;
  if (v->kind != GVal::Map) { *err = true; return; }
  *out = v;
}

// util.go:197-219.  reason: 1 differs, 2 to.spec missing, 4 from.spec missing
int copy_virtualservice(const GVal& from, GVal& to) {
  const GVal* from_spec; bool found, err;
  nested_map(from, "spec", &from_spec, &found, &err);
  if (!found) return 4;                 // :201-203
  if (err) return 4;                    // :204-206
  const GVal* to_spec;
  nested_map(to, "spec", &to_spec, &found, &err);
  if (!found || err) {                  // :209-212
    if (to.kind != GVal::Map) to = GVal::map();
    to.set("spec", *from_spec);
    return 2;
  }
  const bool differs = !strict_equal(*from_spec, *to_spec);   // :214
  if (differs) to.set("spec", *from_spec);
  return differs ? 1 : 0;
}

// --------------------------------------------------------------- Go `time`
struct GoTime {
  int64_t sec = -62135596800ll;  // zero Time: 0001-01-01T00:00:00Z
  int32_t nsec = 0;
  int32_t offset = 0;
};
bool go_after(const GoTime& a, const GoTime& b) { return a.sec > b.sec || (a.sec == b.sec && a.nsec > b.nsec); }
bool go_before(const GoTime& a, const GoTime& b) { return go_after(b, a); }
GoTime go_add_minutes(GoTime t, int64_t minutes) { t.sec += minutes * 60; return t; }

bool leap(int y) { return y % 4 == 0 && (y % 100 != 0 || y % 400 == 0); }
int days_in_month(int m, int y) {
  switch (m) { case 2: return leap(y) ? 29 : 28; case 4: case 6: case 9: case 11: return 30; default: return 31; }
}
// absolute day count by plain loops over 400/100/4/1-year cycles (deliberately
// not the closed form the product's parser uses)
int64_t days_since_epoch(int y, int m, int d) {
  int64_t days = 0;
  // whole years from 1970
  if (y >= 1970) for (int yy = 1970; yy < y; ++yy) days += leap(yy) ? 366 : 365;
  else for (int yy = y; yy < 1970; ++yy) days -= leap(yy) ? 366 : 365;
  for (int mm = 1; mm < m; ++mm) days += days_in_month(mm, y);
  return days + d - 1;
}

// time.Parse(time.RFC3339, value): the layout "2006-01-02T15:04:05Z07:00"
// consumed chunk by chunk as Go's Parse does.
struct Cursor {
  const char* p; size_t n;
  bool digit(size_t k) const { return k < n && p[k] >= '0' && p[k] <= '9'; }
  void skip(size_t k) { p += k; n -= k; }
};
// getnum(s, fixed): one or two digits; fixed requires two
bool getnum(Cursor& c, bool fixed, int* out) {
  if (!c.digit(0)) return false;
  if (!c.digit(1)) {
    if (fixed) return false;
    *out = c.p[0] - '0'; c.skip(1); return true;
  }
  *out = (c.p[0] - '0') * 10 + (c.p[1] - '0'); c.skip(2); return true;
}
bool literal(Cursor& c, char ch) { if (c.n == 0 || c.p[0] != ch) return false; c.skip(1); return true; }

bool go_parse_rfc3339(const std::string& value, GoTime* out) {
  Cursor c{value.data(), value.size()};
  int year = 0, month = 0, day = 0, hour = 0, min = 0, sec = 0;
  int64_t nsec = 0;
  // stdLongYear: exactly four digits
  if (c.n < 4 || !c.digit(0) || !c.digit(1) || !c.digit(2) || !c.digit(3)) return false;
  for (int k = 0; k < 4; ++k) year = year * 10 + (c.p[k] - '0');
  c.skip(4);
  if (!literal(c, '-')) return false;
  if (!getnum(c, true, &month) || month < 1 || month > 12) return false;     // stdZeroMonth
  if (!literal(c, '-')) return false;
  if (!getnum(c, true, &day)) return false;                                   // stdZeroDay (range checked last)
  if (!literal(c, 'T')) return false;
  if (!getnum(c, false, &hour) || hour >= 24) return false;                   // stdHour
  if (!literal(c, ':')) return false;
  if (!getnum(c, true, &min) || min >= 60) return false;                      // stdZeroMinute
  if (!literal(c, ':')) return false;
  if (!getnum(c, true, &sec) || sec >= 60) return false;                      // stdZeroSecond
  // "fractional second in the input but not in the layout"
  if (c.n >= 2 && (c.p[0] == '.' || c.p[0] == ',') && c.digit(1)) {
    size_t k = 2;
    while (c.digit(k)) ++k;
    // parseNanoseconds(value, k): at most nine digits count
    size_t use = k > 10 ? 10 : k;
    for (size_t q = 1; q < use; ++q) nsec = nsec * 10 + (c.p[q] - '0');
    for (size_t q = use; q < 10; ++q) nsec *= 10;
    c.skip(k);
  }
  int offset = 0;
  // stdISO8601ColonTZ
  if (c.n >= 1 && c.p[0] == 'Z') {
    c.skip(1);
  } else {
    if (c.n < 6) return false;
    if (c.p[3] != ':') return false;
    const char sign = c.p[0];
    Cursor hh{c.p + 1, 2}, mm{c.p + 4, 2};
    int h = 0, m = 0;
    if (!getnum(hh, true, &h) || !getnum(mm, true, &m)) return false;
    if (h > 24 || m > 60) return false;     // "> rather than >=": 24:00 and :60 are accepted
    offset = (h * 60 + m) * 60;
    if (sign == '-') offset = -offset; else if (sign != '+') return false;
    c.skip(6);
  }
  if (c.n != 0) return false;               // extra text
  if (day < 1 || day > days_in_month(month, year)) return false;  // "day out of range"
  out->sec = days_since_epoch(year, month, day) * 86400ll + hour * 3600 + min * 60 + sec - offset;
  out->nsec = (int32_t)nsec;
  out->offset = offset;
  return true;
}

// t.Format(time.RFC3339)
std::string go_format_rfc3339(const GoTime& t) {
  int64_t local = t.sec + t.offset;
  int64_t days = local / 86400, rem = local % 86400;
  if (rem < 0) { rem += 86400; days -= 1; }
  int y = 1970;
  while (days < 0) { --y; days += leap(y) ? 366 : 365; }
  while (days >= (leap(y) ? 366 : 365)) { days -= leap(y) ? 366 : 365; ++y; }
  int m = 1;
  while (days >= days_in_month(m, y)) { days -= days_in_month(m, y); ++m; }
  char buf[64];
  int k = snprintf(buf, sizeof buf, "%04d-%02d-%02dT%02d:%02d:%02d", y, m, (int)days + 1, (int)(rem / 3600),
                   (int)(rem / 60 % 60), (int)(rem % 60));
  if (t.offset == 0) snprintf(buf + k, sizeof buf - k, "Z");
  else {
    const int a = t.offset < 0 ? -t.offset : t.offset;
    snprintf(buf + k, sizeof buf - k, "%c%02d:%02d", t.offset < 0 ? '-' : '+', a / 3600, a / 60 % 60);
  }
  return buf;
}

// ------------------------------------------------------------------ culler
const char* STOP_ANNOTATION = "kubeflow-resource-stopped";
const char* LAST_ACTIVITY_ANNOTATION = "notebooks.kubeflow.org/last-activity";
const char* LAST_ACTIVITY_CHECK_TIMESTAMP_ANNOTATION = "notebooks.kubeflow.org/last_activity_check_timestamp";

struct Meta {
  bool has = false;                           // Annotations != nil
  std::map<std::string, std::string> ann;
  std::string get(const char* k) const { auto it = ann.find(k); return it == ann.end() ? std::string() : it->second; }
  bool present(const char* k) const { return has && ann.count(k) != 0; }
};
struct Kernel { std::string last_activity, execution_state; };
struct Env {
  GoTime now;             // time.Now()
  int cull_idle_time;     // CULL_IDLE_TIME (minutes)
  int check_period;       // IDLENESS_CHECK_PERIOD (minutes)
};
std::string create_timestamp(const Env& e) {  // culling_controller.go:512-515
  GoTime t = e.now; t.nsec = 0; return go_format_rfc3339(t);
}

bool stop_annotation_is_set(const Meta& m) {          // :493-502
  if (!m.has) return false;
  return m.ann.count(STOP_ANNOTATION) != 0;
}
bool culling_check_period_has_passed(const Meta& m, const Env& e) {   // :207-217
  if (!m.present(LAST_ACTIVITY_CHECK_TIMESTAMP_ANNOTATION)) return false;
  GoTime stored;                                                       // parse error ignored => zero Time
  GoTime parsed;
  if (go_parse_rfc3339(m.get(LAST_ACTIVITY_CHECK_TIMESTAMP_ANNOTATION), &parsed)) stored = parsed;
  const GoTime next = go_add_minutes(stored, e.check_period);
  return go_before(next, e.now);
}
bool notebook_is_idle(const Meta& m, const Env& e) {                   // :220-241
  if (m.has) {
    if (stop_annotation_is_set(m)) return false;
    GoTime last;
    if (!go_parse_rfc3339(m.get(LAST_ACTIVITY_ANNOTATION), &last)) return false;
    const GoTime cap = go_add_minutes(last, e.cull_idle_time);
    if (go_after(e.now, cap)) return true;
  }
  return false;
}
bool all_kernels_are_idle(const std::vector<Kernel>& ks) {            // :315-327
  for (size_t i = 0; i < ks.size(); ++i)
    if (ks[i].execution_state != "idle") return false;
  return true;
}
std::string get_notebook_recent_time(const std::vector<std::string>& t, size_t* which) {   // :329-349
  GoTime recent;
  if (!go_parse_rfc3339(t[0], &recent)) return "";
  *which = 0;
  for (size_t i = 1; i < t.size(); ++i) {
    GoTime cur;
    if (!go_parse_rfc3339(t[i], &cur)) return "";
    if (go_after(cur, recent)) { recent = cur; *which = i; }
  }
  return go_format_rfc3339(recent);
}
bool compare_annotation_time_to_resource(const Meta& m, const std::string& resource_time) {  // :351-369
  GoTime a, t;
  if (!go_parse_rfc3339(m.get(LAST_ACTIVITY_ANNOTATION), &a)) return false;
  if (!go_parse_rfc3339(resource_time, &t)) return false;
  if (go_after(a, t)) return false;
  return true;
}
struct Trace { int src = 0; uint32_t argmax = 0xFFFFFFFFu; };
bool update_from_kernels(Meta& m, const std::vector<Kernel>& ks, const Env& e, Trace* tr) {   // :371-402
  if (ks.empty()) return false;
  if (!all_kernels_are_idle(ks)) {
    m.ann[LAST_ACTIVITY_ANNOTATION] = create_timestamp(e); m.has = true;
    tr->src = 1; tr->argmax = 0xFFFFFFFFu;
    return false;
  }
  std::vector<std::string> arr;
  for (const auto& k : ks) arr.push_back(k.last_activity);
  size_t which = 0;
  const std::string t = get_notebook_recent_time(arr, &which);
  if (t.empty() || !compare_annotation_time_to_resource(m, t)) return false;
  m.ann[LAST_ACTIVITY_ANNOTATION] = t;
  tr->src = 2; tr->argmax = (uint32_t)which;
  return true;
}
bool update_from_terminals(Meta& m, const std::vector<Kernel>& ts, const Env&, Trace* tr) {   // :404-428
  if (ts.empty()) return false;
  std::vector<std::string> arr;
  for (const auto& k : ts) arr.push_back(k.last_activity);
  size_t which = 0;
  const std::string t = get_notebook_recent_time(arr, &which);
  if (t.empty() || !compare_annotation_time_to_resource(m, t)) return false;
  m.ann[LAST_ACTIVITY_ANNOTATION] = t;
  tr->src = 3; tr->argmax = (uint32_t)which;
  return true;
}
bool annotations_exist(const Meta& m) {                                // :440-447
  return m.present(LAST_ACTIVITY_ANNOTATION) && m.present(LAST_ACTIVITY_CHECK_TIMESTAMP_ANNOTATION);
}
void initialize_annotations(Meta& m, const Env& e) {                   // :449-456
  m.has = true;
  const std::string t = create_timestamp(e);
  m.ann[LAST_ACTIVITY_ANNOTATION] = t;
  m.ann[LAST_ACTIVITY_CHECK_TIMESTAMP_ANNOTATION] = t;
}

bool parse_cull_input(const std::string& text, Meta* m, std::vector<Kernel>* ks, std::vector<Kernel>* ts, bool* pod_missing) {
  GVal j; std::string err;
  if (!nbh::parse_json(text, &j, &err) || j.kind != GVal::Map) return false;
  const GVal* ann = j.find("annotations");
  m->has = ann && ann->kind == GVal::Map;
  m->ann.clear();
  if (m->has) for (const auto& e : ann->kv) m->ann[e.first] = e.second.s;
  *pod_missing = false;
  if (const GVal* p = j.find("pod_missing")) *pod_missing = p->kind == GVal::Bool && p->i;
  for (int which = 0; which < 2; ++which) {
    const GVal* l = j.find(which ? "terminals" : "kernels");
    std::vector<Kernel>* out = which ? ts : ks;
    out->clear();
    if (!l || l->kind != GVal::List) continue;
    for (const auto& e : l->items) {
      Kernel k;
      if (e.kind == GVal::Map) { k.last_activity = e.get_str("last_activity"); k.execution_state = e.get_str("execution_state"); }
      out->push_back(k);
    }
  }
  return true;
}

enum { ACT_PERIOD = 4, ACT_CULL = 8, ACT_STOPPED = 16, ACT_INIT = 32, ACT_NOPOD = 64 };

// mode 0: every predicate evaluated unconditionally; mode 1: the flow of
// CullingReconciler.Reconcile, culling_controller.go:86-203
int cull_decide(Meta& m, const std::vector<Kernel>& ks, const std::vector<Kernel>& ts, bool pod_missing,
                const Env& e, int mode, uint32_t* argmax, std::string* new_la) {
  int act = 0;
  Trace tr;
  *new_la = "";
  if (mode == 1) {
    if (stop_annotation_is_set(m)) { *argmax = 0xFFFFFFFFu; return ACT_STOPPED; }   // :104-117
    if (pod_missing) { *argmax = 0xFFFFFFFFu; return ACT_NOPOD; }                   // :121-135
    if (!annotations_exist(m)) { initialize_annotations(m, e); act |= ACT_INIT; }  // :141-153
  }
  const bool period = culling_check_period_has_passed(m, e);                       // :156
  if (period) act |= ACT_PERIOD;
  if (mode == 1 && !period) { *argmax = 0xFFFFFFFFu; return act; }                 // :157-159
  update_from_kernels(m, ks, e, &tr);                                              // :178
  update_from_terminals(m, ts, e, &tr);                                            // :179
  // :182 updateLastCullingCheckTimestampAnnotation writes the check timestamp; it
  // does not feed any predicate evaluated afterwards in this pass
  if (notebook_is_idle(m, e)) act |= ACT_CULL;                                     // :185
  act |= tr.src;
  *argmax = tr.argmax;
  if (tr.src) *new_la = m.get(LAST_ACTIVITY_ANNOTATION);
  return act;
}

}  // namespace

// ----------------------------------------------------------- owner lookup
// metav1.GetControllerOfNoCopy + IsControlledBy (k8s.io/apimachinery v0.33.7,
// pkg/apis/meta/v1/controller_ref.go — not under /root/reference, restated):
// the controller reference is the first owner reference whose Controller pointer is
// non-nil and true; obj is controlled by owner iff that reference's UID == owner's UID.
const GVal* controller_ref(const GVal& obj) {
  const GVal& refs = field(field(obj, "metadata"), "ownerReferences");
  if (refs.kind != GVal::List) return nullptr;
  for (const auto& r : refs.items) {
    const GVal& c = field(r, "controller");
    if (c.kind == GVal::Ptr && !c.items.empty() && c.items[0].i) return &r;
  }
  return nullptr;
}
bool is_controlled_by(const GVal& obj, const GVal& owner) {
  const GVal* ref = controller_ref(obj);
  if (!ref) return false;
  return field(*ref, "uid").s == field(field(owner, "metadata"), "uid").s;
}
// notebook_controller.go:157-170: List(InNamespace(ns)) then the first controlled item
long find_controlled(const GVal& owner, const nbh_obj* const* list, uint64_t count) {
  const std::string ns = field(field(owner, "metadata"), "namespace").s;
  for (uint64_t j = 0; j < count; ++j) {
    if (field(field(list[j]->v, "metadata"), "namespace").s != ns) continue;   // client.InNamespace
    if (is_controlled_by(list[j]->v, owner)) return (long)j;
  }
  return -1;
}

// ------------------------------------------------- ODH comparator family
// odh-notebook-controller/controllers: CompareNotebookHTTPRoutes notebook_route.go:134-138,
// CompareNotebookNetworkPolicies notebook_network.go:125-129, CompareNotebookReferenceGrants
// notebook_referencegrant.go:72-76 — all `reflect.DeepEqual(a.Labels, b.Labels) &&
// reflect.DeepEqual(a.Spec, b.Spec)`; RoleBinding subjects notebook_rbac.go:131.
const GVal& at_path(const GVal& root, const std::string& path) {
  const GVal* cur = &root;
  size_t pos = 0;
  while (!path.empty()) {
    const size_t dot = path.find('.', pos);
    const std::string key = path.substr(pos, dot == std::string::npos ? std::string::npos : dot - pos);
    while (cur->kind == GVal::Ptr && !cur->items.empty()) cur = &cur->items[0];
    cur = &field(*cur, key.c_str());
    if (dot == std::string::npos) break;
    pos = dot + 1;
  }
  return *cur;
}
bool go_deep_equal(const GVal& a, const GVal& b) {      // an absent field is its zero value; nil != empty
  if (a.kind == GVal::Null && b.kind == GVal::Null) return true;
  if (a.kind == GVal::Null) return b.is_zero();
  if (b.kind == GVal::Null) return a.is_zero();
  return deep_equal(a, b);
}
// notebook_kube_rbac_auth.go:240-266 on one map[string]string pair
bool configmap_map_differs(const GVal& found, const GVal& desired) {
  const size_t nf = found.kind == GVal::Map ? found.kv.size() : 0, nd = desired.kind == GVal::Map ? desired.kv.size() : 0;
  if (nf != nd) return true;
  if (desired.kind == GVal::Map)
    for (const auto& e : desired.kv) {
      const GVal* f = found.kind == GVal::Map ? found.find(e.first) : nullptr;
      const std::string fv = (f && f->kind == GVal::Str) ? f->s : std::string();
      if (fv != e.second.s) return true;
    }
  return false;
}

// ------------------------------------------ equality.Semantic.DeepEqual
// k8s.io/apimachinery v0.33.7 (pkg/api/equality/semantic.go + third_party/forked/golang/reflect/
// deep_equal.go; a go.mod dependency, not under /root/reference — restated from its published
// behaviour): like reflect.DeepEqual except that a nil and an empty slice/map are equal, and
// resource.Quantity values are equal when a.Cmp(b) == 0.  Pointers keep nil != &T{}.
void quantity_value(const GVal& q, std::string* digits, long long* exp) {
  // value x 10^scale (int64Amount) or unscaled x 10^-scale (inf.Dec), normalised: no trailing zeros
  std::string d; long long e = 0;
  const GVal& dec = field(field(q, "d"), "Dec");
  if (dec.kind == GVal::Ptr && !dec.items.empty()) { d = field(dec.items[0], "unscaled").s; e = -field(dec.items[0], "scale").i; }
  else { d = std::to_string((long long)field(field(q, "i"), "value").i); e = field(field(q, "i"), "scale").i; }
  bool neg = false;
  if (!d.empty() && d[0] == '-') { neg = true; d.erase(0, 1); }
  size_t nz = d.find_first_not_of('0');
  d = nz == std::string::npos ? "0" : d.substr(nz);
  while (d.size() > 1 && d.back() == '0') { d.pop_back(); ++e; }
  if (d == "0") { e = 0; neg = false; }
  *digits = (neg ? "-" : "") + d; *exp = e;
}
bool semantic_equal(const GVal& a, const GVal& b) {
  if (a.tag == nbh::kTagQuantity || b.tag == nbh::kTagQuantity) {
    if (a.tag != b.tag) return a.is_zero() && b.is_zero();     // Quantity{} against an absent one
    std::string da, db; long long ea, eb;
    quantity_value(a, &da, &ea); quantity_value(b, &db, &eb);
    return da == db && ea == eb;
  }
  auto empty_seq = [](const GVal& v) { return v.kind == GVal::Null || (v.kind == GVal::List && v.items.empty()) ||
                                              (v.kind == GVal::Map && v.kv.empty()); };
  if ((a.kind == GVal::List || a.kind == GVal::Map || b.kind == GVal::List || b.kind == GVal::Map) &&
      (empty_seq(a) || empty_seq(b)))
    return empty_seq(a) && empty_seq(b);
  if (a.kind == GVal::Struct && b.kind == GVal::Null) return semantic_equal(a, GVal::strct());
  if (a.kind == GVal::Null && b.kind == GVal::Struct) return semantic_equal(GVal::strct(), b);
  if (a.kind != b.kind) {
    if (a.kind == GVal::Null) return b.is_zero();
    if (b.kind == GVal::Null) return a.is_zero();
    return false;
  }
  switch (a.kind) {
    case GVal::List:
      if (a.items.size() != b.items.size()) return false;
      for (size_t k = 0; k < a.items.size(); ++k) if (!semantic_equal(a.items[k], b.items[k])) return false;
      return true;
    case GVal::Map:
      if (a.kv.size() != b.kv.size()) return false;
      for (const auto& e : a.kv) { const GVal* o = b.find(e.first); if (!o || !semantic_equal(e.second, *o)) return false; }
      return true;
    case GVal::Struct: {
      // union of field names; an absent field is the zero value (which may itself equal an empty slice)
      for (const auto& e : a.kv) { const GVal* o = b.find(e.first); if (!semantic_equal(e.second, o ? *o : kZero)) return false; }
      for (const auto& e : b.kv) if (!a.find(e.first) && !semantic_equal(kZero, e.second)) return false;
      return true;
    }
    case GVal::Ptr:
      if (a.items.empty() || b.items.empty()) return a.items.empty() == b.items.empty();
      return semantic_equal(a.items[0], b.items[0]);
    default: return deep_equal(a, b);
  }
}
// notebook_webhook.go:505-564; returns the decision byte of include/nbgraft.h
int restart_block(const GVal& old_nb, const GVal& updated_nb, const GVal& mutated_nb, bool create) {
  if (create) return 0;                                                                    // :515-518
  const GVal& ann = field(field(mutated_nb, "metadata"), "annotations");
  if (ann.kind == GVal::Map && ann.has("kubeflow-resource-stopped")) return 0;            // :521-524
  if (ann.kind == GVal::Map && ann.has("notebooks.opendatahub.io/notebook-restart")) return 0;   // :527-530
  auto pod = [](const GVal& nb) -> const GVal& { return field(field(field(nb, "spec"), "template"), "spec"); };
  if (!semantic_equal(pod(old_nb), pod(updated_nb))) return 2;                             // :548-551
  if (semantic_equal(pod(old_nb), pod(mutated_nb))) return 4;                              // :554-557
  return 1;                                                                                // :559-563
}

// ============================================================ exported API
// metrics.go:82-98 on objects: returns 1 when the StatefulSet counts towards its namespace
ORC_API int orc_counts_as_running(const nbh_obj* sts) {
  const GVal& labels = field(field(field(field(sts->v, "spec"), "template"), "metadata"), "labels");
  const GVal* name = labels.kind == GVal::Map ? labels.find("notebook-name") : nullptr;
  return (name && name->s == field(field(sts->v, "metadata"), "name").s) ? 1 : 0;
}
ORC_API int orc_semantic_equal(const nbh_obj* a, const nbh_obj* b) { return semantic_equal(a->v, b->v) ? 1 : 0; }
ORC_API int orc_restart_block(const nbh_obj* old_nb, const nbh_obj* updated_nb, const nbh_obj* mutated_nb, int create) {
  return restart_block(old_nb->v, updated_nb->v, mutated_nb->v, create != 0);
}
// reason bits as nbg_diff_generic: 1 labels differ, 2 spec/data differ
ORC_API int orc_compare_generic(const nbh_obj* desired, const nbh_obj* found, const char* labels_path, const char* spec_path, int mode) {
  const std::string lp(labels_path), sp(spec_path);
  const GVal& la = at_path(desired->v, lp);
  const GVal& lb = at_path(found->v, lp);
  const GVal& sa = at_path(desired->v, sp);
  const GVal& sb = at_path(found->v, sp);
  int r = 0;
  if (mode == 0) {
    if (!go_deep_equal(la, lb)) r |= 1;
    if (!go_deep_equal(sa, sb)) r |= 2;
  } else {
    if (configmap_map_differs(lb, la)) r |= 1;
    if (configmap_map_differs(sb, sa)) r |= 2;
  }
  return r;
}
ORC_API long orc_find_controlled(const nbh_obj* owner, const nbh_obj* const* list, uint64_t count) {
  return find_controlled(owner->v, list, count);
}

ORC_API int orc_deep_equal(const nbh_obj* a, const nbh_obj* b) { return deep_equal(a->v, b->v) ? 1 : 0; }
ORC_API int orc_strict_equal(const nbh_obj* a, const nbh_obj* b) { return strict_equal(a->v, b->v) ? 1 : 0; }

// These mutate `to` exactly as the Go functions do and return the reason bits
// (include/nbgraft.h encoding).  The Go bool is (reason & 0x0F) != 0 for
// StatefulSet/Service and (reason & 3) != 0 for VirtualService.
ORC_API int orc_copy_statefulset_fields(const nbh_obj* from, nbh_obj* to, int caller_prestep) {
  return copy_statefulset_fields(from->v, to->v, caller_prestep);
}
// CopyDeploymentSetFields (COMMON:136-163) on two separately decoded objects: `from.Spec.Replicas != to.Spec.Replicas`
// compares pointers (:152), which two objects never share, so it holds unless both are nil; then the POINTER is
// assigned (:153), i.e. to.Spec.Replicas takes from's value (or nil).
ORC_API int orc_copy_deployment_fields(const nbh_obj* from_o, nbh_obj* to_o) {
  const GVal& from = from_o->v;
  GVal& to = to_o->v;
  int reason = 0;
  ensure(to, "metadata", GVal::strct());
  ensure(to, "spec", GVal::strct());
  const GVal& from_meta = field(from, "metadata");
  GVal& to_meta = *to.find("metadata");
  GVal& to_spec = *to.find("spec");
  ensure(to_spec, "template", GVal::strct());
  if (live_entry_differs(field(from_meta, "labels"), field(to_meta, "labels"))) reason |= R_LABELS;        // :138-142
  assign_field(to_meta, "labels", field(from_meta, "labels"));                                             // :143
  if (live_entry_differs(field(from_meta, "annotations"), field(to_meta, "annotations"))) reason |= R_ANN; // :145-149
  assign_field(to_meta, "annotations", field(from_meta, "annotations"));                                   // :150
  const GVal& from_rep = field(field(from, "spec"), "replicas");
  const GVal to_rep = field(to_spec, "replicas");
  if (from_rep.kind == GVal::Ptr || to_rep.kind == GVal::Ptr) {                                             // :152
    assign_field(to_spec, "replicas", from_rep);                                                            // :153
    reason |= R_THIRD;
  }
  const GVal& from_pod = field(field(field(from, "spec"), "template"), "spec");
  GVal& to_tmpl = *to_spec.find("template");
  if (!deep_equal(field(to_tmpl, "spec"), from_pod)) reason |= R_FOURTH;                                    // :157
  assign_field(to_tmpl, "spec", from_pod);                                                                  // :160
  return reason;
}
ORC_API int orc_copy_service_fields(const nbh_obj* from, nbh_obj* to) { return copy_service_fields(from->v, to->v); }
ORC_API int orc_copy_virtualservice(const nbh_obj* from, nbh_obj* to) { return copy_virtualservice(from->v, to->v); }

// Batch form for timing: pairs are deep-copied first (the Go call mutates `to`),
// then the predicates run on `threads` threads; returns seconds of the timed part.
ORC_API double orc_bench_pairs(int kind /*0 sts,1 svc,2 vs*/, const nbh_obj* const* from, const nbh_obj* const* to,
                               uint64_t n, int threads, uint8_t* reason_out) {
  std::vector<GVal> live(n);
  for (uint64_t i = 0; i < n; ++i) live[i] = to[i]->v;
  if (threads < 1) threads = 1;
  std::atomic<uint64_t> next{0};
  const uint64_t grain = 256;
  auto work = [&] {
    for (uint64_t lo; (lo = next.fetch_add(grain)) < n;) {
      const uint64_t hi = lo + grain < n ? lo + grain : n;
      for (uint64_t i = lo; i < hi; ++i) {
        int r = kind == 0 ? copy_statefulset_fields(from[i]->v, live[i], 1)
              : kind == 1 ? copy_service_fields(from[i]->v, live[i]) : copy_virtualservice(from[i]->v, live[i]);
        if (reason_out) reason_out[i] = (uint8_t)r;
      }
    }
  };
  struct timespec t0, t1;
  clock_gettime(CLOCK_MONOTONIC, &t0);
  std::vector<std::thread> pool;
  for (int t = 1; t < threads; ++t) pool.emplace_back(work);
  work();
  for (auto& th : pool) th.join();
  clock_gettime(CLOCK_MONOTONIC, &t1);
  return (t1.tv_sec - t0.tv_sec) + (t1.tv_nsec - t0.tv_nsec) * 1e-9;
}

ORC_API int orc_parse_rfc3339(const char* s, int64_t* sec, int32_t* nsec, int32_t* offset) {
  GoTime t;
  if (!go_parse_rfc3339(s, &t)) return 0;
  *sec = t.sec; *nsec = t.nsec; *offset = t.offset;
  return 1;
}
ORC_API int orc_format_rfc3339(int64_t sec, int32_t offset, char* buf, size_t len) {
  GoTime t; t.sec = sec; t.offset = offset;
  const std::string s = go_format_rfc3339(t);
  if (s.size() + 1 > len) return -1;
  memcpy(buf, s.c_str(), s.size() + 1);
  return (int)s.size();
}

// Single functions, for the reference's own table tests.  `json` is the culler
// input object documented in include/nbgraft_host.h.
ORC_API int orc_all_kernels_are_idle(const char* json) {
  Meta m; std::vector<Kernel> ks, ts; bool pm;
  if (!parse_cull_input(json, &m, &ks, &ts, &pm)) return -1;
  return all_kernels_are_idle(ks) ? 1 : 0;
}
ORC_API int orc_stop_annotation_is_set(const char* json) {
  Meta m; std::vector<Kernel> ks, ts; bool pm;
  if (!parse_cull_input(json, &m, &ks, &ts, &pm)) return -1;
  return stop_annotation_is_set(m) ? 1 : 0;
}
ORC_API int orc_notebook_is_idle(const char* json, int64_t now_sec, int32_t now_nsec, int cull_idle_min) {
  Meta m; std::vector<Kernel> ks, ts; bool pm;
  if (!parse_cull_input(json, &m, &ks, &ts, &pm)) return -1;
  Env e; e.now.sec = now_sec; e.now.nsec = now_nsec; e.cull_idle_time = cull_idle_min; e.check_period = 1;
  return notebook_is_idle(m, e) ? 1 : 0;
}
ORC_API int orc_culling_check_period_has_passed(const char* json, int64_t now_sec, int32_t now_nsec, int period_min) {
  Meta m; std::vector<Kernel> ks, ts; bool pm;
  if (!parse_cull_input(json, &m, &ks, &ts, &pm)) return -1;
  Env e; e.now.sec = now_sec; e.now.nsec = now_nsec; e.cull_idle_time = 0; e.check_period = period_min;
  return culling_check_period_has_passed(m, e) ? 1 : 0;
}
// The whole decision for one notebook.  Returns the action byte, or -1.
ORC_API int orc_cull(const char* json, int64_t now_sec, int32_t now_nsec, int32_t now_offset, int cull_idle_min,
                     int period_min, int mode, uint32_t* argmax, char* new_la, size_t new_la_len) {
  Meta m; std::vector<Kernel> ks, ts; bool pm;
  if (!parse_cull_input(json, &m, &ks, &ts, &pm)) return -1;
  Env e; e.now.sec = now_sec; e.now.nsec = now_nsec; e.now.offset = now_offset;
  e.cull_idle_time = cull_idle_min; e.check_period = period_min;
  std::string la; uint32_t am = 0xFFFFFFFFu;
  const int act = cull_decide(m, ks, ts, pm, e, mode, &am, &la);
  if (argmax) *argmax = am;
  if (new_la && new_la_len) snprintf(new_la, new_la_len, "%s", la.c_str());
  return act;
}
// Many notebooks, `threads` threads; returns seconds of the timed part.  The JSON texts are decoded into
// (annotations, kernels, terminals) BEFORE the clock starts: in the reference that decoding belongs to the HTTP
// fetch (CULL:276-313 json.Unmarshal of the two Jupyter answers), outside C1-C9; what is timed is the decision
// on Go-shaped values whose timestamps are still RFC3339 strings, as C2/C3/C7/C8 receive them.  Returns -1 when
// a text does not decode.
ORC_API double orc_bench_cull(const char* const* jsons, uint64_t n, int64_t now_sec, int32_t now_nsec, int cull_idle_min,
                              int period_min, int mode, int threads, uint8_t* action_out, uint32_t* argmax_out) {
  if (threads < 1) threads = 1;
  struct In { Meta m; std::vector<Kernel> ks, ts; bool pm = false; };
  std::vector<In> ins(n);
  std::atomic<bool> bad{false};
  {
    std::atomic<uint64_t> next{0};
    auto decode = [&] {
      for (uint64_t i; (i = next.fetch_add(1)) < n;)
        if (!parse_cull_input(jsons[i], &ins[i].m, &ins[i].ks, &ins[i].ts, &ins[i].pm)) bad = true;
    };
    std::vector<std::thread> pool;
    for (int t = 1; t < threads; ++t) pool.emplace_back(decode);
    decode();
    for (auto& th : pool) th.join();
  }
  if (bad) return -1.0;
  Env e; e.now.sec = now_sec; e.now.nsec = now_nsec; e.now.offset = 0;
  e.cull_idle_time = cull_idle_min; e.check_period = period_min;
  std::atomic<uint64_t> next{0};
  const uint64_t grain = 512;
  auto work = [&] {
    for (uint64_t lo; (lo = next.fetch_add(grain)) < n;) {
      const uint64_t hi = lo + grain < n ? lo + grain : n;
      for (uint64_t i = lo; i < hi; ++i) {
        uint32_t am = 0xFFFFFFFFu;
        std::string la;
        Meta m = ins[i].m;                                   // the decision mutates the annotations (CULL:386,399,425)
        const int a = cull_decide(m, ins[i].ks, ins[i].ts, ins[i].pm, e, mode, &am, &la);
        if (action_out) action_out[i] = (uint8_t)a;
        if (argmax_out) argmax_out[i] = am;
      }
    }
  };
  struct timespec t0, t1;
  clock_gettime(CLOCK_MONOTONIC, &t0);
  std::vector<std::thread> pool;
  for (int t = 1; t < threads; ++t) pool.emplace_back(work);
  work();
  for (auto& th : pool) th.join();
  clock_gettime(CLOCK_MONOTONIC, &t1);
  return (t1.tv_sec - t0.tv_sec) + (t1.tv_nsec - t0.tv_nsec) * 1e-9;
}
