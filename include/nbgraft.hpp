// nbgraft.hpp — the reference's own function names in C++, over the two C ABIs.
//
// The reference is compiled Go; no Go toolchain exists in this image, so the host side that sits
// above libnbgraft.so is C++.  This header mirrors the interface of the path — same names, same
// argument meaning, same side effects and failure behaviour — one object at a time, each call a
// batch of one through the CUDA kernels:
//
//   nbgraft::reconcilehelper::CopyStatefulSetFields / CopyDeploymentSetFields / CopyServiceFields /
//        CopyVirtualService        components/common/reconcilehelper/util.go:107-134, 136-163, 166-195, 199-219
//   nbgraft::culler::Culler::{StopAnnotationIsSet, allKernelsAreIdle, notebookIsIdle,
//        cullingCheckPeriodHasPassed, updateTimestampFromKernelsActivity,
//        updateTimestampFromTerminalsActivity, setStopAnnotation}
//        components/notebook-controller/controllers/culling_controller.go:207-241, 315-428, 475-502
//
// For fleets use the batch API directly (nbgraft.h / nbgraft_host.h); this layer exists so that
// the reference's table tests can be restated call for call (tests/cpp/reference_tables.cpp).
// Header-only, C++17; link libnbgraft.so and libnbgraft_host.so.
#pragma once
#include <cstdint>
#include <ctime>
#include <map>
#include <stdexcept>
#include <string>
#include <utility>
#include <vector>

#include "nbgraft.h"
#include "nbgraft_host.h"

namespace nbgraft {

struct Error : std::runtime_error {
  int status;
  Error(int s, const std::string& what) : std::runtime_error(what), status(s) {}
};

// nbg_ctx.  There is no CPU path: without a CUDA device the constructor throws (status NBG_E_NODEVICE).
class Context {
 public:
  Context() {
    const int rc = nbg_init(nullptr, 0, &h_);
    if (rc != NBG_OK) throw Error(rc, std::string("nbg_init: ") + nbg_strerror(rc));
  }
  ~Context() { nbg_destroy(h_); }
  Context(const Context&) = delete;
  Context& operator=(const Context&) = delete;
  nbg_ctx* get() const { return h_; }

 private:
  nbg_ctx* h_ = nullptr;
};

// A decoded Kubernetes object (the Go value tree, held by libnbgraft_host).
class Object {
 public:
  // type: "Notebook", "StatefulSet", "Service", ... or "Unstructured"
  Object(const char* type, const std::string& json) {
    char err[256] = "";
    if (nbh_obj_from_json(type, json.c_str(), &h_, err, sizeof err) != 0) throw Error(NBG_E_INVAL, err);
  }
  explicit Object(nbh_obj* adopted) : h_(adopted) {}
  Object(const Object& o) : h_(nbh_obj_clone(o.h_)) {}
  Object(Object&& o) noexcept : h_(o.h_) { o.h_ = nullptr; }
  Object& operator=(Object o) { std::swap(h_, o.h_); return *this; }
  ~Object() { if (h_) nbh_obj_free(h_); }
  nbh_obj* get() const { return h_; }
  uint64_t digest() const { return nbh_obj_digest(h_); }          // equal <=> reflect.DeepEqual
  std::string json() const {
    char* s = nbh_obj_to_json(h_);
    std::string out(s ? s : "");
    nbh_str_free(s);
    return out;
  }
  // the sub-object at a dotted path, e.g. "metadata.labels" (digest 0-object when absent)
  bool has(const char* dotted_path) const {
    nbh_obj* sub = nullptr;
    if (nbh_obj_get(h_, dotted_path, &sub) != 0) return false;
    nbh_obj_free(sub);
    return true;
  }
  uint64_t digest_at(const char* dotted_path) const {
    nbh_obj* sub = nullptr;
    if (nbh_obj_get(h_, dotted_path, &sub) != 0) return 0;
    const uint64_t d = nbh_obj_digest(sub);
    nbh_obj_free(sub);
    return d;
  }

 private:
  nbh_obj* h_ = nullptr;
};

inline Object generateStatefulSet(const Object& notebook) {        // notebook_controller.go:433-523
  nbh_obj* out = nullptr;
  char err[256] = "";
  if (nbh_generate_statefulset(notebook.get(), nullptr, &out, err, sizeof err) != 0) throw Error(NBG_E_INVAL, err);
  return Object(out);
}
inline Object generateService(const Object& notebook) {            // notebook_controller.go:525-552
  nbh_obj* out = nullptr;
  char err[256] = "";
  if (nbh_generate_service(notebook.get(), &out, err, sizeof err) != 0) throw Error(NBG_E_INVAL, err);
  return Object(out);
}
inline Object generateVirtualService(const Object& notebook) {     // notebook_controller.go:558-658
  nbh_obj* out = nullptr;
  char err[256] = "";
  if (nbh_generate_virtualservice(notebook.get(), nullptr, &out, err, sizeof err) != 0) throw Error(NBG_E_INVAL, err);
  return Object(out);
}

namespace reconcilehelper {

// CopyStatefulSetFields(from, to *appsv1.StatefulSet) bool — util.go:107-134.  Returns whether `to`
// needed an update and overwrites to.Labels, to.Annotations, to.Spec.Template.Spec (always) and
// *to.Spec.Replicas (when different).  A nil Replicas pointer panics in Go: here it throws
// (status NBG_E_NILREPLICAS) before `to` is touched.
inline bool CopyStatefulSetFields(Context& ctx, const Object& from, Object& to) {
  nbh_sts* b = nbh_sts_new();
  nbh_sts_add(b, from.get(), to.get());
  nbg_sts_batch view;
  nbh_sts_view(b, &view);
  uint8_t reason = 0;
  uint64_t bitmap = 0;
  const int rc = nbg_diff_sts(ctx.get(), &view, 1, &reason, &bitmap);
  nbh_sts_free(b);
  if (rc == NBG_E_NILREPLICAS) throw Error(rc, "panic: runtime error: invalid memory address or nil pointer dereference (util.go:123)");
  if (rc != NBG_OK) throw Error(rc, std::string("nbg_diff_sts: ") + nbg_strerror(rc));
  nbh_apply_sts_fields(from.get(), to.get(), reason, 0);
  return (reason & NBG_STS_UPDATE_MASK) != 0;
}

// CopyDeploymentSetFields(from, to *appsv1.Deployment) bool — util.go:136-163.  As above, except that the Go
// code compares the Replicas POINTERS (:152): an update is reported, and from's replicas assigned, unless both
// are nil; nothing panics.
inline bool CopyDeploymentSetFields(Context& ctx, const Object& from, Object& to) {
  nbh_sts* b = nbh_sts_new();
  nbh_sts_add(b, from.get(), to.get());
  nbg_sts_batch view;
  nbh_sts_view(b, &view);
  uint8_t reason = 0;
  uint64_t bitmap = 0;
  const int rc = nbg_diff_deploy(ctx.get(), &view, 1, &reason, &bitmap);
  nbh_sts_free(b);
  if (rc != NBG_OK) throw Error(rc, std::string("nbg_diff_deploy: ") + nbg_strerror(rc));
  nbh_apply_sts_fields(from.get(), to.get(), reason, 0);
  return (reason & NBG_STS_UPDATE_MASK) != 0;
}

// CopyServiceFields(from, to *corev1.Service) bool — util.go:166-195
inline bool CopyServiceFields(Context& ctx, const Object& from, Object& to) {
  nbh_svc* b = nbh_svc_new();
  nbh_svc_add(b, from.get(), to.get());
  nbg_svc_batch view;
  nbh_svc_view(b, &view);
  uint8_t reason = 0;
  uint64_t bitmap = 0;
  const int rc = nbg_diff_svc(ctx.get(), &view, 1, &reason, &bitmap);
  nbh_svc_free(b);
  if (rc != NBG_OK) throw Error(rc, std::string("nbg_diff_svc: ") + nbg_strerror(rc));
  nbh_apply_svc_fields(from.get(), to.get());
  return (reason & NBG_SVC_UPDATE_MASK) != 0;
}

// CopyVirtualService(from, to *unstructured.Unstructured) bool — util.go:199-219
inline bool CopyVirtualService(Context& ctx, const Object& from, Object& to) {
  nbh_vs* b = nbh_vs_new();
  nbh_vs_add(b, from.get(), to.get());
  nbg_vs_batch view;
  nbh_vs_view(b, &view);
  uint8_t reason = 0;
  uint64_t bitmap = 0;
  const int rc = nbg_diff_vs(ctx.get(), &view, 1, &reason, &bitmap);
  nbh_vs_free(b);
  if (rc != NBG_OK) throw Error(rc, std::string("nbg_diff_vs: ") + nbg_strerror(rc));
  nbh_apply_vs_fields(from.get(), to.get(), reason);
  return (reason & NBG_VS_FROM_MISSING) ? false : (reason & (NBG_VS_SPEC_DIFFERS | NBG_VS_TO_MISSING)) != 0;
}

}  // namespace reconcilehelper

namespace culler {

constexpr const char* STOP_ANNOTATION = "kubeflow-resource-stopped";
constexpr const char* LAST_ACTIVITY_ANNOTATION = "notebooks.kubeflow.org/last-activity";
constexpr const char* LAST_ACTIVITY_CHECK_TIMESTAMP_ANNOTATION = "notebooks.kubeflow.org/last_activity_check_timestamp";
constexpr const char* KERNEL_EXECUTION_STATE_IDLE = "idle";
constexpr const char* KERNEL_EXECUTION_STATE_BUSY = "busy";
constexpr const char* KERNEL_EXECUTION_STATE_STARTING = "starting";

// metav1.ObjectMeta as far as the culler looks at it: Annotations may be a nil map.
struct ObjectMeta {
  bool annotations_nil = true;
  std::map<std::string, std::string> Annotations;
  void set(const std::string& k, const std::string& v) { annotations_nil = false; Annotations[k] = v; }
};
struct KernelStatus { std::string ExecutionState, LastActivity; };   // culling_controller.go:55-61
struct TerminalStatus { std::string LastActivity; };

struct Instant { int64_t sec = 0; int32_t nsec = 0; };

class Culler {
 public:
  int CULL_IDLE_TIME = 1440;          // minutes (culling_controller.go:37, initGlobalVars :525-558)
  int IDLENESS_CHECK_PERIOD = 1;      // minutes (:39)
  explicit Culler(Context& ctx) : ctx_(ctx) {}
  // time.Now(): the wall clock unless a test pins it
  void pin_now(int64_t sec, int32_t nsec = 0) { pinned_ = true; now_.sec = sec; now_.nsec = nsec; }
  Instant now() const {
    if (pinned_) return now_;
    timespec ts;
    clock_gettime(CLOCK_REALTIME, &ts);
    Instant t;
    t.sec = ts.tv_sec; t.nsec = (int32_t)ts.tv_nsec;
    return t;
  }
  static std::string format(int64_t sec) {                          // t.Format(time.RFC3339), UTC
    char buf[64];
    nbh_format_rfc3339(sec, 0, buf, sizeof buf);
    return buf;
  }

  bool StopAnnotationIsSet(const ObjectMeta& meta) const {          // :493-502 — key presence, no device work
    return !meta.annotations_nil && meta.Annotations.count(STOP_ANNOTATION) != 0;
  }
  bool allKernelsAreIdle(const std::vector<KernelStatus>& kernels) {                      // :315-327
    if (kernels.empty()) return true;
    ObjectMeta probe;
    probe.set(LAST_ACTIVITY_ANNOTATION, "0001-01-01T00:00:00Z");
    return (run(probe, &kernels, nullptr).action & NBG_ACT_SRC_MASK) != 1u;               // 1 = "a kernel is not idle"
  }
  bool notebookIsIdle(const ObjectMeta& meta) { return (run(meta, nullptr, nullptr).action & NBG_ACT_CULL) != 0; }            // :220-241
  bool cullingCheckPeriodHasPassed(const ObjectMeta& meta) { return (run(meta, nullptr, nullptr).action & NBG_ACT_PERIOD) != 0; }  // :207-217
  // :371-402 — may rewrite meta.Annotations[LAST_ACTIVITY_ANNOTATION]; true when it took the kernels' time
  bool updateTimestampFromKernelsActivity(ObjectMeta* meta, const std::vector<KernelStatus>& kernels) {
    if (kernels.empty()) return false;
    const Result r = run(*meta, &kernels, nullptr);
    if (r.action & NBG_ACT_SRC_MASK) meta->Annotations[LAST_ACTIVITY_ANNOTATION] = r.new_last_activity;
    return (r.action & NBG_ACT_SRC_MASK) == 2u;
  }
  bool updateTimestampFromTerminalsActivity(ObjectMeta* meta, const std::vector<TerminalStatus>& terminals) {   // :404-428
    if (terminals.empty()) return false;
    const Result r = run(*meta, nullptr, &terminals);
    if (r.action & NBG_ACT_SRC_MASK) meta->Annotations[LAST_ACTIVITY_ANNOTATION] = r.new_last_activity;
    return (r.action & NBG_ACT_SRC_MASK) == 3u;
  }
  void setStopAnnotation(ObjectMeta* meta) {                                              // :475-491 (metrics out of scope)
    if (!meta) return;
    meta->set(STOP_ANNOTATION, format(now().sec));
  }

 private:
  struct Result { unsigned action = 0; uint32_t argmax = 0; std::string new_last_activity; };

  static void quote(const std::string& s, std::string* o) {
    o->push_back('"');
    for (char c : s) {
      if (c == '"' || c == '\\') { o->push_back('\\'); o->push_back(c); }
      else if ((unsigned char)c < 0x20) { char b[8]; snprintf(b, sizeof b, "\\u%04x", c); *o += b; }
      else o->push_back(c);
    }
    o->push_back('"');
  }
  Result run(const ObjectMeta& meta, const std::vector<KernelStatus>* kernels, const std::vector<TerminalStatus>* terminals) {
    std::string j = "{\"annotations\":";
    if (meta.annotations_nil) j += "null";
    else {
      j += "{";
      bool first = true;
      for (const auto& kv : meta.Annotations) {
        if (!first) j += ",";
        first = false;
        quote(kv.first, &j); j += ":"; quote(kv.second, &j);
      }
      j += "}";
    }
    j += ",\"kernels\":[";
    if (kernels)
      for (size_t k = 0; k < kernels->size(); ++k) {
        if (k) j += ",";
        j += "{\"last_activity\":"; quote((*kernels)[k].LastActivity, &j);
        j += ",\"execution_state\":"; quote((*kernels)[k].ExecutionState, &j); j += "}";
      }
    j += "],\"terminals\":[";
    if (terminals)
      for (size_t k = 0; k < terminals->size(); ++k) {
        if (k) j += ",";
        j += "{\"last_activity\":"; quote((*terminals)[k].LastActivity, &j); j += "}";
      }
    j += "]}";
    nbh_cull* b = nbh_cull_new();
    char err[256] = "";
    if (nbh_cull_add_json(b, j.c_str(), err, sizeof err) != 0) { nbh_cull_free(b); throw Error(NBG_E_INVAL, err); }
    nbg_cull_batch view;
    nbh_cull_view(b, &view);
    const Instant t = now();
    nbg_cull_params p;
    p.now_sec = t.sec; p.now_nsec = t.nsec; p.cull_idle_min = CULL_IDLE_TIME; p.check_period_min = IDLENESS_CHECK_PERIOD;
    p.mode = NBG_CULL_PREDICATES;
    uint8_t action = 0;
    uint32_t argmax = 0;
    uint64_t cull_bits = 0, period_bits = 0;
    const int rc = nbg_cull(ctx_.get(), &view, 1, &p, &action, &argmax, &cull_bits, &period_bits);
    Result r;
    if (rc == NBG_OK) {
      r.action = action; r.argmax = argmax;
      char buf[64] = "";
      nbh_cull_new_last_activity(b, 0, action, argmax, t.sec, 0, buf, sizeof buf);
      r.new_last_activity = buf;
    }
    nbh_cull_free(b);
    if (rc != NBG_OK) throw Error(rc, std::string("nbg_cull: ") + nbg_strerror(rc));
    return r;
  }

  Context& ctx_;
  bool pinned_ = false;
  Instant now_;
};

}  // namespace culler
}  // namespace nbgraft
