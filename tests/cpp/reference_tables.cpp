// reference_tables.cpp — the reference's own table tests, restated call for call on the C++ mirror of its
// interface (include/nbgraft.hpp).  Every call below is a batch of one through the CUDA kernels.
//
//   culling_controller_test.go:14-56    TestSetStopAnnotation
//   culling_controller_test.go:58-94    TestStopAnnotationIsSet
//   culling_controller_test.go:96-141   TestAllKernelsAreIdle
//   culling_controller_test.go:143-263  TestNotebookIsIdle   (CULL_IDLE_TIME env per case, as the Go test sets it)
//   notebook_controller_bdd_test.go:49-90  StatefulSet labels == Notebook labels
// plus the quirks of util.go:107-219 that the reference has no test for (expected values derived from the cited
// lines; the same cases as tests/cases.py).
//
// Without a CUDA device the context cannot be created (no CPU path): prints "nodevice" and exits 0.
// Built and run by tests/test_cpp_mirror.py.
#include <cstdio>
#include <functional>
#include <string>
#include <vector>

#include "nbgraft.hpp"

using namespace nbgraft;
using culler::Culler;
using culler::KernelStatus;
using culler::ObjectMeta;

static int checks = 0, failures = 0;
static void expect(bool ok, const std::string& name) {
  ++checks;
  if (!ok) { ++failures; fprintf(stderr, "FAIL %s\n", name.c_str()); }
}

static ObjectMeta meta(std::initializer_list<std::pair<const char*, std::string>> ann, bool nil_map = false) {
  ObjectMeta m;
  if (!nil_map) m.annotations_nil = false;
  for (const auto& kv : ann) m.set(kv.first, kv.second);
  return m;
}

static void culler_tables(Context& ctx) {
  Culler c(ctx);
  const culler::Instant now = c.now();
  auto stamp = [&](int minutes) { return Culler::format(now.sec + 60 * minutes); };   // createTimestamp() shifted

  {  // TestSetStopAnnotation
    c.setStopAnnotation(nullptr);                                                      // "Nil Metadata": must not crash
    std::vector<std::pair<const char*, ObjectMeta>> cases = {
        {"No existing Annotations", meta({}, true)}, {"Basic case", meta({})},
        {"Annotation is already set", meta({{culler::STOP_ANNOTATION, stamp(0)}})}};
    for (auto& tc : cases) {
      c.setStopAnnotation(&tc.second);
      expect(tc.second.Annotations.count(culler::STOP_ANNOTATION) == 1, std::string("TestSetStopAnnotation/") + tc.first);
    }
  }
  {  // TestStopAnnotationIsSet
    struct Case { const char* name; ObjectMeta meta; bool result; };
    const std::vector<Case> cases = {{"No existing Annotations", meta({}, true), false},
                                     {"Stop Annotation not Set", meta({}), false},
                                     {"Annotation is already set", meta({{culler::STOP_ANNOTATION, stamp(0)}}), true}};
    for (const auto& tc : cases) expect(c.StopAnnotationIsSet(tc.meta) == tc.result, std::string("TestStopAnnotationIsSet/") + tc.name);
  }
  {  // TestAllKernelsAreIdle
    struct Case { const char* name; std::vector<KernelStatus> kernels; bool result; };
    const std::vector<Case> cases = {
        {"/api/kernels returns an empty list of kernels.", {}, true},
        {"/api/kernels returns an list of idle kernels.",
         {{culler::KERNEL_EXECUTION_STATE_IDLE, ""}, {culler::KERNEL_EXECUTION_STATE_IDLE, ""}}, true},
        {"/api/kernels returns an list of kernels, with one kernel in busy state.",
         {{culler::KERNEL_EXECUTION_STATE_IDLE, ""}, {culler::KERNEL_EXECUTION_STATE_BUSY, ""}}, false}};
    for (const auto& tc : cases) expect(c.allKernelsAreIdle(tc.kernels) == tc.result, std::string("TestAllKernelsAreIdle/") + tc.name);
  }
  {  // TestNotebookIsIdle: env {} keeps whatever CULL_IDLE_TIME an earlier case left (the Go cases run in order)
    struct Case { const char* name; ObjectMeta meta; int cull_idle_time; bool result; };
    const std::vector<Case> cases = {
        {"No existing Annotations", meta({}, true), 0, false},
        {"Basic case", meta({}), 0, false},
        {"Stop Annotation already set", meta({{culler::STOP_ANNOTATION, stamp(0)}}), 0, false},
        {"LAST_ACTIVITY_ANNOTATION is not set", meta({}), 0, false},
        {"LAST_ACTIVITY_ANNOTATION is not RF3339", meta({{culler::LAST_ACTIVITY_ANNOTATION, "should-fail"}}), 0, false},
        {"LAST_ACTIVITY_ANNOTATION is old", meta({{culler::LAST_ACTIVITY_ANNOTATION, "2021-08-30T15:37:36.990063Z"}}), 0, true},
        {"LAST_ACTIVITY_ANNOTATION is too old", meta({{culler::LAST_ACTIVITY_ANNOTATION, "1900-08-30T15:37:36.990063Z"}}), 0, true},
        {"LAST_ACTIVITY_ANNOTATION is the current time", meta({{culler::LAST_ACTIVITY_ANNOTATION, stamp(0)}}), 5, false},
        {"LAST_ACTIVITY_ANNOTATION is 1 minute MORE than the deadline.", meta({{culler::LAST_ACTIVITY_ANNOTATION, stamp(-6)}}), 5, true},
        {"LAST_ACTIVITY_ANNOTATION is 1 minute LESS than the deadline.", meta({{culler::LAST_ACTIVITY_ANNOTATION, stamp(-3)}}), 5, false}};
    c.CULL_IDLE_TIME = 1440;                                                           // initGlobalVars default
    for (const auto& tc : cases) {
      if (tc.cull_idle_time) c.CULL_IDLE_TIME = tc.cull_idle_time;                     // t.Setenv + initGlobalVars
      expect(c.notebookIsIdle(tc.meta) == tc.result, std::string("TestNotebookIsIdle/") + tc.name);
    }
  }
  {  // derived from culling_controller.go:329-402: the kernels' newest activity replaces an older annotation, fraction dropped
    Culler d(ctx);
    d.pin_now(1767225600, 123456789);
    ObjectMeta m = meta({{culler::LAST_ACTIVITY_ANNOTATION, "2025-12-31T10:00:00Z"}});
    const std::vector<KernelStatus> ks = {{"idle", "2025-12-31T11:00:00.5Z"}, {"idle", "2025-12-31T12:30:00.990063Z"}};
    expect(d.updateTimestampFromKernelsActivity(&m, ks), "kernels' activity is newer");
    expect(m.Annotations[culler::LAST_ACTIVITY_ANNOTATION] == "2025-12-31T12:30:00Z", "Format(RFC3339) drops the fraction");
    const std::vector<KernelStatus> busy = {{"idle", "2025-12-31T11:00:00Z"}, {"busy", "2025-12-31T09:00:00Z"}};
    expect(!d.updateTimestampFromKernelsActivity(&m, busy), "a busy kernel: annotation := now, returns false");
    expect(m.Annotations[culler::LAST_ACTIVITY_ANNOTATION] == "2026-01-01T00:00:00Z", "now at second precision");
  }
}

static const char* kNotebook =                                                        // notebook_controller_bdd_test.go:49-63
    R"({"apiVersion":"kubeflow.org/v1","kind":"Notebook","metadata":{"name":"test-notebook","namespace":"default",
        "labels":{"testLabel":"testLabelValue"}},
        "spec":{"template":{"spec":{"containers":[{"name":"busybox","image":"busybox"}]}}}})";

static void reconcilehelper_tables(Context& ctx) {
  const Object nb("Notebook", kNotebook);
  {  // BDD: the StatefulSet built for the Notebook carries exactly its labels (:90), so nothing is to update
    const Object desired = generateStatefulSet(nb);
    Object live = desired;
    expect(desired.digest_at("metadata.labels") == nb.digest_at("metadata.labels"), "BDD: StatefulSet labels == Notebook labels");
    expect(!reconcilehelper::CopyStatefulSetFields(ctx, desired, live), "BDD: desired vs itself needs no update");
    expect(live.digest() == desired.digest(), "and the copies leave it as it was");
  }
  struct Edit { const char* name; std::function<void(Object&)> on_live; bool want; };
  auto set = [](const char* path, const char* v) { return [=](Object& o) { nbh_obj_set_str(o.get(), path, v); }; };
  auto erase = [](const char* path) { return [=](Object& o) { nbh_obj_erase(o.get(), path); }; };
  {  // CopyStatefulSetFields, util.go:107-134
    const std::vector<Edit> cases = {
        {"live label changed", set("metadata.labels.testLabel", "someone-else"), true},
        {"live label whose value is empty and that desired lacks: reads as \"\" == \"\"", set("metadata.labels.extra", ""), false},
        {"live-only label with a value", set("metadata.labels.extra", "x"), true},
        {"desired-only label: the loop ranges over live keys only", erase("metadata.labels.testLabel"), false},
        {"image changed", set("spec.template.spec.containers.0.image", "busybox:2"), true},
        {"a defaulted field on the live side (reflect.DeepEqual is exact)", set("spec.template.spec.dnsPolicy", "ClusterFirst"), true}};
    for (const auto& tc : cases) {
      const Object desired = generateStatefulSet(nb);
      Object live = desired;
      tc.on_live(live);
      expect(reconcilehelper::CopyStatefulSetFields(ctx, desired, live) == tc.want, std::string("CopyStatefulSetFields/") + tc.name);
      expect(live.digest_at("spec.template.spec") == desired.digest_at("spec.template.spec") &&
                 live.digest_at("metadata.labels") == desired.digest_at("metadata.labels"),
             std::string("CopyStatefulSetFields copies unconditionally/") + tc.name);
    }
  }
  {  // CopyDeploymentSetFields, util.go:136-163: Replicas are compared as pointers (:152)
    const char* dep = R"({"apiVersion":"apps/v1","kind":"Deployment","metadata":{"name":"d","labels":{"app":"d"}},
                          "spec":{"replicas":2,"template":{"spec":{"containers":[{"name":"c","image":"nginx"}]}}}})";
    const Object desired("Deployment", dep);
    Object live = desired;
    expect(reconcilehelper::CopyDeploymentSetFields(ctx, desired, live), "CopyDeploymentSetFields/equal numbers, different pointers => update");
    Object d_nil = desired, l_nil = desired;
    nbh_obj_erase(d_nil.get(), "spec.replicas");
    nbh_obj_erase(l_nil.get(), "spec.replicas");
    expect(!reconcilehelper::CopyDeploymentSetFields(ctx, d_nil, l_nil), "CopyDeploymentSetFields/both nil => nothing to do");
    Object l3 = desired;
    expect(reconcilehelper::CopyDeploymentSetFields(ctx, d_nil, l3) && !l3.has("spec.replicas"),
           "CopyDeploymentSetFields/desired nil: no panic, the nil pointer is assigned");
  }
  {  // CopyServiceFields, util.go:166-195
    const std::vector<Edit> cases = {{"identical", [](Object&) {}, false},
                                     {"selector changed", set("spec.selector.statefulset", "other"), true},
                                     {"clusterIP is deliberately not compared (:182)", set("spec.clusterIP", "10.0.0.7"), false},
                                     {"port name changed", set("spec.ports.0.name", "http"), true}};
    for (const auto& tc : cases) {
      const Object desired = generateService(nb);
      Object live = desired;
      tc.on_live(live);
      expect(reconcilehelper::CopyServiceFields(ctx, desired, live) == tc.want, std::string("CopyServiceFields/") + tc.name);
      expect(live.digest_at("spec.ports") == desired.digest_at("spec.ports") &&
                 live.digest_at("spec.selector") == desired.digest_at("spec.selector"),
             std::string("CopyServiceFields copies/") + tc.name);
    }
    const Object desired = generateService(nb);
    Object live = desired;
    nbh_obj_set_str(live.get(), "spec.clusterIP", "10.0.0.7");
    reconcilehelper::CopyServiceFields(ctx, desired, live);
    expect(live.has("spec.clusterIP"), "ClusterIP survives the copy");
  }
  {  // CopyVirtualService, util.go:199-219
    const Object desired = generateVirtualService(nb);
    Object same = desired;
    expect(!reconcilehelper::CopyVirtualService(ctx, desired, same), "CopyVirtualService/identical");
    Object changed = desired;
    nbh_obj_set_str(changed.get(), "spec.http.0.rewrite.uri", "/elsewhere/");
    expect(reconcilehelper::CopyVirtualService(ctx, desired, changed), "CopyVirtualService/rewrite changed");
    expect(changed.digest_at("spec") == desired.digest_at("spec"), "CopyVirtualService/spec copied");
    Object no_spec = desired;
    nbh_obj_erase(no_spec.get(), "spec");
    expect(reconcilehelper::CopyVirtualService(ctx, desired, no_spec), "CopyVirtualService/to.spec missing => copied, true (:208-212)");
    expect(no_spec.digest_at("spec") == desired.digest_at("spec"), "CopyVirtualService/missing spec filled in");
    Object from_no_spec = desired;
    nbh_obj_erase(from_no_spec.get(), "spec");
    Object live = desired;
    expect(!reconcilehelper::CopyVirtualService(ctx, from_no_spec, live), "CopyVirtualService/from.spec missing => false (:200-206)");
    expect(live.digest() == desired.digest(), "CopyVirtualService/and nothing is written");
  }
}

int main() {
  try {
    Context ctx;
    culler_tables(ctx);
    reconcilehelper_tables(ctx);
    if (failures) { fprintf(stderr, "%d of %d checks failed\n", failures, checks); return 1; }
    printf("ok %d checks, %llu kernel launches\n", checks, (unsigned long long)nbg_launch_count(ctx.get()));
    return 0;
  } catch (const Error& e) {
    if (e.status == NBG_E_NODEVICE) { printf("nodevice\n"); return 0; }
    fprintf(stderr, "error: %s\n", e.what());
    return 1;
  }
}
