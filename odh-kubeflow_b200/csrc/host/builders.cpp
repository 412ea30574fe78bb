// builders.cpp — the desired-object builders of the notebook controller,
// restated on GVal.  They define the `from` side of every comparison:
//   generateStatefulSet     NBCTL:433-523 (+ setPrefixEnvVar NBCTL:417-431)
//   generateService         NBCTL:525-552
//   generateVirtualService  NBCTL:558-658
// and, for the ODH-shaped fleet (BASELINE config #4), the fields the ODH
// mutating webhook injects into a Notebook before the controller sees it:
//   InjectReconciliationLock ODHWH:105-114, InjectKubeRbacProxy ODHWH:177-326,
//   InjectCertConfig ODHWH:666-781, MountPipelineRuntimeImages
//   notebook_runtime.go:216-285, MountElyraRuntimeConfigSecret
//   notebook_dspa_secret.go:375-450.
#include "json.hpp"
#include "k8s.hpp"
#include "quantity.hpp"

namespace nbh {
namespace {

const GVal kZero;

const GVal& field(const GVal& s, const char* name) {
  const GVal* p = s.find(name);
  return p ? *p : kZero;
}
// s.a.b.c for nested structs
const GVal& path(const GVal& s, std::initializer_list<const char*> names) {
  const GVal* cur = &s;
  for (const char* n : names) cur = &field(*cur, n);
  return *cur;
}
// mutable nested struct field, created (as an empty struct) when absent
GVal& ensure_struct(GVal& s, const char* name) {
  if (!s.find(name)) {
    // bypass set(): an empty struct is "zero" and would be dropped
    auto it = std::lower_bound(s.kv.begin(), s.kv.end(), std::string(name),
                               [](const std::pair<std::string, GVal>& e, const std::string& n) { return e.first < n; });
    s.kv.insert(it, std::make_pair(std::string(name), GVal::strct()));
  }
  return *s.find(name);
}

bool contains(const std::string& hay, const char* needle) { return hay.find(needle) != std::string::npos; }

GVal env_var(const std::string& name, const std::string& value) {
  GVal e = GVal::strct();
  e.set("name", GVal::str(name));
  e.set("value", GVal::str(value));
  return e;
}
GVal volume_mount(const std::string& name, const std::string& mount_path, bool read_only = false,
                  const std::string& sub_path = std::string()) {
  GVal m = GVal::strct();
  m.set("name", GVal::str(name));
  m.set("mountPath", GVal::str(mount_path));
  m.set("readOnly", GVal::boolean(read_only));
  m.set("subPath", GVal::str(sub_path));
  return m;
}
GVal int_or_string_int(int64_t v) {
  GVal g = GVal::strct();
  g.set("IntVal", GVal::integer(v));
  return g;
}
void append(GVal* list_or_null, GVal item) {  // Go append(): nil slice becomes non-nil
  if (list_or_null->kind != GVal::List) *list_or_null = GVal::list();
  list_or_null->items.push_back(std::move(item));
}
// the slice field `name` of struct `s`, inserted as Null when absent (caller appends)
GVal& slice_field(GVal& s, const char* name) {
  if (!s.find(name)) {
    auto it = std::lower_bound(s.kv.begin(), s.kv.end(), std::string(name),
                               [](const std::pair<std::string, GVal>& e, const std::string& n) { return e.first < n; });
    s.kv.insert(it, std::make_pair(std::string(name), GVal()));
  }
  return *s.find(name);
}

}  // namespace

// --------------------------------------------------------------------- G1
bool generate_statefulset(const GVal& nb, const BuilderEnv& env, GVal* out, std::string* err) {
  const GVal& meta = field(nb, "metadata");
  const std::string name = meta.get_str("name"), ns = meta.get_str("namespace");
  const GVal& nb_labels = field(meta, "labels");
  const GVal& nb_ann = field(meta, "annotations");
  const bool generate_name = name.size() > 52;            // NBCTL:145-149

  // NBCTL:434-437: HasAnnotation is key presence, whatever the value
  const int64_t replicas = (nb_ann.kind == GVal::Map && nb_ann.has("kubeflow-resource-stopped")) ? 0 : 1;

  GVal ss_meta = GVal::strct();
  if (!generate_name) {
    ss_meta.set("name", GVal::str(name));
    ss_meta.set("namespace", GVal::str(ns));
    GVal labels = GVal::map();                            // Labels: map[string]string{}  (NBCTL:442)
    if (nb_labels.kind == GVal::Map)
      for (const auto& e : nb_labels.kv) labels.set(e.first, e.second);   // NBCTL:474-477
    ss_meta.set("labels", std::move(labels));
  } else {
    ss_meta.set("generateName", GVal::str("nb-"));
    ss_meta.set("namespace", GVal::str(ns));
    if (nb_labels.kind == GVal::Map && !nb_labels.kv.empty()) {
      if (err) *err = "assignment to entry in nil map (NBCTL:476 with isGenerateName)";
      return false;
    }
  }

  GVal tmpl_labels = GVal::map();                         // NBCTL:462-466
  tmpl_labels.set("statefulset", GVal::str(name));
  tmpl_labels.set("notebook-name", GVal::str(name));
  tmpl_labels.set("opendatahub.io/workbenches", GVal::str("true"));
  if (nb_labels.kind == GVal::Map)
    for (const auto& e : nb_labels.kv) tmpl_labels.set(e.first, e.second);  // NBCTL:480-483
  GVal tmpl_ann = GVal::map();                            // NBCTL:467, :486-491
  if (nb_ann.kind == GVal::Map)
    for (const auto& e : nb_ann.kv)
      if (!contains(e.first, "kubectl") && !contains(e.first, "notebook")) tmpl_ann.set(e.first, e.second);

  GVal pod = path(nb, {"spec", "template", "spec"});     // DeepCopy (NBCTL:469)
  if (pod.kind != GVal::Struct) pod = GVal::strct();
  GVal* containers = pod.find("containers");
  if (!containers || containers->kind != GVal::List || containers->items.empty()) {
    if (err) *err = "index out of range [0]: Notebook has no containers (NBCTL:494)";
    return false;
  }
  GVal& c0 = containers->items[0];
  if (c0.kind != GVal::Struct) c0 = GVal::strct();
  if (c0.get_str("workingDir").empty()) c0.set("workingDir", GVal::str("/home/jovyan"));  // NBCTL:495-497
  if (!c0.find("ports")) {                                // == nil, not len 0 (NBCTL:498)
    GVal p = GVal::strct();
    p.set("containerPort", GVal::integer(8888));
    p.set("name", GVal::str("notebook-port"));
    p.set("protocol", GVal::str("TCP"));
    GVal ports = GVal::list();
    ports.items.push_back(std::move(p));
    c0.set("ports", std::move(ports));
  }
  {  // setPrefixEnvVar (NBCTL:417-431): an existing NB_PREFIX is left as it is
    bool found = false;
    const GVal* envs = c0.find("env");
    if (envs && envs->kind == GVal::List)
      for (const auto& e : envs->items)
        if (e.get_str("name") == "NB_PREFIX") { found = true; break; }
    if (!found) append(&slice_field(c0, "env"), env_var("NB_PREFIX", "/notebook/" + ns + "/" + name));
  }
  if (env.add_fsgroup && !pod.find("securityContext")) {  // NBCTL:514-521
    GVal sc = GVal::strct();
    sc.set("fsGroup", GVal::ptr(GVal::integer(100)));
    pod.set("securityContext", GVal::ptr(std::move(sc)));
  }

  GVal tmpl_meta = GVal::strct();
  tmpl_meta.set("labels", std::move(tmpl_labels));
  tmpl_meta.set("annotations", std::move(tmpl_ann));
  GVal tmpl = GVal::strct();
  tmpl.set("metadata", std::move(tmpl_meta));
  pod.canon();
  tmpl.set("spec", std::move(pod));

  GVal match = GVal::map();
  match.set("statefulset", GVal::str(name));
  GVal selector = GVal::strct();
  selector.set("matchLabels", std::move(match));

  GVal spec = GVal::strct();
  spec.set("replicas", GVal::ptr(GVal::integer(replicas)));
  spec.set("selector", GVal::ptr(std::move(selector)));
  spec.set("template", std::move(tmpl));

  GVal ss = GVal::strct();
  ss.set("metadata", std::move(ss_meta));
  ss.set("spec", std::move(spec));
  *out = std::move(ss);
  return true;
}

// --------------------------------------------------------------------- G2
bool generate_service(const GVal& nb, GVal* out, std::string* err) {
  const GVal& meta = field(nb, "metadata");
  const std::string name = meta.get_str("name"), ns = meta.get_str("namespace");
  const GVal& containers = path(nb, {"spec", "template", "spec", "containers"});
  if (containers.kind != GVal::List || containers.items.empty()) {
    if (err) *err = "index out of range [0]: Notebook has no containers (NBCTL:528)";
    return false;
  }
  int64_t port = 8888;                                    // DefaultContainerPort
  const GVal& ports = field(containers.items[0], "ports");
  if (ports.kind == GVal::List) {                         // != nil (NBCTL:529)
    if (ports.items.empty()) {
      if (err) *err = "index out of range [0]: containers[0].ports is empty but not nil (NBCTL:530)";
      return false;
    }
    port = field(ports.items[0], "containerPort").i;
  }
  GVal sp = GVal::strct();
  sp.set("name", GVal::str("http-notebook"));
  sp.set("port", GVal::integer(80));
  sp.set("targetPort", int_or_string_int(port));
  sp.set("protocol", GVal::str("TCP"));
  GVal plist = GVal::list();
  plist.items.push_back(std::move(sp));
  GVal sel = GVal::map();
  sel.set("statefulset", GVal::str(name));
  GVal spec = GVal::strct();
  spec.set("type", GVal::str("ClusterIP"));
  spec.set("selector", std::move(sel));
  spec.set("ports", std::move(plist));
  GVal m = GVal::strct();
  m.set("name", GVal::str(name));
  m.set("namespace", GVal::str(ns));
  GVal svc = GVal::strct();
  svc.set("metadata", std::move(m));
  svc.set("spec", std::move(spec));
  *out = std::move(svc);
  return true;
}

// --------------------------------------------------------------------- G3
bool generate_virtualservice(const GVal& nb, const BuilderEnv& env, GVal* out, std::string* /*err*/) {
  const GVal& meta = field(nb, "metadata");
  const std::string name = meta.get_str("name"), ns = meta.get_str("namespace");
  const GVal& ann = field(meta, "annotations");
  const std::string prefix = "/notebook/" + ns + "/" + name + "/";
  std::string rewrite = prefix;
  const std::string rw = ann.kind == GVal::Map ? ann.get_str("notebooks.kubeflow.org/http-rewrite-uri") : "";
  if (!rw.empty()) rewrite = rw;                           // NBCTL:571-574
  const std::string service = name + "." + ns + ".svc." + env.cluster_domain;

  GVal headers = GVal::map();                              // NBCTL:602-618
  const std::string hj = ann.kind == GVal::Map ? ann.get_str("notebooks.kubeflow.org/http-headers-request-set") : "";
  if (!hj.empty()) {
    GVal parsed;
    std::string perr;
    bool ok = parse_json(hj, &parsed, &perr) && (parsed.kind == GVal::Map || parsed.is_null());
    if (ok && parsed.kind == GVal::Map) {
      // json.Unmarshal into map[string]string: any non-string, non-null value is an
      // error => the reference falls back to an empty map (NBCTL:608-611)
      for (const auto& e : parsed.kv) {
        if (e.second.kind == GVal::Str) headers.set(e.first, e.second);
        else if (e.second.is_null()) headers.set(e.first, GVal::str(""));
        else { ok = false; break; }
      }
    }
    if (!ok) headers = GVal::map();
  }

  auto obj = [](std::initializer_list<std::pair<const char*, GVal>> kvs) {
    GVal m = GVal::map();
    for (const auto& e : kvs) m.set(e.first, e.second);
    return m;
  };
  auto lst = [](std::initializer_list<GVal> vs) {
    GVal l = GVal::list();
    for (const auto& v : vs) l.items.push_back(v);
    return l;
  };
  GVal http = lst({obj({
      {"headers", obj({{"request", obj({{"set", headers}})}})},
      {"match", lst({obj({{"uri", obj({{"prefix", GVal::str(prefix)}})}})})},
      {"rewrite", obj({{"uri", GVal::str(rewrite)}})},
      {"route", lst({obj({{"destination", obj({{"host", GVal::str(service)},
                                                {"port", obj({{"number", GVal::integer(80)}})}})}})})},
  })});
  GVal spec = obj({{"hosts", lst({GVal::str(env.istio_host)})},
                   {"gateways", lst({GVal::str(env.istio_gateway)})},
                   {"http", http}});
  GVal vs = obj({{"apiVersion", GVal::str("networking.istio.io/v1alpha3")},
                 {"kind", GVal::str("VirtualService")},
                 {"metadata", obj({{"name", GVal::str("notebook-" + ns + "-" + name)},
                                   {"namespace", GVal::str(ns)}})},
                 {"spec", spec}});
  *out = std::move(vs);
  return true;
}

// --------------------------------------------------------------------- G4
// parseAndValidateAuthSidecarResources (ODHWH:124-172): defaults 100m / 64Mi, overridden by the four
// auth-sidecar annotations (trimmed; empty = absent).  false where the Go function returns an error: a
// value that does not parse or is negative, or a request above its limit.
static bool auth_sidecar_resources(const GVal& nb, GVal* cpu_req, GVal* mem_req, GVal* cpu_lim, GVal* mem_lim) {
  parse_quantity("100m", cpu_req); parse_quantity("64Mi", mem_req);
  parse_quantity("100m", cpu_lim); parse_quantity("64Mi", mem_lim);
  const GVal* meta = nb.find("metadata");
  const GVal* ann = meta ? meta->find("annotations") : nullptr;
  if (ann && ann->kind == GVal::Map) {
    const std::pair<const char*, GVal*> keys[4] = {
        {"notebooks.opendatahub.io/auth-sidecar-cpu-request", cpu_req},
        {"notebooks.opendatahub.io/auth-sidecar-memory-request", mem_req},
        {"notebooks.opendatahub.io/auth-sidecar-cpu-limit", cpu_lim},
        {"notebooks.opendatahub.io/auth-sidecar-memory-limit", mem_lim}};
    for (const auto& k : keys) {
      std::string v = ann->get_str(k.first);
      if (v.empty()) continue;
      const size_t b = v.find_first_not_of(" \t\n\v\f\r"), e = v.find_last_not_of(" \t\n\v\f\r");
      v = b == std::string::npos ? std::string() : v.substr(b, e - b + 1);      // strings.TrimSpace
      GVal parsed;
      if (!parse_quantity(v, &parsed) || quantity_sign(parsed) < 0) return false;
      *k.second = std::move(parsed);
    }
  }
  return quantity_cmp(*cpu_req, *cpu_lim) <= 0 && quantity_cmp(*mem_req, *mem_lim) <= 0;
}

bool odh_webhook_mutate(GVal* nb, const OdhInject& what) {
  // create every container we are going to reference BEFORE taking references:
  // inserting a field may move its siblings
  ensure_struct(*nb, "metadata");
  ensure_struct(*nb, "spec");
  const std::string name = nb->find("metadata")->get_str("name");
  if (what.lock) {                                         // ODHWH:105-114
    GVal& meta = *nb->find("metadata");
    GVal* ann = meta.find("annotations");
    if (!ann || ann->kind != GVal::Map) { meta.set("annotations", GVal::map()); ann = meta.find("annotations"); }
    ann->set("kubeflow-resource-stopped", GVal::str("odh-notebook-controller-lock"));
  }
  GVal& pod = ensure_struct(ensure_struct(*nb->find("spec"), "template"), "spec");
  slice_field(pod, "volumes");
  slice_field(pod, "serviceAccountName");
  auto upsert_volume = [&](GVal vol, bool replace) {
    GVal& vols = slice_field(pod, "volumes");
    const std::string vname = vol.get_str("name");
    if (vols.kind == GVal::List)
      for (auto& v : vols.items)
        if (v.get_str("name") == vname) { if (replace) v = std::move(vol); return; }
    append(&vols, std::move(vol));
  };
  auto configmap_volume = [](const std::string& vname, const std::string& cm, bool optional, int64_t mode) {
    GVal src = GVal::strct();
    src.set("name", GVal::str(cm));
    if (optional) src.set("optional", GVal::ptr(GVal::boolean(true)));
    if (mode >= 0) src.set("defaultMode", GVal::ptr(GVal::integer(mode)));
    GVal v = GVal::strct();
    v.set("name", GVal::str(vname));
    v.set("configMap", GVal::ptr(std::move(src)));
    return v;
  };
  GVal& containers = slice_field(pod, "containers");

  if (what.ca_bundle) {                                    // InjectCertConfig ODHWH:666-781
    const std::string mp = "/etc/pki/tls/custom-certs/ca-bundle.crt";
    GVal vol = configmap_volume("trusted-ca", "workbench-trusted-ca-bundle", true, -1);
    GVal item = GVal::strct();
    item.set("key", GVal::str("ca-bundle.crt"));
    item.set("path", GVal::str("ca-bundle.crt"));
    GVal items = GVal::list();
    items.items.push_back(std::move(item));
    vol.find("configMap")->items[0].set("items", std::move(items));
    upsert_volume(std::move(vol), true);
    if (containers.kind == GVal::List)
      for (auto& c : containers.items) {
        if (c.get_str("name") != name) continue;
        // Go ranges over a map here (random order); this restatement uses sorted order.
        static const char* keys[] = {"GIT_SSL_CAINFO", "KF_PIPELINES_SSL_SA_CERTS", "PIPELINES_SSL_SA_CERTS",
                                     "PIP_CERT",       "REQUESTS_CA_BUNDLE",        "SSL_CERT_FILE"};
        for (const char* k : keys) {
          bool exists = false;
          const GVal* envs = c.find("env");
          if (envs && envs->kind == GVal::List)
            for (const auto& e : envs->items) if (e.get_str("name") == k) exists = true;
          if (!exists) append(&slice_field(c, "env"), env_var(k, mp));
        }
        GVal& mounts = slice_field(c, "volumeMounts");
        bool replaced = false;
        if (mounts.kind == GVal::List)
          for (auto& m : mounts.items)
            if (m.get_str("name") == "trusted-ca") { m = volume_mount("trusted-ca", mp, true, "ca-bundle.crt"); replaced = true; break; }
        if (!replaced) append(&mounts, volume_mount("trusted-ca", mp, true, "ca-bundle.crt"));
        break;
      }
  }
  if (what.runtime_images) {                               // notebook_runtime.go:216-285
    upsert_volume(configmap_volume("runtime-images", "pipeline-runtime-images", true, -1), false);
    if (containers.kind == GVal::List)
      for (auto& c : containers.items) {
        bool exists = false;
        const GVal* ms = c.find("volumeMounts");
        if (ms && ms->kind == GVal::List)
          for (const auto& m : ms->items) if (m.get_str("name") == "runtime-images") exists = true;
        if (!exists) append(&slice_field(c, "volumeMounts"), volume_mount("runtime-images", "/opt/app-root/pipeline-runtimes/"));
      }
  }
  if (what.elyra_secret) {                                 // notebook_dspa_secret.go:375-450
    GVal src = GVal::strct();
    src.set("secretName", GVal::str("ds-pipeline-config"));
    src.set("optional", GVal::ptr(GVal::boolean(true)));
    GVal v = GVal::strct();
    v.set("name", GVal::str("elyra-dsp-details"));
    v.set("secret", GVal::ptr(std::move(src)));
    upsert_volume(std::move(v), false);
    if (containers.kind == GVal::List)
      for (auto& c : containers.items) {
        bool exists = false;
        const GVal* ms = c.find("volumeMounts");
        if (ms && ms->kind == GVal::List)
          for (const auto& m : ms->items)
            if (m.get_str("name") == "elyra-dsp-details" || m.get_str("mountPath") == "/opt/app-root/runtimes") exists = true;
        if (!exists) append(&slice_field(c, "volumeMounts"), volume_mount("elyra-dsp-details", "/opt/app-root/runtimes"));
      }
  }
  if (what.feast) {                                        // Handle, ODHWH:431-444
    const GVal* labels = nb->find("metadata")->find("labels");
    const GVal* flag = (labels && labels->kind == GVal::Map) ? labels->find("opendatahub.io/feast-integration") : nullptr;
    const bool enabled = flag && flag->kind == GVal::Str && flag->s == "true";          // isFeastEnabled, feast:34-41
    if (enabled) {                                         // mountFeastConfig, feast:53-117 (all keys: no Items)
      upsert_volume(configmap_volume("odh-feast-config", name + "-feast-config", false, -1), true);
      // a missing image container is an error the webhook only logs (ODHWH:435-437): the volume stays
      if (containers.kind == GVal::List)
        for (auto& c : containers.items) {
          if (c.get_str("name") != name) continue;
          GVal& mounts = slice_field(c, "volumeMounts");
          bool replaced = false;
          if (mounts.kind == GVal::List)
            for (auto& m : mounts.items)
              if (m.get_str("name") == "odh-feast-config") { m = volume_mount("odh-feast-config", "/opt/app-root/src/feast-config", true); replaced = true; break; }
          if (!replaced) append(&mounts, volume_mount("odh-feast-config", "/opt/app-root/src/feast-config", true));
          break;
        }
    } else {                                               // isFeastMounted + unmountFeastConfig, feast:44-51,120-147
      GVal* vols = pod.find("volumes");
      bool mounted = false;
      if (vols && vols->kind == GVal::List)
        for (size_t k = 0; k < vols->items.size(); ++k)
          if (vols->items[k].get_str("name") == "odh-feast-config") { vols->items.erase(vols->items.begin() + k); mounted = true; break; }
      if (mounted && containers.kind == GVal::List)
        for (auto& c : containers.items) {
          if (c.get_str("name") != name) continue;
          GVal* mounts = c.find("volumeMounts");
          if (mounts && mounts->kind == GVal::List)
            for (size_t k = 0; k < mounts->items.size(); ++k)
              if (mounts->items[k].get_str("name") == "odh-feast-config") { mounts->items.erase(mounts->items.begin() + k); break; }
          break;
        }
      // append(s[:i], s[i+1:]...) keeps a non-nil slice: removing the last element leaves an EMPTY list, not nil
    }
  }
  if (what.rbac_proxy) {                                   // InjectKubeRbacProxy ODHWH:177-326
    GVal cpu_req, mem_req, cpu_lim, mem_lim;
    if (!auth_sidecar_resources(*nb, &cpu_req, &mem_req, &cpu_lim, &mem_lim)) { nb->canon(); return false; }   // nothing injected
    GVal req = GVal::map(), lim = GVal::map();
    req.set("cpu", std::move(cpu_req)); req.set("memory", std::move(mem_req));
    lim.set("cpu", std::move(cpu_lim)); lim.set("memory", std::move(mem_lim));
    GVal res = GVal::strct();
    res.set("requests", std::move(req));
    res.set("limits", std::move(lim));
    auto probe = [](int64_t initial_delay) {
      GVal get = GVal::strct();
      get.set("path", GVal::str("/healthz"));
      get.set("port", int_or_string_int(8444));
      get.set("scheme", GVal::str("HTTPS"));
      GVal p = GVal::strct();
      p.set("httpGet", GVal::ptr(std::move(get)));
      p.set("initialDelaySeconds", GVal::integer(initial_delay));
      p.set("timeoutSeconds", GVal::integer(1));
      p.set("periodSeconds", GVal::integer(5));
      p.set("successThreshold", GVal::integer(1));
      p.set("failureThreshold", GVal::integer(3));
      return GVal::ptr(std::move(p));
    };
    GVal args = GVal::list();
    for (const char* a : {"--secure-listen-address=0.0.0.0:8443", "--upstream=http://127.0.0.1:8888/",
                          "--logtostderr=true", "--v=10", "--proxy-endpoints-port=8444",
                          "--config-file=/etc/kube-rbac-proxy/config-file.yaml",
                          "--tls-cert-file=/etc/tls/private/tls.crt", "--tls-private-key-file=/etc/tls/private/tls.key",
                          "--auth-header-fields-enabled=true", "--auth-header-user-field-name=X-Auth-Request-User",
                          "--auth-header-groups-field-name=X-Auth-Request-Groups"})
      args.items.push_back(GVal::str(a));
    GVal port = GVal::strct();
    port.set("name", GVal::str("kube-rbac-proxy"));
    port.set("containerPort", GVal::integer(8443));
    port.set("protocol", GVal::str("TCP"));
    GVal ports = GVal::list();
    ports.items.push_back(std::move(port));
    GVal mounts = GVal::list();
    mounts.items.push_back(volume_mount("kube-rbac-proxy-config", "/etc/kube-rbac-proxy"));
    mounts.items.push_back(volume_mount("kube-rbac-proxy-tls-certificates", "/etc/tls/private"));
    GVal pc = GVal::strct();
    pc.set("name", GVal::str("kube-rbac-proxy"));
    pc.set("image", GVal::str("registry.redhat.io/openshift4/ose-kube-rbac-proxy:latest"));
    pc.set("imagePullPolicy", GVal::str("Always"));
    pc.set("args", std::move(args));
    pc.set("ports", std::move(ports));
    pc.set("livenessProbe", probe(30));
    pc.set("readinessProbe", probe(5));
    pc.set("resources", std::move(res));
    pc.set("volumeMounts", std::move(mounts));
    bool replaced = false;
    if (containers.kind == GVal::List)
      for (auto& c : containers.items)
        if (c.get_str("name") == "kube-rbac-proxy") { c = pc; replaced = true; break; }
    if (!replaced) append(&containers, std::move(pc));
    upsert_volume(configmap_volume("kube-rbac-proxy-config", name + "-kube-rbac-proxy-config", false, 420), true);
    GVal ssrc = GVal::strct();
    ssrc.set("secretName", GVal::str(name + "-kube-rbac-proxy-tls"));
    ssrc.set("defaultMode", GVal::ptr(GVal::integer(420)));
    GVal tls = GVal::strct();
    tls.set("name", GVal::str("kube-rbac-proxy-tls-certificates"));
    tls.set("secret", GVal::ptr(std::move(ssrc)));
    upsert_volume(std::move(tls), true);
    pod.set("serviceAccountName", GVal::str(name));
  }
  if (what.proxy_env && containers.kind == GVal::List) {   // InjectProxyConfigEnvVars, ODHWH:566-616
    for (auto& c : containers.items) {
      if (c.get_str("name") != name) continue;
      // Go ranges over a 3-key map (random order); this restatement appends in sorted key order.  A variable that
      // is already present keeps its value: the "update" at ODHWH:583-585 assigns to the range copy, not to the slice.
      const std::pair<const char*, const std::string*> vars[] = {
          {"HTTPS_PROXY", &what.https_proxy}, {"HTTP_PROXY", &what.http_proxy}, {"NO_PROXY", &what.no_proxy}};
      for (const auto& kv : vars) {
        bool exists = false;
        const GVal* envs = c.find("env");
        if (envs && envs->kind == GVal::List)
          for (const auto& e : envs->items) if (e.get_str("name") == kv.first) exists = true;
        if (!exists) append(&slice_field(c, "env"), env_var(kv.first, *kv.second));
      }
      break;
    }
  }
  nb->canon();
  return true;
}

// ------------------------------------------------------------- ODH desired objects
// The desired side of the ODH controller's comparators, restated from the builders it calls before
// CompareNotebookHTTPRoutes / NetworkPolicies / ReferenceGrants and the RoleBinding check.  Pointer fields that the
// Go code leaves nil stay absent (group, kind, weight ... are defaults the API server adds to the LIVE object, which
// is why a freshly created route already differs from the desired one under reflect.DeepEqual).
bool odh_build(const GVal& nb, const std::string& what, const OdhBuildEnv& env, GVal* out) {
  const GVal& meta = field(nb, "metadata");
  const std::string name = meta.get_str("name"), ns = meta.get_str("namespace");
  auto labels_of = [](std::initializer_list<std::pair<const char*, std::string>> kv) {
    GVal m = GVal::map();
    for (const auto& e : kv) m.set(e.first, GVal::str(e.second));
    return m;
  };
  auto match_labels = [&](const char* key, const std::string& val) {
    GVal sel = GVal::strct();
    sel.set("matchLabels", labels_of({{key, val}}));
    return sel;
  };
  auto str_ptr = [](const std::string& v) { return GVal::ptr(GVal::str(v)); };
  GVal o = GVal::strct(), m = GVal::strct(), spec = GVal::strct();
  if (what == "HTTPRoute") {                               // NewNotebookHTTPRoute, notebook_route.go:50-131
    const std::string route = "nb-" + ns + "-" + name;
    if (route.size() > 63)                                 // HTTPRouteSubDomainMaxLen: generateName instead (:68-79)
      m.set("generateName", GVal::str("nb-" + ns.substr(0, 10) + "-" + name.substr(0, 10) + "-"));
    else
      m.set("name", GVal::str(route));
    m.set("namespace", GVal::str(env.central_namespace));
    m.set("labels", labels_of({{"notebook-name", name}, {"notebook-namespace", ns}}));
    GVal parent = GVal::strct();
    parent.set("name", GVal::str(env.gateway_name.empty() ? "data-science-gateway" : env.gateway_name));
    parent.set("namespace", str_ptr(env.gateway_namespace.empty() ? "openshift-ingress" : env.gateway_namespace));
    GVal parents = GVal::list();
    parents.items.push_back(std::move(parent));
    GVal pathm = GVal::strct();
    pathm.set("type", str_ptr("PathPrefix"));
    pathm.set("value", str_ptr("/notebook/" + ns + "/" + name));
    GVal match = GVal::strct();
    match.set("path", GVal::ptr(std::move(pathm)));
    GVal matches = GVal::list();
    matches.items.push_back(std::move(match));
    GVal backend = GVal::strct();
    backend.set("name", GVal::str(name));
    backend.set("namespace", str_ptr(ns));
    backend.set("port", GVal::ptr(GVal::integer(8888)));
    GVal backends = GVal::list();
    backends.items.push_back(std::move(backend));
    GVal rule = GVal::strct();
    rule.set("matches", std::move(matches));
    rule.set("backendRefs", std::move(backends));
    GVal rules = GVal::list();
    rules.items.push_back(std::move(rule));
    spec.set("parentRefs", std::move(parents));
    spec.set("rules", std::move(rules));
  } else if (what == "NetworkPolicy" || what == "KubeRbacProxyNetworkPolicy") {
    // NewNotebookNetworkPolicy / NewKubeRbacProxyNetworkPolicy, notebook_network.go:132-211
    const bool proxy = what != "NetworkPolicy";
    m.set("name", GVal::str(name + (proxy ? "-kube-rbac-proxy-np" : "-ctrl-np")));
    m.set("namespace", GVal::str(ns));
    GVal port = GVal::strct();
    port.set("protocol", str_ptr("TCP"));
    {
      GVal ios = GVal::strct();                            // &intstr.IntOrString{IntVal: port}: Type stays Int (0)
      ios.set("IntVal", GVal::integer(proxy ? 8443 : 8888));
      port.set("port", GVal::ptr(std::move(ios)));
    }
    GVal ports = GVal::list();
    ports.items.push_back(std::move(port));
    GVal rule = GVal::strct();
    rule.set("ports", std::move(ports));
    if (!proxy) {                                          // only the controller's namespace may reach the notebook port
      GVal peer = GVal::strct();
      peer.set("namespaceSelector", GVal::ptr(match_labels("kubernetes.io/metadata.name", env.central_namespace)));
      GVal from = GVal::list();
      from.items.push_back(std::move(peer));
      rule.set("from", std::move(from));
    }
    GVal ingress = GVal::list();
    ingress.items.push_back(std::move(rule));
    GVal types = GVal::list();
    types.items.push_back(GVal::str("Ingress"));
    spec.set("podSelector", match_labels("notebook-name", name));
    spec.set("ingress", std::move(ingress));
    spec.set("policyTypes", std::move(types));
  } else if (what == "ReferenceGrant") {                   // NewNotebookReferenceGrant, notebook_referencegrant.go:39-69
    m.set("name", GVal::str("notebook-httproute-access"));
    m.set("namespace", GVal::str(ns));
    m.set("labels", labels_of({{"app.kubernetes.io/managed-by", "odh-notebook-controller"},
                               {"opendatahub.io/component", "notebook-controller"}}));
    GVal from = GVal::strct();
    from.set("group", GVal::str("gateway.networking.k8s.io"));
    from.set("kind", GVal::str("HTTPRoute"));
    from.set("namespace", GVal::str(env.central_namespace));
    GVal froms = GVal::list();
    froms.items.push_back(std::move(from));
    GVal to = GVal::strct();                               // Group "" (core), Name unset: any Service of the namespace
    to.set("kind", GVal::str("Service"));
    GVal tos = GVal::list();
    tos.items.push_back(std::move(to));
    spec.set("from", std::move(froms));
    spec.set("to", std::move(tos));
  } else if (what == "RoleBinding") {                      // NewRoleBinding, notebook_rbac.go:36-59
    m.set("name", GVal::str(env.rolebinding_name));
    m.set("namespace", GVal::str(ns));
    m.set("labels", labels_of({{"notebook-name", name}}));
    GVal subject = GVal::strct();
    subject.set("kind", GVal::str("ServiceAccount"));
    subject.set("name", GVal::str(name));
    subject.set("namespace", GVal::str(ns));
    GVal subjects = GVal::list();
    subjects.items.push_back(std::move(subject));
    GVal ref = GVal::strct();
    ref.set("kind", GVal::str(env.role_kind));
    ref.set("name", GVal::str(env.role_name));
    ref.set("apiGroup", GVal::str("rbac.authorization.k8s.io"));
    o.set("metadata", std::move(m));
    o.set("subjects", std::move(subjects));
    o.set("roleRef", std::move(ref));
    *out = std::move(o);
    return true;
  } else {
    return false;
  }
  o.set("metadata", std::move(m));
  o.set("spec", std::move(spec));
  *out = std::move(o);
  return true;
}

}  // namespace nbh
