// k8s.cpp — typed decoding of Kubernetes JSON into GVal.
//
// The field tables restate the shapes of k8s.io/api v0.33.7 (core/v1, apps/v1)
// and the Notebook CRD (reference: components/notebook-controller/api/v1beta1/
// notebook_types.go:26-33, NotebookSpec.Template.Spec = corev1.PodSpec).  The
// k8s.io/api module is a go.mod dependency (NBC/go.mod:13) and is not under the
// reference tree; names and kinds are the published API.  HTTPRoute / ReferenceGrant
// are sigs.k8s.io/gateway-api v1.2.1 (ODH/go.mod:24; inlined CommonRouteSpec, BackendRef
// and BackendObjectReference are flattened), NetworkPolicy is networking/v1, RoleBinding
// rbac/v1 — the objects the ODH comparators look at.
//
// Type syntax of the table: s string, i integer, b bool, f float, q Quantity,
// ios IntOrString, any untyped JSON, [T slice, *T pointer, {s map[string]string,
// {q map[string]Quantity (ResourceList), Name another struct.
#include "k8s.hpp"

#include <map>
#include <memory>
#include <mutex>
#include <sstream>

#include "json.hpp"
#include "quantity.hpp"

namespace nbh {
namespace {

const char* kSchema = R"(
ObjectMeta: name:s generateName:s namespace:s uid:s resourceVersion:s generation:i creationTimestamp:any deletionTimestamp:any deletionGracePeriodSeconds:*i labels:{s annotations:{s ownerReferences:[OwnerReference finalizers:[s managedFields:any selfLink:s
OwnerReference: apiVersion:s kind:s name:s uid:s controller:*b blockOwnerDeletion:*b
LocalObjectReference: name:s
KeyToPath: key:s path:s mode:*i
ObjectFieldSelector: apiVersion:s fieldPath:s
ResourceFieldSelector: containerName:s resource:s divisor:q
ConfigMapKeySelector: name:s key:s optional:*b
SecretKeySelector: name:s key:s optional:*b
EnvVarSource: fieldRef:*ObjectFieldSelector resourceFieldRef:*ResourceFieldSelector configMapKeyRef:*ConfigMapKeySelector secretKeyRef:*SecretKeySelector
EnvVar: name:s value:s valueFrom:*EnvVarSource
ConfigMapEnvSource: name:s optional:*b
SecretEnvSource: name:s optional:*b
EnvFromSource: prefix:s configMapRef:*ConfigMapEnvSource secretRef:*SecretEnvSource
ContainerPort: name:s hostPort:i containerPort:i protocol:s hostIP:s
ResourceClaim: name:s request:s
ResourceRequirements: limits:{q requests:{q claims:[ResourceClaim
VolumeMount: name:s readOnly:b recursiveReadOnly:*s mountPath:s subPath:s mountPropagation:*s subPathExpr:s
VolumeDevice: name:s devicePath:s
ExecAction: command:[s
HTTPHeader: name:s value:s
HTTPGetAction: path:s port:ios host:s scheme:s httpHeaders:[HTTPHeader
TCPSocketAction: port:ios host:s
GRPCAction: port:i service:*s
SleepAction: seconds:i
Probe: exec:*ExecAction httpGet:*HTTPGetAction tcpSocket:*TCPSocketAction grpc:*GRPCAction initialDelaySeconds:i timeoutSeconds:i periodSeconds:i successThreshold:i failureThreshold:i terminationGracePeriodSeconds:*i
LifecycleHandler: exec:*ExecAction httpGet:*HTTPGetAction tcpSocket:*TCPSocketAction sleep:*SleepAction
Lifecycle: postStart:*LifecycleHandler preStop:*LifecycleHandler stopSignal:*s
Capabilities: add:[s drop:[s
SELinuxOptions: user:s role:s type:s level:s
SeccompProfile: type:s localhostProfile:*s
AppArmorProfile: type:s localhostProfile:*s
WindowsSecurityContextOptions: gmsaCredentialSpecName:*s gmsaCredentialSpec:*s runAsUserName:*s hostProcess:*b
SecurityContext: capabilities:*Capabilities privileged:*b seLinuxOptions:*SELinuxOptions windowsOptions:*WindowsSecurityContextOptions runAsUser:*i runAsGroup:*i runAsNonRoot:*b readOnlyRootFilesystem:*b allowPrivilegeEscalation:*b procMount:*s seccompProfile:*SeccompProfile appArmorProfile:*AppArmorProfile
Sysctl: name:s value:s
PodSecurityContext: seLinuxOptions:*SELinuxOptions windowsOptions:*WindowsSecurityContextOptions runAsUser:*i runAsGroup:*i runAsNonRoot:*b supplementalGroups:[i supplementalGroupsPolicy:*s fsGroup:*i sysctls:[Sysctl fsGroupChangePolicy:*s seccompProfile:*SeccompProfile appArmorProfile:*AppArmorProfile seLinuxChangePolicy:*s
ContainerResizePolicy: resourceName:s restartPolicy:s
Container: name:s image:s command:[s args:[s workingDir:s ports:[ContainerPort envFrom:[EnvFromSource env:[EnvVar resources:ResourceRequirements resizePolicy:[ContainerResizePolicy restartPolicy:*s volumeMounts:[VolumeMount volumeDevices:[VolumeDevice livenessProbe:*Probe readinessProbe:*Probe startupProbe:*Probe lifecycle:*Lifecycle terminationMessagePath:s terminationMessagePolicy:s imagePullPolicy:s securityContext:*SecurityContext stdin:b stdinOnce:b tty:b
HostPathVolumeSource: path:s type:*s
EmptyDirVolumeSource: medium:s sizeLimit:*q
SecretVolumeSource: secretName:s items:[KeyToPath defaultMode:*i optional:*b
ConfigMapVolumeSource: name:s items:[KeyToPath defaultMode:*i optional:*b
PersistentVolumeClaimVolumeSource: claimName:s readOnly:b
NFSVolumeSource: server:s path:s readOnly:b
DownwardAPIVolumeFile: path:s fieldRef:*ObjectFieldSelector resourceFieldRef:*ResourceFieldSelector mode:*i
DownwardAPIVolumeSource: items:[DownwardAPIVolumeFile defaultMode:*i
SecretProjection: name:s items:[KeyToPath optional:*b
ConfigMapProjection: name:s items:[KeyToPath optional:*b
DownwardAPIProjection: items:[DownwardAPIVolumeFile
ServiceAccountTokenProjection: audience:s expirationSeconds:*i path:s
VolumeProjection: secret:*SecretProjection downwardAPI:*DownwardAPIProjection configMap:*ConfigMapProjection serviceAccountToken:*ServiceAccountTokenProjection clusterTrustBundle:any
ProjectedVolumeSource: sources:[VolumeProjection defaultMode:*i
Volume: name:s hostPath:*HostPathVolumeSource emptyDir:*EmptyDirVolumeSource secret:*SecretVolumeSource configMap:*ConfigMapVolumeSource persistentVolumeClaim:*PersistentVolumeClaimVolumeSource nfs:*NFSVolumeSource downwardAPI:*DownwardAPIVolumeSource projected:*ProjectedVolumeSource csi:any ephemeral:any
Toleration: key:s operator:s value:s effect:s tolerationSeconds:*i
HostAlias: ip:s hostnames:[s
PodDNSConfigOption: name:s value:*s
PodDNSConfig: nameservers:[s searches:[s options:[PodDNSConfigOption
PodReadinessGate: conditionType:s
PodOS: name:s
PodSchedulingGate: name:s
PodSpec: volumes:[Volume initContainers:[Container containers:[Container ephemeralContainers:any restartPolicy:s terminationGracePeriodSeconds:*i activeDeadlineSeconds:*i dnsPolicy:s nodeSelector:{s serviceAccountName:s serviceAccount:s automountServiceAccountToken:*b nodeName:s hostNetwork:b hostPID:b hostIPC:b shareProcessNamespace:*b securityContext:*PodSecurityContext imagePullSecrets:[LocalObjectReference hostname:s subdomain:s affinity:any schedulerName:s tolerations:[Toleration hostAliases:[HostAlias priorityClassName:s priority:*i dnsConfig:*PodDNSConfig readinessGates:[PodReadinessGate runtimeClassName:*s enableServiceLinks:*b preemptionPolicy:*s overhead:{q topologySpreadConstraints:any setHostnameAsFQDN:*b os:*PodOS hostUsers:*b schedulingGates:[PodSchedulingGate resourceClaims:any resources:*ResourceRequirements
PodTemplateSpec: metadata:ObjectMeta spec:PodSpec
LabelSelector: matchLabels:{s matchExpressions:any
StatefulSetSpec: replicas:*i selector:*LabelSelector template:PodTemplateSpec volumeClaimTemplates:any serviceName:s podManagementPolicy:s updateStrategy:any revisionHistoryLimit:*i minReadySeconds:i persistentVolumeClaimRetentionPolicy:any ordinals:any
StatefulSet: apiVersion:s kind:s metadata:ObjectMeta spec:StatefulSetSpec status:any
DeploymentSpec: replicas:*i selector:*LabelSelector template:PodTemplateSpec strategy:any minReadySeconds:i revisionHistoryLimit:*i paused:b progressDeadlineSeconds:*i
Deployment: apiVersion:s kind:s metadata:ObjectMeta spec:DeploymentSpec status:any
ServicePort: name:s protocol:s appProtocol:*s port:i targetPort:ios nodePort:i
ServiceSpec: ports:[ServicePort selector:{s clusterIP:s clusterIPs:[s type:s externalIPs:[s sessionAffinity:s loadBalancerIP:s loadBalancerSourceRanges:[s externalName:s externalTrafficPolicy:s healthCheckNodePort:i publishNotReadyAddresses:b sessionAffinityConfig:any ipFamilies:[s ipFamilyPolicy:*s allocateLoadBalancerNodePorts:*b loadBalancerClass:*s internalTrafficPolicy:*s trafficDistribution:*s
Service: apiVersion:s kind:s metadata:ObjectMeta spec:ServiceSpec status:any
ParentReference: group:*s kind:*s namespace:*s name:s sectionName:*s port:*i
HTTPPathMatch: type:*s value:*s
HTTPHeaderMatch: type:*s name:s value:s
HTTPQueryParamMatch: type:*s name:s value:s
HTTPRouteMatch: path:*HTTPPathMatch headers:[HTTPHeaderMatch queryParams:[HTTPQueryParamMatch method:*s
HTTPBackendRef: group:*s kind:*s name:s namespace:*s port:*i weight:*i filters:any
HTTPRouteRule: name:*s matches:[HTTPRouteMatch filters:any backendRefs:[HTTPBackendRef timeouts:any retry:any sessionPersistence:any
HTTPRouteSpec: parentRefs:[ParentReference hostnames:[s rules:[HTTPRouteRule
HTTPRoute: apiVersion:s kind:s metadata:ObjectMeta spec:HTTPRouteSpec status:any
NetworkPolicyPort: protocol:*s port:*ios endPort:*i
IPBlock: cidr:s except:[s
NetworkPolicyPeer: podSelector:*LabelSelector namespaceSelector:*LabelSelector ipBlock:*IPBlock
NetworkPolicyIngressRule: ports:[NetworkPolicyPort from:[NetworkPolicyPeer
NetworkPolicyEgressRule: ports:[NetworkPolicyPort to:[NetworkPolicyPeer
NetworkPolicySpec: podSelector:LabelSelector ingress:[NetworkPolicyIngressRule egress:[NetworkPolicyEgressRule policyTypes:[s
NetworkPolicy: apiVersion:s kind:s metadata:ObjectMeta spec:NetworkPolicySpec
ReferenceGrantFrom: group:s kind:s namespace:s
ReferenceGrantTo: group:s kind:s name:*s
ReferenceGrantSpec: from:[ReferenceGrantFrom to:[ReferenceGrantTo
ReferenceGrant: apiVersion:s kind:s metadata:ObjectMeta spec:ReferenceGrantSpec
ConfigMap: apiVersion:s kind:s metadata:ObjectMeta immutable:*b data:{s binaryData:any
Subject: kind:s apiGroup:s name:s namespace:s
RoleRef: apiGroup:s kind:s name:s
RoleBinding: apiVersion:s kind:s metadata:ObjectMeta subjects:[Subject roleRef:RoleRef
NotebookTemplateSpec: spec:PodSpec
NotebookSpec: template:NotebookTemplateSpec
Notebook: apiVersion:s kind:s metadata:ObjectMeta spec:NotebookSpec status:any
)";

struct Type {
  enum K { STR, INT, BOOL, FLOAT, QUANTITY, INTORSTR, ANY, SLICE, PTR, MAPSTR, MAPQ, STRUCT } k = ANY;
  const Type* elem = nullptr;
  std::string name;
  std::vector<std::pair<std::string, const Type*>> fields;
};

struct Registry {
  std::map<std::string, std::unique_ptr<Type>> named;
  std::vector<std::unique_ptr<Type>> anon;
  Type prim[7];

  const Type* parse_type(const std::string& t) {
    if (t == "s") return &prim[0];
    if (t == "i") return &prim[1];
    if (t == "b") return &prim[2];
    if (t == "f") return &prim[3];
    if (t == "q") return &prim[4];
    if (t == "ios") return &prim[5];
    if (t == "any") return &prim[6];
    if (t[0] == '[' || t[0] == '*') {
      auto ty = std::make_unique<Type>();
      ty->k = t[0] == '[' ? Type::SLICE : Type::PTR;
      ty->elem = parse_type(t.substr(1));
      anon.push_back(std::move(ty));
      return anon.back().get();
    }
    if (t == "{s" || t == "{q") {
      auto ty = std::make_unique<Type>();
      ty->k = t == "{s" ? Type::MAPSTR : Type::MAPQ;
      anon.push_back(std::move(ty));
      return anon.back().get();
    }
    auto& slot = named[t];
    if (!slot) { slot = std::make_unique<Type>(); slot->k = Type::STRUCT; slot->name = t; }
    return slot.get();
  }

  Registry() {
    prim[0].k = Type::STR; prim[1].k = Type::INT; prim[2].k = Type::BOOL; prim[3].k = Type::FLOAT;
    prim[4].k = Type::QUANTITY; prim[5].k = Type::INTORSTR; prim[6].k = Type::ANY;
    std::istringstream in(kSchema);
    std::string line;
    while (std::getline(in, line)) {
      if (line.empty()) continue;
      std::istringstream ls(line);
      std::string head;
      ls >> head;
      head.pop_back();  // ':'
      Type* ty = const_cast<Type*>(parse_type(head));
      std::string f;
      while (ls >> f) {
        const size_t c = f.find(':');
        ty->fields.emplace_back(f.substr(0, c), parse_type(f.substr(c + 1)));
      }
    }
  }
  const Type* get(const std::string& n) const {
    auto it = named.find(n);
    return (it == named.end() || it->second->fields.empty()) ? nullptr : it->second.get();
  }
};

const Registry& registry() {
  static const Registry r;
  return r;
}

// untyped JSON kept for fields outside the table: null members vanish (a JSON
// null decodes to the zero value of whatever Go type the field has)
GVal strip_nulls(const GVal& j) {
  GVal o = j;
  if (o.kind == GVal::Map) {
    o.kv.clear();
    for (const auto& e : j.kv)
      if (!e.second.is_null()) o.kv.emplace_back(e.first, strip_nulls(e.second));
  } else if (o.kind == GVal::List) {
    for (auto& c : o.items) c = strip_nulls(c);
  }
  return o;
}

bool convert(const GVal& j, const Type* t, GVal* out, std::string* err, const std::string& path) {
  auto bad = [&](const char* want) {
    if (err) *err = path + ": expected " + want;
    return false;
  };
  if (j.is_null()) { *out = GVal(); return true; }  // JSON null => Go zero value
  switch (t->k) {
    case Type::STR:
      if (j.kind != GVal::Str) return bad("string");
      *out = j; return true;
    case Type::INT:
      if (j.kind != GVal::Int) return bad("integer");
      *out = j; return true;
    case Type::BOOL:
      if (j.kind != GVal::Bool) return bad("bool");
      *out = j; return true;
    case Type::FLOAT:
      if (j.kind == GVal::Int) { *out = GVal::real((double)j.i); return true; }
      if (j.kind != GVal::Float) return bad("number");
      *out = j; return true;
    case Type::QUANTITY: {
      std::string text;
      if (j.kind == GVal::Str) {
        text = j.s;
        const size_t b = text.find_first_not_of(" \t\r\n");
        const size_t e = text.find_last_not_of(" \t\r\n");
        text = b == std::string::npos ? std::string() : text.substr(b, e - b + 1);
      } else if (j.kind == GVal::Int) {
        text = std::to_string(j.i);
      } else if (j.kind == GVal::Float) {
        std::ostringstream os; os << j.f; text = os.str();
      } else {
        return bad("quantity");
      }
      if (!parse_quantity(text, out)) return bad("a parseable quantity");
      return true;
    }
    case Type::INTORSTR: {  // intstr.IntOrString{Type, IntVal, StrVal}
      GVal v = GVal::strct();
      if (j.kind == GVal::Int) { v.set("IntVal", GVal::integer(j.i)); }
      else if (j.kind == GVal::Str) { v.set("Type", GVal::integer(1)); v.set("StrVal", GVal::str(j.s)); }
      else return bad("int or string");
      *out = std::move(v); return true;
    }
    case Type::ANY:
      *out = strip_nulls(j); return true;
    case Type::SLICE: {
      if (j.kind != GVal::List) return bad("array");
      GVal l = GVal::list();
      l.items.reserve(j.items.size());
      for (size_t k = 0; k < j.items.size(); ++k) {
        GVal e;
        if (!convert(j.items[k], t->elem, &e, err, path + "[" + std::to_string(k) + "]")) return false;
        // a null element of a struct slice is a zero struct, keep its position
        if (e.is_null() && t->elem->k == Type::STRUCT) e = GVal::strct();
        if (e.is_null() && t->elem->k == Type::STR) e = GVal::str("");
        if (e.is_null() && t->elem->k == Type::INT) e = GVal::integer(0);
        l.items.push_back(std::move(e));
      }
      *out = std::move(l); return true;
    }
    case Type::PTR: {
      GVal e;
      if (!convert(j, t->elem, &e, err, path)) return false;
      if (e.is_null()) {  // pointee decoded to a zero leaf/struct: still a non-nil pointer
        switch (t->elem->k) {
          case Type::STR: e = GVal::str(""); break;
          case Type::INT: e = GVal::integer(0); break;
          case Type::BOOL: e = GVal::boolean(false); break;
          default: break;
        }
      }
      *out = GVal::ptr(std::move(e)); return true;
    }
    case Type::MAPSTR: {
      if (j.kind != GVal::Map) return bad("object");
      GVal m = GVal::map();
      for (const auto& e : j.kv) {
        if (e.second.is_null()) { m.set(e.first, GVal::str("")); continue; }
        if (e.second.kind != GVal::Str) return bad("string map values");
        m.set(e.first, e.second);
      }
      *out = std::move(m); return true;
    }
    case Type::MAPQ: {
      if (j.kind != GVal::Map) return bad("object");
      GVal m = GVal::map();
      for (const auto& e : j.kv) {
        GVal q;
        if (!convert(e.second, &registry().prim[4], &q, err, path + "." + e.first)) return false;
        if (q.is_null()) q = GVal::strct();
        m.set(e.first, std::move(q));
      }
      *out = std::move(m); return true;
    }
    case Type::STRUCT: {
      if (j.kind != GVal::Map) return bad("object");
      GVal s = GVal::strct();
      for (const auto& e : j.kv) {
        const Type* ft = nullptr;
        for (const auto& f : t->fields) if (f.first == e.first) { ft = f.second; break; }
        GVal v;
        if (ft) {
          if (!convert(e.second, ft, &v, err, path + "." + e.first)) return false;
          s.set(e.first, std::move(v));
        } else if (!e.second.is_null()) {
          s.set("?" + e.first, strip_nulls(e.second));
        }
      }
      *out = std::move(s); return true;
    }
  }
  return bad("a known type");
}

}  // namespace

bool decode_typed(const GVal& j, const std::string& type, GVal* out, std::string* err) {
  const Type* t = registry().get(type);
  if (!t) { if (err) *err = "unknown type " + type; return false; }
  if (!convert(j, t, out, err, type)) return false;
  if (out->is_null()) *out = GVal::strct();
  return true;
}

bool decode_typed_json(const std::string& json, const std::string& type, GVal* out, std::string* err) {
  GVal j;
  if (!parse_json(json, &j, err)) return false;
  if (type == "any" || type == "Unstructured") { *out = std::move(j); return true; }
  return decode_typed(j, type, out, err);
}

}  // namespace nbh
