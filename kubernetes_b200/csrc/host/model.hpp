// Flattened object model of the parts of v1.Pod / v1.Node / Service / RC / RS /
// StatefulSet that genericScheduler.Schedule reads, plus the JSON
// interchange loader used by the test harness.
//
// Reference types mirrored (staging/src/k8s.io/api/core/v1/types.go):
//   Pod.Spec{Containers,InitContainers,NodeSelector,Affinity,Tolerations,NodeName,
//   ExtendedResources(fork :2885, :4012-4056),Volumes}, Node.Spec{Taints,Unschedulable},
//   Node.Status{Allocatable,Conditions,ExtendedResources(fork :3848-3850)}.
// Quantities are carried as int64 in the unit the scheduler uses
// (schedulercache/node_info.go:90-111): cpu in milli-cores, everything else Value().
#pragma once
#include <cstdint>
#include <map>
#include <string>
#include <utility>
#include <vector>

#include "json.hpp"

namespace ksb {

using StrMap = std::map<std::string, std::string>;
// name -> quantity. "cpu" is in milli-units. Presence of a key matters
// (priorities/util/non_zero.go:36-50), so this is a map and not a struct.
using ResourceList = std::map<std::string, int64_t>;

// Well-known names (api/core/v1/types.go:3990-4003, kubelet/apis/well_known_labels.go).
static const char* const kResCPU = "cpu";
static const char* const kResMemory = "memory";
static const char* const kResPods = "pods";
static const char* const kResEphemeral = "ephemeral-storage";
static const char* const kResAlphaGPU = "alpha.kubernetes.io/nvidia-gpu";
static const char* const kLabelHostname = "kubernetes.io/hostname";
static const char* const kLabelZone = "failure-domain.beta.kubernetes.io/zone";
static const char* const kLabelRegion = "failure-domain.beta.kubernetes.io/region";

struct SelectorRequirement {  // NodeSelectorRequirement / LabelSelectorRequirement / ResourceSelector
  std::string key;
  std::string op;  // In NotIn Exists DoesNotExist Gt Lt
  std::vector<std::string> values;
};

struct LabelSelector {  // metav1.LabelSelector; `present=false` is a nil pointer
  bool present = false;
  StrMap match_labels;
  std::vector<SelectorRequirement> match_expressions;
};

struct NodeSelectorTerm {
  std::vector<SelectorRequirement> match_expressions;
};

struct PreferredSchedulingTerm {
  int32_t weight = 0;
  NodeSelectorTerm preference;
};

struct NodeAffinity {
  bool present = false;
  bool has_required = false;  // RequiredDuringSchedulingIgnoredDuringExecution != nil
  std::vector<NodeSelectorTerm> required_terms;
  std::vector<PreferredSchedulingTerm> preferred;
};

struct PodAffinityTerm {
  LabelSelector selector;
  std::vector<std::string> namespaces;
  std::string topology_key;
};

struct WeightedPodAffinityTerm {
  int32_t weight = 0;
  PodAffinityTerm term;
};

struct PodAffinity {  // shared shape for PodAffinity and PodAntiAffinity
  bool present = false;
  std::vector<PodAffinityTerm> required;
  std::vector<WeightedPodAffinityTerm> preferred;
};

struct Affinity {
  bool present = false;
  NodeAffinity node;
  PodAffinity pod;
  PodAffinity anti;
};

struct Toleration {
  std::string key, op, value, effect;
};

struct Taint {
  std::string key, value, effect;
};

struct ContainerPort {
  int32_t host_port = 0;
  std::string protocol;
  std::string host_ip;
};

struct Container {
  ResourceList requests;
  ResourceList limits;
  std::vector<ContainerPort> ports;
};

struct PodExtendedResource {  // fork: types.go:4036-4056
  std::string name;
  ResourceList limits;  // Resources.Limits; must have exactly one key
  std::vector<SelectorRequirement> required;  // Affinity.Required
  std::vector<std::string> assigned;
};

struct OwnerRef {
  bool present = false;  // a controller=true owner reference exists
  std::string kind, uid;
};

struct Pod {
  std::string ns, name;
  StrMap labels;
  bool deleting = false;  // DeletionTimestamp != nil
  std::string node_name;
  OwnerRef controller;
  std::vector<Container> containers;
  std::vector<Container> init_containers;
  StrMap node_selector;
  Affinity affinity;
  std::vector<Toleration> tolerations;
  std::vector<PodExtendedResource> extended;
  int32_t num_volumes = 0;
  int32_t priority = 0;  // util.GetPodPriority (plugin/pkg/scheduler/util/utils.go:76-87): *Spec.Priority, 0 when unset
  std::vector<std::string> images;  // Spec.Containers[i].Image in order (priorities/image_locality.go:66-80)
  std::string key() const { return ns + "/" + name; }
};

struct NodeCondition {
  std::string type, status;
};

struct ExtendedDevice {  // fork: v1.ExtendedResource (types.go:4012-4020)
  std::string id;
  std::string health;
  StrMap attributes;
};

struct AvoidEntry {
  std::string kind, uid;
};

struct NodeImage {  // v1.ContainerImage
  std::vector<std::string> names;
  int64_t size_bytes = 0;
};

struct Node {
  std::string name;
  bool has_labels = true;  // false: Labels == nil (matters for topology helpers)
  StrMap labels;
  ResourceList allocatable;
  std::vector<Taint> taints;
  std::vector<NodeCondition> conditions;
  bool unschedulable = false;
  std::vector<AvoidEntry> avoid;  // parsed preferAvoidPods annotation
  bool avoid_annotation_invalid = false;
  // resource name -> devices
  std::map<std::string, std::vector<ExtendedDevice>> extended;
  std::vector<NodeImage> images;  // Status.Images (read by ImageLocalityPriority only)
};

struct SelectorOwner {  // Service / ReplicationController (map selector) or RS / StatefulSet (LabelSelector)
  std::string ns;
  std::string name;
  StrMap map_selector;     // services, RCs
  LabelSelector selector;  // RSs, StatefulSets
};

struct Cluster {
  std::vector<Node> nodes;
  std::vector<SelectorOwner> services, rcs, rss, sss;
  std::vector<Pod> existing_pods;  // already bound (node_name set)
  std::vector<Pod> pods;           // the stream, FIFO
};

// ---------------------------------------------------------------- JSON loading

namespace detail {

inline StrMap strmap(const json::Value* v) {
  StrMap m;
  if (!v || v->is_null()) return m;
  for (auto& kv : v->obj()) m[kv.first] = kv.second.s;
  return m;
}
inline ResourceList reslist(const json::Value* v) {
  ResourceList m;
  if (!v || v->is_null()) return m;
  for (auto& kv : v->obj()) {
    const json::Value& q = kv.second;
    m[kv.first] = q.kind == json::Value::Double ? (int64_t)q.d : q.i;
  }
  return m;
}
inline std::vector<std::string> strlist(const json::Value* v) {
  std::vector<std::string> out;
  if (!v || v->is_null()) return out;
  for (auto& e : v->arr()) out.push_back(e.s);
  return out;
}
inline std::vector<SelectorRequirement> reqs(const json::Value* v) {
  std::vector<SelectorRequirement> out;
  if (!v || v->is_null()) return out;
  for (auto& e : v->arr()) {
    SelectorRequirement r;
    r.key = e.get_str("key");
    r.op = e.get_str("op");
    r.values = strlist(e.find("values"));
    out.push_back(std::move(r));
  }
  return out;
}
inline LabelSelector label_selector(const json::Value* v) {
  LabelSelector s;
  if (!v || v->is_null()) return s;
  s.present = true;
  s.match_labels = strmap(v->find("matchLabels"));
  s.match_expressions = reqs(v->find("matchExpressions"));
  return s;
}
inline PodAffinityTerm pod_term(const json::Value& v) {
  PodAffinityTerm t;
  t.selector = label_selector(v.find("selector"));
  t.namespaces = strlist(v.find("namespaces"));
  t.topology_key = v.get_str("topologyKey");
  return t;
}
inline PodAffinity pod_affinity(const json::Value* v) {
  PodAffinity a;
  if (!v || v->is_null()) return a;
  a.present = true;
  for (auto& e : v->get_arr("required")) a.required.push_back(pod_term(e));
  for (auto& e : v->get_arr("preferred")) {
    WeightedPodAffinityTerm w;
    w.weight = (int32_t)e.get_int("weight");
    const json::Value* t = e.find("term");
    if (t) w.term = pod_term(*t);
    a.preferred.push_back(std::move(w));
  }
  return a;
}
inline Container container(const json::Value& v) {
  Container c;
  c.requests = reslist(v.find("requests"));
  c.limits = reslist(v.find("limits"));
  for (auto& p : v.get_arr("ports")) {
    ContainerPort cp;
    cp.host_port = (int32_t)p.get_int("hostPort");
    cp.protocol = p.get_str("protocol");
    cp.host_ip = p.get_str("hostIP");
    c.ports.push_back(std::move(cp));
  }
  return c;
}

}  // namespace detail

inline Pod pod_from_json(const json::Value& v) {
  using namespace detail;
  Pod p;
  p.ns = v.get_str("namespace");
  p.name = v.get_str("name");
  p.labels = strmap(v.find("labels"));
  p.deleting = v.get_bool("deleting");
  p.node_name = v.get_str("nodeName");
  if (const json::Value* o = v.find("controller"); o && !o->is_null()) {
    p.controller.present = true;
    p.controller.kind = o->get_str("kind");
    p.controller.uid = o->get_str("uid");
  }
  for (auto& c : v.get_arr("containers")) {
    p.containers.push_back(container(c));
    p.images.push_back(c.get_str("image"));
  }
  for (auto& c : v.get_arr("initContainers")) p.init_containers.push_back(container(c));
  p.node_selector = strmap(v.find("nodeSelector"));
  if (const json::Value* a = v.find("affinity"); a && !a->is_null()) {
    p.affinity.present = true;
    if (const json::Value* na = a->find("node"); na && !na->is_null()) {
      p.affinity.node.present = true;
      if (const json::Value* r = na->find("required"); r && !r->is_null()) {
        p.affinity.node.has_required = true;
        for (auto& t : r->arr()) {
          NodeSelectorTerm term;
          term.match_expressions = reqs(t.find("matchExpressions"));
          p.affinity.node.required_terms.push_back(std::move(term));
        }
      }
      for (auto& t : na->get_arr("preferred")) {
        PreferredSchedulingTerm pt;
        pt.weight = (int32_t)t.get_int("weight");
        pt.preference.match_expressions = reqs(t.find("matchExpressions"));
        p.affinity.node.preferred.push_back(std::move(pt));
      }
    }
    p.affinity.pod = pod_affinity(a->find("podAffinity"));
    p.affinity.anti = pod_affinity(a->find("podAntiAffinity"));
  }
  for (auto& t : v.get_arr("tolerations")) {
    Toleration tol;
    tol.key = t.get_str("key");
    tol.op = t.get_str("operator");
    tol.value = t.get_str("value");
    tol.effect = t.get_str("effect");
    p.tolerations.push_back(std::move(tol));
  }
  for (auto& e : v.get_arr("extendedResources")) {
    PodExtendedResource r;
    r.name = e.get_str("name");
    r.limits = reslist(e.find("limits"));
    r.required = reqs(e.find("required"));
    r.assigned = strlist(e.find("assigned"));
    p.extended.push_back(std::move(r));
  }
  p.num_volumes = (int32_t)v.get_int("volumes");
  p.priority = (int32_t)v.get_int("priority");
  return p;
}

inline Node node_from_json(const json::Value& v) {
  using namespace detail;
  Node n;
  n.name = v.get_str("name");
  const json::Value* l = v.find("labels");
  n.has_labels = l && !l->is_null();
  n.labels = strmap(l);
  n.allocatable = reslist(v.find("allocatable"));
  for (auto& t : v.get_arr("taints")) {
    Taint tt;
    tt.key = t.get_str("key");
    tt.value = t.get_str("value");
    tt.effect = t.get_str("effect");
    n.taints.push_back(std::move(tt));
  }
  for (auto& c : v.get_arr("conditions")) {
    NodeCondition nc;
    nc.type = c.get_str("type");
    nc.status = c.get_str("status");
    n.conditions.push_back(std::move(nc));
  }
  n.unschedulable = v.get_bool("unschedulable");
  for (auto& a : v.get_arr("avoid")) {
    AvoidEntry e;
    e.kind = a.get_str("kind");
    e.uid = a.get_str("uid");
    n.avoid.push_back(std::move(e));
  }
  n.avoid_annotation_invalid = v.get_bool("avoidInvalid");
  if (const json::Value* er = v.find("extendedResources"); er && !er->is_null()) {
    for (auto& kv : er->obj()) {
      std::vector<ExtendedDevice> devs;
      for (auto& d : kv.second.arr()) {
        ExtendedDevice dev;
        dev.id = d.get_str("id");
        dev.health = d.get_str("health");
        dev.attributes = strmap(d.find("attributes"));
        devs.push_back(std::move(dev));
      }
      n.extended[kv.first] = std::move(devs);
    }
  }
  for (auto& im : v.get_arr("images")) {
    NodeImage x;
    x.names = strlist(im.find("names"));
    x.size_bytes = im.get_int("sizeBytes");
    n.images.push_back(std::move(x));
  }
  return n;
}

inline SelectorOwner owner_from_json(const json::Value& v, bool label_selector_kind) {
  SelectorOwner s;
  s.ns = v.get_str("namespace");
  s.name = v.get_str("name");
  if (label_selector_kind)
    s.selector = detail::label_selector(v.find("selector"));
  else
    s.map_selector = detail::strmap(v.find("selector"));
  return s;
}

inline Cluster cluster_from_json(const json::Value& v) {
  Cluster c;
  for (auto& n : v.get_arr("nodes")) c.nodes.push_back(node_from_json(n));
  for (auto& s : v.get_arr("services")) c.services.push_back(owner_from_json(s, false));
  for (auto& s : v.get_arr("rcs")) c.rcs.push_back(owner_from_json(s, false));
  for (auto& s : v.get_arr("rss")) c.rss.push_back(owner_from_json(s, true));
  for (auto& s : v.get_arr("sss")) c.sss.push_back(owner_from_json(s, true));
  for (auto& p : v.get_arr("existingPods")) c.existing_pods.push_back(pod_from_json(p));
  for (auto& p : v.get_arr("pods")) c.pods.push_back(pod_from_json(p));
  return c;
}

}  // namespace ksb
