// ORACLE — TEST INFRASTRUCTURE ONLY. CPU restatement ("literal mode": the
// reference's own control flow and complexity) of genericScheduler.Schedule with
// the DefaultProvider predicates and priorities plus the fork's device matching.
// Nothing under kubernetes_b200/ may include or link this.
//
// Follows (paths relative to /root/reference/plugin/pkg/scheduler unless noted):
//   core/generic_scheduler.go:109-161 (Schedule), :177-192 (selectHost),
//     :283-362 (findNodesThatFit), :402-501 (podFitsOnNode, single pass: FIFO queue),
//     :509-640 (PrioritizeNodes); :70-88 (FitError.Error)
//   core/extended_resources.go:42-183 (fork; device iteration fixed to ascending ID)
//   algorithm/predicates/predicates.go:583-687 (resources), :691-776 (selector, host),
//     :925-945 (ports), :965-1029 (GeneralPredicates), :1047-1346 (inter-pod affinity),
//     :1349-1441 (taints, pressure, conditions); metadata.go:80-101; utils.go:138-202
//   algorithm/priorities/{least_requested,balanced_resource_allocation,selector_spreading,
//     node_affinity,taint_toleration,node_prefer_avoid_pods,interpod_affinity,reduce,metadata}.go
//   algorithm/priorities/util/{non_zero,topologies,util}.go; util/utils.go:33-62
//   api/types.go:32 (MaxPriority), :204-227 (HostPriorityList.Less)
//   algorithmprovider/defaults/defaults.go:119-184, :217-257 (sets + weights)
//   testing/fake_lister.go:33-178 (in-memory listers the harness mirrors)
//   /root/reference/staging/src/k8s.io/api/core/v1/toleration.go:37-56
//   /root/reference/pkg/apis/core/v1/helper/helpers.go:280-308, qos/qos.go:38-100
//   /root/reference/pkg/util/node/node.go:116-133 (GetZoneKey)
//
// float64 arithmetic: compile with -ffp-contract=off so no FMA is formed (Go/amd64
// never fuses); int(f) is truncation toward zero.
#pragma once
#include <algorithm>
#include <atomic>
#include <cmath>
#include <functional>
#include <map>
#include <set>
#include <string>
#include <thread>
#include <vector>

#include "cache.hpp"
#include "labels.hpp"
#include "model.hpp"

namespace oracle {

static const int kMaxPriority = 10;  // api/types.go:32

struct HostPriority {
  std::string host;
  int64_t score = 0;
};
using HostPriorityList = std::vector<HostPriority>;

using NodeInfoMap = std::map<std::string, NodeInfo>;
using NodePtr = std::shared_ptr<const Node>;
using PodPtr = std::shared_ptr<const Pod>;

// workqueue.Parallelize (client-go/util/workqueue/parallelizer.go:29-53)
inline void parallelize(int workers, int pieces, const std::function<void(int)>& fn) {
  if (workers <= 1 || pieces <= 1) {
    for (int i = 0; i < pieces; ++i) fn(i);
    return;
  }
  std::atomic<int> next(0);
  std::vector<std::thread> th;
  int w = std::min(workers, pieces);
  for (int t = 0; t < w; ++t)
    th.emplace_back([&] {
      for (;;) {
        int i = next.fetch_add(1);
        if (i >= pieces) return;
        fn(i);
      }
    });
  for (auto& t : th) t.join();
}

// ------------------------------------------------------------------ helpers

// toleration.go:37-56
inline bool tolerates_taint(const ksb::Toleration& t, const ksb::Taint& taint) {
  if (!t.effect.empty() && t.effect != taint.effect) return false;
  if (!t.key.empty() && t.key != taint.key) return false;
  if (t.op.empty() || t.op == "Equal") return t.value == taint.value;
  if (t.op == "Exists") return true;
  return false;
}
inline bool tolerations_tolerate_taint(const std::vector<ksb::Toleration>& tols, const ksb::Taint& taint) {
  for (auto& t : tols)
    if (tolerates_taint(t, taint)) return true;
  return false;
}

// pkg/util/node/node.go:116-133
inline std::string get_zone_key(const Node& n) {
  if (!n.has_labels) return "";
  auto r = n.labels.find(ksb::kLabelRegion);
  auto z = n.labels.find(ksb::kLabelZone);
  std::string region = r == n.labels.end() ? "" : r->second;
  std::string zone = z == n.labels.end() ? "" : z->second;
  if (region.empty() && zone.empty()) return "";
  return region + std::string(":\x00:", 3) + zone;
}

// priorities/util/topologies.go:51-71
inline bool nodes_have_same_topology_key(const Node& a, const Node& b, const std::string& key) {
  if (key.empty()) return false;
  if (!a.has_labels || !b.has_labels) return false;
  auto ia = a.labels.find(key);
  auto ib = b.labels.find(key);
  if (ia != a.labels.end() && ib != b.labels.end()) return ia->second == ib->second;
  return false;
}

// topologies.go:28-37
inline std::set<std::string> namespaces_from_term(const Pod& pod, const ksb::PodAffinityTerm& term) {
  std::set<std::string> names;
  if (term.namespaces.empty()) names.insert(pod.ns);
  else names.insert(term.namespaces.begin(), term.namespaces.end());
  return names;
}
// topologies.go:41-50
inline bool pod_matches_ns_and_selector(const Pod& pod, const std::set<std::string>& nss, const Selector& sel) {
  if (!nss.count(pod.ns)) return false;
  return sel_matches(sel, pod.labels);
}

// util/utils.go:33-62
inline std::set<std::string> get_used_ports(const Pod& pod) {
  std::set<std::string> ports;
  for (auto& c : pod.containers)
    for (auto& p : c.ports) {
      if (p.host_port == 0) continue;
      std::string proto = p.protocol.empty() ? "TCP" : p.protocol;
      std::string ip = p.host_ip.empty() ? "0.0.0.0" : p.host_ip;
      ports.insert(proto + "/" + ip + "/" + std::to_string(p.host_port));
    }
  return ports;
}

// predicates/utils.go:138-202
struct HostPortInfo {
  std::string protocol, hostIP, hostPort;
};
inline HostPortInfo decode_port(const std::string& s) {
  HostPortInfo h;
  size_t a = s.find('/');
  size_t b = s.find('/', a + 1);
  h.protocol = s.substr(0, a);
  h.hostIP = s.substr(a + 1, b - a - 1);
  h.hostPort = s.substr(b + 1);
  return h;
}
inline bool special_port_conflict_check(const std::string& special, const std::set<std::string>& others) {
  HostPortInfo sp = decode_port(special);
  if (sp.hostIP == "0.0.0.0") {
    for (auto& item : others) {
      HostPortInfo h = decode_port(item);
      if (sp.hostPort == h.hostPort && sp.protocol == h.protocol) return true;
    }
  }
  return false;
}
inline bool ports_conflict(const std::set<std::string>& existing, const std::set<std::string>& want) {
  for (auto& e : existing)
    if (special_port_conflict_check(e, want)) return true;
  for (auto& w : want) {
    if (special_port_conflict_check(w, existing)) return true;
    if (existing.count(w)) return true;
  }
  return false;
}

// qos.go:38-100 reduced to the BestEffort test the scheduler uses (predicates.go:1376-1379):
// no container has a cpu / memory / hugepages-* request or limit > 0.
inline bool is_pod_best_effort(const Pod& pod) {
  auto supported = [](const std::string& n) { return n == "cpu" || n == "memory" || is_hugepage_resource(n); };
  for (auto& c : pod.containers) {
    for (auto& kv : c.requests)
      if (supported(kv.first) && kv.second > 0) return false;
    for (auto& kv : c.limits)
      if (supported(kv.first) && kv.second > 0) return false;
  }
  return true;
}

// predicates.go:583-621
inline Resource get_resource_request(const Pod& pod) {
  Resource result;
  for (auto& c : pod.containers) resource_add(result, c.requests);
  for (auto& c : pod.init_containers)
    for (auto& kv : c.requests) {
      const std::string& n = kv.first;
      int64_t q = kv.second;
      if (n == ksb::kResMemory) {
        if (q > result.Memory) result.Memory = q;
      } else if (n == ksb::kResEphemeral) {
        if (q > result.EphemeralStorage) result.EphemeralStorage = q;
      } else if (n == ksb::kResCPU) {
        if (q > result.MilliCPU) result.MilliCPU = q;
      } else if (n == ksb::kResAlphaGPU) {
        if (q > result.NvidiaGPU) result.NvidiaGPU = q;
      } else if (is_scalar_resource_name(n)) {
        auto it = result.Scalar.find(n);
        int64_t cur = it == result.Scalar.end() ? 0 : it->second;
        if (q > cur) result.Scalar[n] = q;
      }
    }
  return result;
}

// ------------------------------------------------------------------ listers (testing/fake_lister.go)

struct Listers {
  std::vector<ksb::SelectorOwner> services, rcs, rss, sss;
};

// priorities/metadata.go:69-96 over the fake listers (fake_lister.go:80-178)
inline std::vector<Selector> get_selectors(const Pod& pod, const Listers& L) {
  std::vector<Selector> out;
  for (auto& s : L.services) {
    if (s.ns != pod.ns) continue;
    Selector sel = selector_from_validated_set(s.map_selector);
    if (sel_matches(sel, pod.labels)) out.push_back(selector_from_set(s.map_selector));
  }
  for (auto& s : L.rcs) {
    if (s.ns != pod.ns) continue;
    Selector sel = selector_from_validated_set(s.map_selector);
    if (sel_matches(sel, pod.labels)) out.push_back(selector_from_set(s.map_selector));
  }
  {
    // GetPodReplicaSets: a selector conversion error aborts the whole lookup with err => no RS selectors.
    std::vector<Selector> tmp;
    bool ok = true;
    for (auto& s : L.rss) {
      if (s.ns != pod.ns) continue;
      Selector sel;
      std::string err;
      if (!label_selector_as_selector(s.selector, &sel, &err)) {
        ok = false;
        break;
      }
      if (sel_matches(sel, pod.labels)) tmp.push_back(sel);
    }
    if (ok)
      for (auto& s : tmp) out.push_back(s);
  }
  {
    std::vector<Selector> tmp;
    bool ok = true;
    for (auto& s : L.sss) {
      if (s.ns != pod.ns) continue;
      Selector sel;
      std::string err;
      if (!label_selector_as_selector(s.selector, &sel, &err)) {
        ok = false;
        break;
      }
      if (sel_matches(sel, pod.labels)) tmp.push_back(sel);
    }
    if (ok)
      for (auto& s : tmp) out.push_back(s);
  }
  return out;
}

// ------------------------------------------------------------------ predicate metadata

struct MatchingAntiAffinityTerm {
  std::string topology_key;
  NodePtr node;
};

struct PredicateMeta {
  bool bestEffort = false;
  Resource podRequest;
  std::set<std::string> podPorts;
  // flattened: the reference keys by existing-pod name but only iterates all entries
  std::vector<MatchingAntiAffinityTerm> matchingAntiAffinityTerms;
  bool error = false;
  std::string error_msg;
};

// predicates.go:1108-1119 — required terms only
inline const std::vector<ksb::PodAffinityTerm>& required_terms(const ksb::PodAffinity& a) {
  static const std::vector<ksb::PodAffinityTerm> empty;
  return a.present ? a.required : empty;
}

struct Oracle;  // fwd

// ------------------------------------------------------------------ result types

struct ScheduleResult {
  enum Status { OK = 0, NoNodes = 1, FitError = 2, Error = 3 } status = OK;
  std::string host;
  // ExtendedResourceBinding: pod.Spec.ExtendedResources[i].Name -> device ids
  std::map<std::string, std::vector<std::string>> binding;
  std::string message;  // FitError.Error() text or error text
  int num_filtered = 0;
  int64_t score = 0;  // total score of the chosen host (0 when the 1-fit shortcut is taken)
};

struct PriorityWeights {
  int64_t least_requested = 1, balanced = 1, selector_spread = 1, node_affinity = 1, taint_toleration = 1,
          inter_pod_affinity = 1, prefer_avoid = 10000;
  // not in the DefaultProvider set (registered by name, enabled through a Policy: defaults.go:93-116, api/types.go:36-125)
  int64_t most_requested = 0, resource_limits = 0, image_locality = 0;
  int64_t service_spreading = 0;  // defaults.go:90-102: SelectorSpread's Map / Reduce; with a priorityMetadata (always, in PrioritizeNodes) the same selectors too (selector_spreading.go:71-76)
  int64_t equal = 0;              // generic_scheduler.go:643-652: 1 on every node
};

// Policy-configured plugins that take arguments (api/types.go:64-125): NodeLabelPriority instances
// (priorities/node_label.go:29-60) and the CheckNodeLabelPresence predicate (predicates.go:778-826).
struct NodeLabelPriorityArg {
  std::string label;
  bool presence = true;
  int64_t weight = 1;
};
// The verdict of the scheduler extenders for the pods of a run (core/extender.go; the HTTP services themselves are outside
// this path): nodes their Filter calls kept (empty + has_filter=false: no Filter verdict) and their weighted scores.
struct ExtenderVerdict {
  bool has_filter = false;
  std::set<std::string> keep;
  std::map<std::string, int64_t> scores;  // combinedScores: sum of weight * score
};

struct Policy {
  std::vector<NodeLabelPriorityArg> node_label;
  std::vector<std::string> presence_labels;
  bool presence = true;
};

struct Oracle {
  SchedulerCache cache;
  Listers listers;
  NodeInfoMap cachedNodeInfoMap;  // snapshot taken at Schedule entry (generic_scheduler.go:122)
  uint64_t lastNodeIndex = 0;
  int workers = 1;  // 16 mirrors Parallelize(16, ...)
  int hardPodAffinityWeight = 1;
  PriorityWeights weights;
  Policy policy;
  ExtenderVerdict extender;

  // ---------------- predicates

  // predicates.go:630-687
  static void pod_fits_resources(const PredicateMeta& meta, const NodeInfo& ni, std::vector<std::string>* fails) {
    int allowed = ni.allocatable.AllowedPodNumber;
    if ((int)ni.pods.size() + 1 > allowed) fails->push_back("Insufficient pods");
    const Resource& pr = meta.podRequest;
    if (pr.MilliCPU == 0 && pr.Memory == 0 && pr.NvidiaGPU == 0 && pr.EphemeralStorage == 0 && pr.Scalar.empty()) return;
    const Resource& al = ni.allocatable;
    const Resource& rq = ni.requested;
    if (al.MilliCPU < pr.MilliCPU + rq.MilliCPU) fails->push_back("Insufficient cpu");
    if (al.Memory < pr.Memory + rq.Memory) fails->push_back("Insufficient memory");
    if (al.NvidiaGPU < pr.NvidiaGPU + rq.NvidiaGPU) fails->push_back(std::string("Insufficient ") + ksb::kResAlphaGPU);
    if (al.EphemeralStorage < pr.EphemeralStorage + rq.EphemeralStorage)
      fails->push_back(std::string("Insufficient ") + ksb::kResEphemeral);
    for (auto& kv : pr.Scalar) {
      auto a = al.Scalar.find(kv.first);
      auto r = rq.Scalar.find(kv.first);
      int64_t av = a == al.Scalar.end() ? 0 : a->second;
      int64_t rv = r == rq.Scalar.end() ? 0 : r->second;
      if (av < kv.second + rv) fails->push_back("Insufficient " + kv.first);
    }
  }

  // predicates.go:691-703
  static bool node_matches_node_selector_terms(const Node& node, const std::vector<ksb::NodeSelectorTerm>& terms) {
    for (auto& t : terms) {
      Selector sel;
      std::string err;
      if (!requirements_as_selector(t.match_expressions, &sel, &err)) return false;
      if (sel_matches(sel, node.labels)) return true;
    }
    return false;
  }
  // predicates.go:706-749
  static bool pod_matches_node_labels(const Pod& pod, const Node& node) {
    if (!pod.node_selector.empty()) {
      Selector sel = selector_from_set(pod.node_selector);
      if (!sel_matches(sel, node.labels)) return false;
    }
    bool matches = true;
    if (pod.affinity.present && pod.affinity.node.present) {
      const ksb::NodeAffinity& na = pod.affinity.node;
      if (!na.has_required) return true;
      matches = matches && node_matches_node_selector_terms(node, na.required_terms);
    }
    return matches;
  }

  // predicates.go:1414-1441
  static void check_node_condition(const NodeInfo& ni, std::vector<std::string>* fails) {
    if (!ni.node) {
      fails->push_back("NodeUnknownCondition");
      return;
    }
    for (auto& c : ni.node->conditions) {
      if (c.type == "Ready" && c.status != "True") fails->push_back("NodeNotReady");
      else if (c.type == "OutOfDisk" && c.status != "False") fails->push_back("NodeOutOfDisk");
      else if (c.type == "NetworkUnavailable" && c.status != "False") fails->push_back("NodeNetworkUnavailable");
    }
    if (ni.node->unschedulable) fails->push_back("NodeUnschedulable");
  }

  // all pods in the live cache (schedulerCache.FilteredList, cache.go:104-116). The
  // nodeInfo.Filter argument only drops pods bound to this node that the snapshot does
  // not hold, which cannot happen in the single-threaded harness.
  std::vector<PodPtr> all_pods() const {
    std::vector<PodPtr> out;
    for (auto& kv : cache.nodes)
      for (auto& p : kv.second.pods) out.push_back(p);
    return out;
  }
  NodePtr get_node(const std::string& name) const {  // predicates.NodeInfo.GetNodeInfo
    auto it = cachedNodeInfoMap.find(name);
    if (it == cachedNodeInfoMap.end()) return nullptr;
    return it->second.node;
  }

  // metadata.go:80-101 + predicates.go:1132-1190. Reproduces the reference's aliasing of
  // the range variable: every entry recorded for one existing pod points at the LAST
  // term of that pod's list (`&term` of a Go <1.22 loop variable, predicates.go:1166-1177).
  PredicateMeta make_predicate_meta(const Pod& pod) const {
    PredicateMeta m;
    m.bestEffort = is_pod_best_effort(pod);
    m.podRequest = get_resource_request(pod);
    m.podPorts = get_used_ports(pod);
    for (auto& kv : cachedNodeInfoMap) {
      const NodeInfo& ni = kv.second;
      if (!ni.node) {
        m.error = true;
        m.error_msg = "node not found";
        return m;
      }
      for (auto& ep : ni.podsWithAffinity) {
        if (!ep->affinity.present) continue;
        const auto& terms = required_terms(ep->affinity.anti);
        for (size_t i = 0; i < terms.size(); ++i) {
          auto nss = namespaces_from_term(*ep, terms[i]);
          Selector sel;
          std::string err;
          if (!label_selector_as_selector(terms[i].selector, &sel, &err)) {
            m.error = true;
            m.error_msg = err;
            return m;
          }
          if (pod_matches_ns_and_selector(pod, nss, sel))
            m.matchingAntiAffinityTerms.push_back({terms.back().topology_key, ni.node});
        }
      }
    }
    return m;
  }

  struct TermProbe {  // a pod-affinity term with its selector parsed once per Schedule
    const ksb::PodAffinityTerm* term;
    std::set<std::string> nss;
    Selector sel;
    bool sel_err = false;
  };
  static TermProbe probe(const Pod& owner, const ksb::PodAffinityTerm& t) {
    TermProbe p;
    p.term = &t;
    p.nss = namespaces_from_term(owner, t);
    std::string err;
    p.sel_err = !label_selector_as_selector(t.selector, &p.sel, &err);
    return p;
  }

  // predicates.go:1080-1104. returns 0 ok; sets err on error
  bool any_pod_matches_term(const TermProbe& tp, const std::vector<PodPtr>& allPods, const Node& node,
                            bool* termMatches, bool* matchingPodExists) const {
    *termMatches = false;
    *matchingPodExists = false;
    if (tp.term->topology_key.empty()) return false;  // error
    if (tp.sel_err) return false;
    for (auto& ep : allPods) {
      if (!pod_matches_ns_and_selector(*ep, tp.nss, tp.sel)) continue;
      *matchingPodExists = true;
      NodePtr en = get_node(ep->node_name);
      if (!en) return false;  // GetNodeInfo error
      if (nodes_have_same_topology_key(node, *en, tp.term->topology_key)) {
        *termMatches = true;
        return true;
      }
    }
    return true;
  }

  // predicates.go:1047-1074 (+ :1238-1346). Returns false with *err set on a hard error.
  bool inter_pod_affinity_matches(const Pod& pod, const PredicateMeta& meta, const NodeInfo& ni,
                                  const std::vector<PodPtr>& allPods, const std::vector<TermProbe>& affTerms,
                                  const std::vector<TermProbe>& antiTerms, std::vector<std::string>* fails,
                                  std::string* err) const {
    const Node& node = *ni.node;
    // satisfiesExistingPodsAntiAffinity :1238-1283
    for (auto& t : meta.matchingAntiAffinityTerms) {
      if (t.topology_key.empty()) {
        *err = "Empty topologyKey is not allowed except for PreferredDuringScheduling pod anti-affinity";
        return false;
      }
      if (nodes_have_same_topology_key(node, *t.node, t.topology_key)) {
        fails->push_back("MatchInterPodAffinity");
        fails->push_back("ExistingPodsAntiAffinityRulesNotMatch");
        return true;
      }
    }
    if (!pod.affinity.present || (!pod.affinity.pod.present && !pod.affinity.anti.present)) return true;
    // satisfiesPodsAffinityAntiAffinity :1286-1346
    for (auto& tp : affTerms) {
      bool termMatches, exists;
      if (!any_pod_matches_term(tp, allPods, node, &termMatches, &exists)) {
        // :1293-1297 returns ErrPodAffinityRulesNotMatch WITH the error; InterPodAffinityMatches (:1155-1160) puts
        // ErrPodAffinityNotMatch in front and hands both on (podFitsOnNode then fails the whole Schedule call)
        fails->push_back("MatchInterPodAffinity");
        fails->push_back("PodAffinityRulesNotMatch");
        *err = "Cannot schedule pod onto node, because of PodAffinityTerm";
        return false;
      }
      if (!termMatches) {
        if (exists) {
          fails->push_back("MatchInterPodAffinity");
          fails->push_back("PodAffinityRulesNotMatch");
          return true;
        }
        if (!pod_matches_ns_and_selector(pod, tp.nss, tp.sel)) {
          fails->push_back("MatchInterPodAffinity");
          fails->push_back("PodAffinityRulesNotMatch");
          return true;
        }
      }
    }
    for (auto& tp : antiTerms) {
      bool termMatches, exists;
      bool ok = any_pod_matches_term(tp, allPods, node, &termMatches, &exists);
      if (!ok || termMatches) {
        fails->push_back("MatchInterPodAffinity");
        fails->push_back("PodAntiAffinityRulesNotMatch");
        return true;
      }
    }
    return true;
  }

  // ---------------- indexed mode (SURVEY.md §8c "two oracle modes")
  // The literal functions above cost O(nodes x pods) per scheduled pod: every node rescans every pod. Indexed mode
  // scans the pods once per Schedule and keeps, per affinity term, the set of topology values that hold a matching
  // pod; a node then needs one label lookup per term. Same decisions by construction; any situation in which the
  // literal code's answer depends on its scan order (a hard error: empty topology key, unparsable selector, a pod on
  // an unknown node) switches that Schedule call back to the literal functions. tests/test_oracle_indexed.py
  // checks that both modes give the same results on seeded instances.
  bool indexed = false;
  struct TermIndex {
    bool exists = false;           // some pod matches the term's namespaces + selector
    std::set<std::string> values;  // topology values (under the term's key) of the nodes holding such pods
  };
  struct IpaIndex {
    bool usable = false;
    std::map<std::string, std::set<std::string>> anti_by_key;  // existing pods' matching anti-affinity terms
    std::vector<TermIndex> aff, anti;                         // the incoming pod's own required terms
  };
  static const std::string* topo_value(const Node& n, const std::string& key) {
    if (key.empty() || !n.has_labels) return nullptr;
    auto it = n.labels.find(key);
    return it == n.labels.end() ? nullptr : &it->second;
  }
  IpaIndex build_ipa_index(const PredicateMeta& meta, const std::vector<PodPtr>& allPods, const std::vector<TermProbe>& affTerms,
                           const std::vector<TermProbe>& antiTerms) const {
    IpaIndex ix;
    for (auto& t : meta.matchingAntiAffinityTerms) {
      if (t.topology_key.empty()) return ix;  // the literal path raises an error at the first such term it meets
      if (const std::string* v = topo_value(*t.node, t.topology_key)) ix.anti_by_key[t.topology_key].insert(*v);
    }
    auto index_terms = [&](const std::vector<TermProbe>& probes, std::vector<TermIndex>* out) {
      for (auto& tp : probes) {
        if (tp.term->topology_key.empty() || tp.sel_err) return false;
        TermIndex ti;
        for (auto& ep : allPods) {
          if (!pod_matches_ns_and_selector(*ep, tp.nss, tp.sel)) continue;
          ti.exists = true;
          NodePtr en = get_node(ep->node_name);
          if (!en) return false;
          if (const std::string* v = topo_value(*en, tp.term->topology_key)) ti.values.insert(*v);
        }
        out->push_back(std::move(ti));
      }
      return true;
    };
    if (!index_terms(affTerms, &ix.aff) || !index_terms(antiTerms, &ix.anti)) return ix;
    ix.usable = true;
    return ix;
  }
  // inter_pod_affinity_matches over the index (never errors: the index is only usable when the literal path cannot)
  void inter_pod_affinity_matches_indexed(const Pod& pod, const NodeInfo& ni, const IpaIndex& ix,
                                          const std::vector<TermProbe>& affTerms, const std::vector<TermProbe>& antiTerms,
                                          std::vector<std::string>* fails) const {
    const Node& node = *ni.node;
    auto in_domain = [&](const std::string& key, const std::set<std::string>& values) {
      const std::string* v = topo_value(node, key);
      return v && values.count(*v);
    };
    for (auto& kv : ix.anti_by_key)
      if (in_domain(kv.first, kv.second)) {
        fails->push_back("MatchInterPodAffinity");
        fails->push_back("ExistingPodsAntiAffinityRulesNotMatch");
        return;
      }
    if (!pod.affinity.present || (!pod.affinity.pod.present && !pod.affinity.anti.present)) return;
    for (size_t i = 0; i < ix.aff.size(); ++i) {
      const TermProbe& tp = affTerms[i];
      if (in_domain(tp.term->topology_key, ix.aff[i].values)) continue;
      if (ix.aff[i].exists || !pod_matches_ns_and_selector(pod, tp.nss, tp.sel)) {
        fails->push_back("MatchInterPodAffinity");
        fails->push_back("PodAffinityRulesNotMatch");
        return;
      }
    }
    for (size_t i = 0; i < ix.anti.size(); ++i)
      if (in_domain(antiTerms[i].term->topology_key, ix.anti[i].values)) {
        fails->push_back("MatchInterPodAffinity");
        fails->push_back("PodAntiAffinityRulesNotMatch");
        return;
      }
  }

  // ---------------- fork: device matching (core/extended_resources.go)

  // :152-167
  static bool is_device_a_match(const ksb::ExtendedDevice& res, const std::vector<ksb::SelectorRequirement>& required,
                                bool* ok, std::string* err) {
    if (required.empty()) {
      *ok = true;
      return true;
    }
    Selector sel;
    if (!requirements_as_selector(required, &sel, err)) return false;
    *ok = sel_matches(sel, res.attributes);
    return true;
  }
  // :113-150 — device order: ascending ID (canonical order for the reference's random map order)
  static bool allocate_resources(const std::string& rname, int64_t num, const std::vector<ksb::SelectorRequirement>& required,
                                 const ExtendedResourceMap& available, std::vector<std::string>* selected,
                                 std::string* fail) {
    auto dom = available.find(rname);
    if (dom == available.end()) {
      *fail = "Insufficient " + rname;
      return false;
    }
    for (auto& kv : dom->second) {
      if ((int64_t)selected->size() == num) break;
      bool ok;
      std::string err;
      if (!is_device_a_match(kv.second, required, &ok, &err)) {
        *fail = err;
        return false;
      }
      if (ok) selected->push_back(kv.first);
    }
    if ((int64_t)selected->size() != num) {
      *fail = "Insufficient " + rname;
      return false;
    }
    return true;
  }
  // :83-111
  static bool has_extended_resources(const Pod& p, const NodeInfo& ni,
                                     std::map<std::string, std::vector<std::string>>* binding, std::string* fail) {
    ExtendedResourceMap available = ni.ext.available;  // nI.Available(): deep copy
    binding->clear();
    for (auto& pres : p.extended) {
      std::string rname, err;
      if (!pod_extended_resource_name(pres, &rname, &err)) {
        *fail = err;
        return false;
      }
      std::vector<std::string> res;
      int64_t val = pres.limits.begin()->second;
      if (!allocate_resources(rname, val, pres.required, available, &res, fail)) {
        binding->clear();
        return false;
      }
      (*binding)[pres.name] = res;
      auto dom = available.find(rname);
      if (dom != available.end())
        for (auto& id : res) dom->second.erase(id);
    }
    return true;
  }

  // ---------------- priorities

  // least_requested.go:50-61
  static int64_t unused_score(int64_t requested, int64_t capacity) {
    if (capacity == 0) return 0;
    if (requested > capacity) return 0;
    return ((capacity - requested) * kMaxPriority) / capacity;
  }
  // balanced_resource_allocation.go:95-100
  static double fraction_of_capacity(int64_t requested, int64_t capacity) {
    if (capacity == 0) return 1;
    return (double)requested / (double)capacity;
  }

  struct PriorityMeta {  // priorities/metadata.go:45-66
    int64_t nonZeroCPU = 0, nonZeroMem = 0;
    std::vector<ksb::Toleration> tolerationsPreferNoSchedule;
    std::vector<Selector> podSelectors;
  };
  PriorityMeta make_priority_meta(const Pod& pod) const {
    PriorityMeta m;
    for (auto& c : pod.containers) {
      int64_t a, b;
      get_nonzero_requests(c.requests, &a, &b);
      m.nonZeroCPU += a;
      m.nonZeroMem += b;
    }
    for (auto& t : pod.tolerations)
      if (t.effect.empty() || t.effect == "PreferNoSchedule") m.tolerationsPreferNoSchedule.push_back(t);
    m.podSelectors = get_selectors(pod, listers);
    return m;
  }

  static int64_t least_requested_map(const PriorityMeta& pm, const NodeInfo& ni) {  // :63-91
    int64_t cpu = pm.nonZeroCPU + ni.nonzero.MilliCPU;
    int64_t mem = pm.nonZeroMem + ni.nonzero.Memory;
    int64_t cs = unused_score(cpu, ni.allocatable.MilliCPU);
    int64_t ms = unused_score(mem, ni.allocatable.Memory);
    return (cs + ms) / 2;
  }
  // most_requested.go:32-88: calculateUsedScore per resource, integer average
  static int64_t used_score(int64_t requested, int64_t capacity) {
    if (capacity == 0) return 0;
    if (requested > capacity) return 0;
    return (requested * kMaxPriority) / capacity;
  }
  static int64_t most_requested_map(const PriorityMeta& pm, const NodeInfo& ni) {
    int64_t cpu = pm.nonZeroCPU + ni.nonzero.MilliCPU;
    int64_t mem = pm.nonZeroMem + ni.nonzero.Memory;
    return (used_score(cpu, ni.allocatable.MilliCPU) + used_score(mem, ni.allocatable.Memory)) / 2;
  }
  // node_label.go:46-60
  static int64_t node_label_map(const NodeLabelPriorityArg& a, const NodeInfo& ni) {
    const bool exists = ni.node->has_labels && ni.node->labels.count(a.label) != 0;
    return (exists && a.presence) || (!exists && !a.presence) ? kMaxPriority : 0;
  }
  // resource_limits.go:34-126: 1 when the pod's cpu or memory limit is set and fits the node's allocatable
  static void pod_limits(const Pod& pod, int64_t* cpu, int64_t* mem) {
    *cpu = 0;
    *mem = 0;
    for (auto& c : pod.containers)
      for (auto& kv : c.limits) {
        if (kv.first == ksb::kResCPU) *cpu += kv.second;
        else if (kv.first == ksb::kResMemory) *mem += kv.second;
      }
    for (auto& c : pod.init_containers)
      for (auto& kv : c.limits) {
        if (kv.first == ksb::kResCPU) *cpu = std::max(*cpu, kv.second);
        else if (kv.first == ksb::kResMemory) *mem = std::max(*mem, kv.second);
      }
  }
  // image_locality.go:32-82: the sizes of the pod's container images already on the node, bucketed into 0..10
  static int64_t image_locality_map(const Pod& pod, const NodeInfo& ni) {
    constexpr int64_t mb = 1024 * 1024, minImgSize = 23 * mb, maxImgSize = 1000 * mb;  // balanced_resource_allocation.go:32-36
    std::map<std::string, int64_t> imageSizes;  // totalImageSize :66-82 (a later entry of the same name wins)
    for (auto& im : ni.node->images)
      for (auto& name : im.names) imageSizes[name] = im.size_bytes;
    int64_t total = 0;
    for (auto& image : pod.images) {
      auto it = imageSizes.find(image);
      if (it != imageSizes.end()) total += it->second;
    }
    if (total == 0 || total < minImgSize) return 0;  // calculateScoreFromSize :49-62
    if (total >= maxImgSize) return kMaxPriority;
    return kMaxPriority * (total - minImgSize) / (maxImgSize - minImgSize) + 1;
  }
  static int64_t resource_limits_map(const Pod& pod, const NodeInfo& ni) {
    int64_t lc, lm;
    pod_limits(pod, &lc, &lm);
    auto one = [](int64_t limit, int64_t alloc) { return limit != 0 && alloc != 0 && limit <= alloc; };
    return one(lc, ni.allocatable.MilliCPU) || one(lm, ni.allocatable.Memory) ? 1 : 0;
  }
  // predicates.go:800-826
  static void check_node_label_presence(const Policy& pol, const NodeInfo& ni, std::vector<std::string>* fails) {
    for (auto& l : pol.presence_labels) {
      const bool exists = ni.node->has_labels && ni.node->labels.count(l) != 0;
      if ((exists && !pol.presence) || (!exists && pol.presence)) {
        fails->push_back("CheckNodeLabelPresence");
        return;
      }
    }
  }

  static int64_t balanced_map(const PriorityMeta& pm, const NodeInfo& ni) {  // :51-93
    int64_t cpu = pm.nonZeroCPU + ni.nonzero.MilliCPU;
    int64_t mem = pm.nonZeroMem + ni.nonzero.Memory;
    double cf = fraction_of_capacity(cpu, ni.allocatable.MilliCPU);
    double mf = fraction_of_capacity(mem, ni.allocatable.Memory);
    if (cf >= 1 || mf >= 1) return 0;
    double diff = std::fabs(cf - mf);
    return (int64_t)((1 - diff) * (double)kMaxPriority);
  }
  static int64_t selector_spread_map(const Pod& pod, const PriorityMeta& pm, const NodeInfo& ni) {  // :64-113
    if (pm.podSelectors.empty()) return 0;
    int64_t count = 0;
    for (auto& np : ni.pods) {
      if (pod.ns != np->ns) continue;
      if (np->deleting) continue;
      bool matches = false;
      for (auto& s : pm.podSelectors)
        if (sel_matches(s, np->labels)) {
          matches = true;
          break;
        }
      if (matches) count++;
    }
    return count;
  }
  void selector_spread_reduce(HostPriorityList& result) const {  // :119-174
    std::map<std::string, int64_t> countsByZone;
    int64_t maxCountByZone = 0, maxCountByNodeName = 0;
    for (auto& r : result) {
      if (r.score > maxCountByNodeName) maxCountByNodeName = r.score;
      std::string z = get_zone_key(*cachedNodeInfoMap.at(r.host).node);
      if (z.empty()) continue;
      countsByZone[z] += r.score;
    }
    for (auto& kv : countsByZone)
      if (kv.second > maxCountByZone) maxCountByZone = kv.second;
    bool haveZones = !countsByZone.empty();
    double maxNodeF = (double)maxCountByNodeName;
    double maxZoneF = (double)maxCountByZone;
    double maxPrioF = (double)kMaxPriority;
    const double zoneWeighting = 2.0 / 3.0;
    for (auto& r : result) {
      double fScore = maxPrioF;
      if (maxCountByNodeName > 0) fScore = maxPrioF * ((double)(maxCountByNodeName - r.score) / maxNodeF);
      if (haveZones) {
        std::string z = get_zone_key(*cachedNodeInfoMap.at(r.host).node);
        if (!z.empty()) {
          double zoneScore = maxPrioF;
          if (maxCountByZone > 0) zoneScore = maxPrioF * ((double)(maxCountByZone - countsByZone[z]) / maxZoneF);
          fScore = (fScore * (1.0 - zoneWeighting)) + (zoneWeighting * zoneScore);
        }
      }
      r.score = (int64_t)fScore;
    }
  }
  // node_affinity.go:34-75; returns false + err on selector parse error (aborts Schedule)
  static bool node_affinity_map(const Pod& pod, const NodeInfo& ni, int64_t* out, std::string* err) {
    int32_t count = 0;
    if (pod.affinity.present && pod.affinity.node.present) {
      for (auto& pt : pod.affinity.node.preferred) {
        if (pt.weight == 0) continue;
        Selector sel;
        if (!requirements_as_selector(pt.preference.match_expressions, &sel, err)) return false;
        if (sel_matches(sel, ni.node->labels)) count += pt.weight;
      }
    }
    *out = count;
    return true;
  }
  static int64_t taint_toleration_map(const PriorityMeta& pm, const NodeInfo& ni) {  // :29-73
    int64_t n = 0;
    for (auto& t : ni.node->taints) {
      if (t.effect != "PreferNoSchedule") continue;
      if (!tolerations_tolerate_taint(pm.tolerationsPreferNoSchedule, t)) n++;
    }
    return n;
  }
  static void normalize_reduce(HostPriorityList& result, bool reverse) {  // reduce.go:29-63
    int64_t maxCount = 0;
    for (auto& r : result)
      if (r.score > maxCount) maxCount = r.score;
    if (maxCount == 0) {
      if (reverse)
        for (auto& r : result) r.score = kMaxPriority;
      return;
    }
    for (auto& r : result) {
      int64_t s = kMaxPriority * r.score / maxCount;
      if (reverse) s = kMaxPriority - s;
      r.score = s;
    }
  }
  static int64_t prefer_avoid_map(const Pod& pod, const NodeInfo& ni) {  // node_prefer_avoid_pods.go:29-59
    bool has = pod.controller.present &&
               (pod.controller.kind == "ReplicationController" || pod.controller.kind == "ReplicaSet");
    if (!has) return kMaxPriority;
    if (ni.node->avoid_annotation_invalid) return kMaxPriority;
    for (auto& a : ni.node->avoid)
      if (a.kind == pod.controller.kind && a.uid == pod.controller.uid) return 0;
    return kMaxPriority;
  }

  // interpod_affinity.go:121-245
  bool inter_pod_affinity_priority(const Pod& pod, const std::vector<NodePtr>& nodes, HostPriorityList* out,
                                   std::string* err) const {
    bool hasAff = pod.affinity.present && pod.affinity.pod.present;
    bool hasAnti = pod.affinity.present && pod.affinity.anti.present;
    std::map<std::string, double> counts;
    std::map<std::string, std::map<std::string, double>> domains;  // indexed mode: weight per (topology key, value)
    bool failed = false;
    auto process_term = [&](const ksb::PodAffinityTerm& term, const Pod& defining, const Pod& toCheck,
                            const Node& fixedNode, double weight) {
      auto nss = namespaces_from_term(defining, term);
      Selector sel;
      std::string e;
      if (!label_selector_as_selector(term.selector, &sel, &e)) {
        if (!failed) *err = e;
        failed = true;
        return;
      }
      if (pod_matches_ns_and_selector(toCheck, nss, sel)) {
        if (indexed) {  // one entry per (topology key, value) instead of one pass over the nodes per matching term
          if (const std::string* v = topo_value(fixedNode, term.topology_key)) domains[term.topology_key][*v] += weight;
        } else {
          for (auto& n : nodes)
            if (nodes_have_same_topology_key(*n, fixedNode, term.topology_key)) counts[n->name] += weight;
        }
      }
    };
    auto process_terms = [&](const std::vector<ksb::WeightedPodAffinityTerm>& terms, const Pod& defining,
                             const Pod& toCheck, const Node& fixedNode, int mult) {
      for (auto& t : terms) process_term(t.term, defining, toCheck, fixedNode, (double)(t.weight * mult));
    };
    auto process_pod = [&](const Pod& ep) {
      NodePtr en = get_node(ep.node_name);
      if (!en) return;  // NotFound is logged and skipped (:164-170)
      bool eAff = ep.affinity.present && ep.affinity.pod.present;
      bool eAnti = ep.affinity.present && ep.affinity.anti.present;
      if (hasAff) process_terms(pod.affinity.pod.preferred, pod, ep, *en, 1);
      if (hasAnti) process_terms(pod.affinity.anti.preferred, pod, ep, *en, -1);
      if (eAff) {
        if (hardPodAffinityWeight > 0)
          for (auto& t : ep.affinity.pod.required) process_term(t, ep, pod, *en, (double)hardPodAffinityWeight);
        process_terms(ep.affinity.pod.preferred, ep, pod, *en, 1);
      }
      if (eAnti) process_terms(ep.affinity.anti.preferred, ep, pod, *en, -1);
    };
    for (auto& kv : cachedNodeInfoMap) {
      const NodeInfo& ni = kv.second;
      if (!ni.node) continue;
      if (hasAff || hasAnti) {
        for (auto& ep : ni.pods) process_pod(*ep);
      } else {
        for (auto& ep : ni.podsWithAffinity) process_pod(*ep);
      }
    }
    if (failed) return false;
    if (indexed)  // the weights are integers: their sum is exact in any order
      for (auto& n : nodes)
        for (auto& kv : domains)
          if (const std::string* v = topo_value(*n, kv.first)) {
            auto hit = kv.second.find(*v);
            if (hit != kv.second.end()) counts[n->name] += hit->second;
          }
    double maxCount = 0, minCount = 0;
    for (auto& n : nodes) {
      double c = counts.count(n->name) ? counts[n->name] : 0.0;
      if (c > maxCount) maxCount = c;
      if (c < minCount) minCount = c;
    }
    out->clear();
    for (auto& n : nodes) {
      double fScore = 0;
      double c = counts.count(n->name) ? counts[n->name] : 0.0;
      if ((maxCount - minCount) > 0) fScore = (double)kMaxPriority * ((c - minCount) / (maxCount - minCount));
      out->push_back({n->name, (int64_t)fScore});
    }
    return true;
  }

  // ---------------- core

  // FitError.Error(), generic_scheduler.go:70-88
  static std::string fit_error_message(int numAllNodes, const std::map<std::string, std::vector<std::string>>& failed) {
    std::map<std::string, int> reasons;
    for (auto& kv : failed)
      for (auto& r : kv.second) reasons[r]++;
    std::vector<std::string> strs;
    for (auto& kv : reasons) strs.push_back(std::to_string(kv.second) + " " + kv.first);
    std::sort(strs.begin(), strs.end());
    std::string msg = "0/" + std::to_string(numAllNodes) + " nodes are available: ";
    for (size_t i = 0; i < strs.size(); ++i) {
      if (i) msg += ", ";
      msg += strs[i];
    }
    msg += ".";
    return msg;
  }

  // Evaluate the DefaultProvider predicate set on one node (podFitsOnNode, single pass).
  // Returns false on a hard error.
  bool pod_fits_on_node(const Pod& pod, const PredicateMeta& meta, const NodeInfo& ni, const std::vector<PodPtr>& allPods,
                        const std::vector<TermProbe>& affTerms, const std::vector<TermProbe>& antiTerms,
                        std::vector<std::string>* fails, std::string* err, const IpaIndex* ix = nullptr) const {
    if (!ni.node) {
      *err = "node not found";
      return false;
    }
    // GeneralPredicates :965-1029
    pod_fits_resources(meta, ni, fails);
    if (!pod.node_name.empty() && pod.node_name != ni.node->name) fails->push_back("HostName");
    if (!meta.podPorts.empty() && ports_conflict(ni.usedPorts, meta.podPorts)) fails->push_back("PodFitsHostPorts");
    if (!pod_matches_node_labels(pod, *ni.node)) fails->push_back("MatchNodeSelector");
    // PodToleratesNodeTaints :1349-1373
    {
      bool ok = true;
      for (auto& t : ni.taints) {
        if (t.effect != "NoSchedule" && t.effect != "NoExecute") continue;
        if (!tolerations_tolerate_taint(pod.tolerations, t)) {
          ok = false;
          break;
        }
      }
      if (!ok) fails->push_back("PodToleratesNodeTaints");
    }
    if (meta.bestEffort && ni.memoryPressure == "True") fails->push_back("NodeUnderMemoryPressure");
    if (ni.diskPressure == "True") fails->push_back("NodeUnderDiskPressure");
    check_node_condition(ni, fails);
    if (!policy.presence_labels.empty()) check_node_label_presence(policy, ni, fails);
    if (ix && ix->usable) inter_pod_affinity_matches_indexed(pod, ni, *ix, affTerms, antiTerms, fails);
    else if (!inter_pod_affinity_matches(pod, meta, ni, allPods, affTerms, antiTerms, fails, err)) return false;
    return true;
  }

  // genericScheduler.Schedule, generic_scheduler.go:109-161
  ScheduleResult schedule(const Pod& pod, const std::vector<NodePtr>& nodes) {
    ScheduleResult res;
    if (nodes.empty()) {
      res.status = ScheduleResult::NoNodes;
      res.message = "no nodes available to schedule pods";
      return res;
    }
    if (pod.num_volumes > 0) {  // volume predicates are out of scope: reject loudly
      res.status = ScheduleResult::Error;
      res.message = "pods with volumes are not supported";
      return res;
    }
    // UpdateNodeNameToInfoMap (cache.go:83-97): clone only NodeInfos whose generation moved
    for (auto& kv : cache.nodes) {
      auto cur = cachedNodeInfoMap.find(kv.first);
      if (cur == cachedNodeInfoMap.end() || cur->second.generation != kv.second.generation)
        cachedNodeInfoMap[kv.first] = kv.second;
    }
    for (auto it = cachedNodeInfoMap.begin(); it != cachedNodeInfoMap.end();) {
      if (!cache.nodes.count(it->first)) it = cachedNodeInfoMap.erase(it);
      else ++it;
    }

    // findNodesThatFit :283-362
    PredicateMeta meta = make_predicate_meta(pod);
    if (meta.error) {
      // GetMetadata returns nil; every predicate then recomputes its inputs from the pod.
      // The only observable difference is for inter-pod anti-affinity, which the harness
      // never drives through this branch (all cached nodes have Node objects).
      res.status = ScheduleResult::Error;
      res.message = meta.error_msg;
      return res;
    }
    std::vector<PodPtr> allPods;
    std::vector<TermProbe> affTerms, antiTerms;
    if (pod.affinity.present && (pod.affinity.pod.present || pod.affinity.anti.present)) {
      allPods = all_pods();
      for (auto& t : required_terms(pod.affinity.pod)) affTerms.push_back(probe(pod, t));
      for (auto& t : required_terms(pod.affinity.anti)) antiTerms.push_back(probe(pod, t));
    }
    IpaIndex ipaIndex;
    if (indexed) ipaIndex = build_ipa_index(meta, allPods, affTerms, antiTerms);
    int N = (int)nodes.size();
    std::vector<char> fits(N, 0);
    std::vector<std::vector<std::string>> fails(N);
    std::vector<std::string> errs(N);
    std::vector<char> has_err(N, 0);
    parallelize(workers, N, [&](int i) {
      const NodeInfo& ni = cachedNodeInfoMap.at(nodes[i]->name);
      std::string err;
      if (!pod_fits_on_node(pod, meta, ni, allPods, affTerms, antiTerms, &fails[i], &err, indexed ? &ipaIndex : nullptr)) {
        has_err[i] = 1;
        errs[i] = err;
        return;
      }
      fits[i] = fails[i].empty();
    });
    for (int i = 0; i < N; ++i)
      if (has_err[i]) {
        res.status = ScheduleResult::Error;
        res.message = errs[i];
        return res;
      }
    std::map<std::string, std::vector<std::string>> failedPredicateMap;
    std::vector<NodePtr> filtered;
    for (int i = 0; i < N; ++i) {
      if (fits[i]) filtered.push_back(nodes[i]);
      else failedPredicateMap[nodes[i]->name] = fails[i];
    }
    // extenders' Filter :333-351, before the fork's device selection
    if (!filtered.empty() && extender.has_filter) {
      std::vector<NodePtr> kept;
      for (auto& nd : filtered) {
        if (extender.keep.count(nd->name)) kept.push_back(nd);
        else failedPredicateMap[nd->name].push_back("extender");
      }
      filtered.swap(kept);
    }
    // GetExtendedResources (fork) :357 -> extended_resources.go:42-81
    std::map<std::string, std::map<std::string, std::vector<std::string>>> extRes;
    if (!filtered.empty() && !pod.extended.empty()) {
      int F = (int)filtered.size();
      std::vector<char> ok(F, 0);
      std::vector<std::map<std::string, std::vector<std::string>>> bind(F);
      std::vector<std::string> why(F);
      parallelize(workers, F, [&](int i) {
        const NodeInfo& ni = cachedNodeInfoMap.at(filtered[i]->name);
        ok[i] = has_extended_resources(pod, ni, &bind[i], &why[i]);
      });
      std::vector<NodePtr> refiltered;
      for (int i = 0; i < F; ++i) {
        if (ok[i]) {
          refiltered.push_back(filtered[i]);
          extRes[filtered[i]->name] = bind[i];
        } else {
          failedPredicateMap[filtered[i]->name] = {why[i]};
        }
      }
      filtered.swap(refiltered);
    }
    res.num_filtered = (int)filtered.size();
    if (filtered.empty()) {
      res.status = ScheduleResult::FitError;
      res.message = fit_error_message(N, failedPredicateMap);
      return res;
    }
    if (filtered.size() == 1) {  // :144-146: no selectHost, lastNodeIndex untouched
      res.host = filtered[0]->name;
      res.binding = extRes[res.host];
      return res;
    }
    // PrioritizeNodes :509-640
    HostPriorityList total;
    std::string err;
    if (!prioritize(pod, filtered, &total, &err)) {
      res.status = ScheduleResult::Error;
      res.message = err;
      return res;
    }
    // selectHost :177-192 with HostPriorityList.Less (api/types.go:218-223):
    // descending (score, host); pick lastNodeIndex % (#max-score entries).
    std::sort(total.begin(), total.end(), [](const HostPriority& a, const HostPriority& b) {
      if (a.score != b.score) return a.score > b.score;
      return a.host > b.host;
    });
    int64_t maxScore = total[0].score;
    size_t firstAfter = 0;
    while (firstAfter < total.size() && total[firstAfter].score == maxScore) ++firstAfter;
    size_t ix = (size_t)(lastNodeIndex % (uint64_t)firstAfter);
    lastNodeIndex++;
    res.host = total[ix].host;
    res.score = maxScore;
    res.binding = extRes[res.host];
    return res;
  }

  bool prioritize(const Pod& pod, const std::vector<NodePtr>& filtered, HostPriorityList* total, std::string* err) const {
    int F = (int)filtered.size();
    PriorityMeta pm = make_priority_meta(pod);
    HostPriorityList ipa;
    if (!inter_pod_affinity_priority(pod, filtered, &ipa, err)) return false;
    HostPriorityList lr(F), bra(F), ss(F), na(F), tt(F), av(F);
    std::vector<char> bad(F, 0);
    std::vector<std::string> errs(F);
    parallelize(workers, F, [&](int i) {
      const NodeInfo& ni = cachedNodeInfoMap.at(filtered[i]->name);
      const std::string& h = filtered[i]->name;
      ss[i] = {h, selector_spread_map(pod, pm, ni)};
      // DefaultProvider order (defaults.go:217-257): SelectorSpread, InterPodAffinity, LeastRequested,
      // Balanced, NodePreferAvoidPods, NodeAffinity, TaintToleration
      lr[i] = {h, least_requested_map(pm, ni)};
      bra[i] = {h, balanced_map(pm, ni)};
      av[i] = {h, prefer_avoid_map(pod, ni)};
      int64_t c;
      if (!node_affinity_map(pod, ni, &c, &errs[i])) {
        bad[i] = 1;
        return;
      }
      na[i] = {h, c};
      tt[i] = {h, taint_toleration_map(pm, ni)};
    });
    for (int i = 0; i < F; ++i)
      if (bad[i]) {
        *err = errs[i];
        return false;
      }
    selector_spread_reduce(ss);
    normalize_reduce(na, false);
    normalize_reduce(tt, true);
    total->assign(F, HostPriority());
    for (int i = 0; i < F; ++i) {
      (*total)[i].host = filtered[i]->name;
      (*total)[i].score = ss[i].score * weights.selector_spread + ipa[i].score * weights.inter_pod_affinity +
                          lr[i].score * weights.least_requested + bra[i].score * weights.balanced +
                          av[i].score * weights.prefer_avoid + na[i].score * weights.node_affinity +
                          tt[i].score * weights.taint_toleration;
      // Policy-enabled priorities (Map only, no Reduce)
      const NodeInfo& ni = cachedNodeInfoMap.at(filtered[i]->name);
      if (weights.most_requested) (*total)[i].score += most_requested_map(pm, ni) * weights.most_requested;
      if (weights.resource_limits) (*total)[i].score += resource_limits_map(pod, ni) * weights.resource_limits;
      if (weights.image_locality) (*total)[i].score += image_locality_map(pod, ni) * weights.image_locality;
      (*total)[i].score += ss[i].score * weights.service_spreading + 1 * weights.equal;
      for (auto& a : policy.node_label) (*total)[i].score += node_label_map(a, ni) * a.weight;
      // extenders' Prioritize :608-632
      auto es = extender.scores.find(filtered[i]->name);
      if (es != extender.scores.end()) (*total)[i].score += es->second;
    }
    return true;
  }

  // ---------------- harness: the scheduleOne loop (scheduler.go:430-497) with all binds
  // succeeding and the apiserver's bind echo of device IDs emulated.

  std::vector<NodePtr> node_list() const {
    std::vector<NodePtr> out;
    for (auto& kv : cache.nodes)
      if (kv.second.node) out.push_back(kv.second.node);
    return out;
  }

  void load(const ksb::Cluster& c) {
    listers.services = c.services;
    listers.rcs = c.rcs;
    listers.rss = c.rss;
    listers.sss = c.sss;
    for (auto& n : c.nodes) cache.AddNode(n);
    for (auto& p : c.existing_pods) cache.AddPod(p);
  }

  ScheduleResult schedule_one(const Pod& pod_in) {
    std::vector<NodePtr> nodes = node_list();
    ScheduleResult r = schedule(pod_in, nodes);
    if (r.status != ScheduleResult::OK) return r;
    Pod assumed = pod_in;  // scheduler.go:478: assumedPod.Spec.NodeName = host
    assumed.node_name = r.host;
    for (auto& e : assumed.extended) {  // bind echo: storage.go:180-187 sets Assigned
      auto it = r.binding.find(e.name);
      if (it != r.binding.end()) e.assigned = it->second;
    }
    cache.AssumePod(assumed);
    return r;
  }
};

}  // namespace oracle
