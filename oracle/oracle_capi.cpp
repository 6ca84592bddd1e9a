// ORACLE — TEST INFRASTRUCTURE ONLY. C entry points (JSON in, JSON out) so the
// Python tests and bench.py's cpu_baseline leg can drive the CPU restatement.
// Nothing under kubernetes_b200/ may include or link this.
#include <algorithm>
#include <chrono>
#include <cstdlib>
#include <cstring>
#include <string>

#include "json.hpp"
#include "model.hpp"
#include "scheduler.hpp"
#include "preempt.hpp"

using namespace oracle;
namespace json = ksb::json;

static void put_kv(std::string& out, const char* k, const std::string& v, bool comma = true) {
  json::escape_to(out, k);
  out += ":";
  json::escape_to(out, v);
  if (comma) out += ",";
}

static void result_to_json(std::string& out, const ScheduleResult& r) {
  out += "{";
  put_kv(out, "host", r.host);
  out += "\"status\":" + std::to_string((int)r.status) + ",";
  out += "\"filtered\":" + std::to_string(r.num_filtered) + ",";
  out += "\"score\":" + std::to_string(r.score) + ",";
  put_kv(out, "message", r.message);
  out += "\"devices\":{";
  bool first = true;
  for (auto& kv : r.binding) {
    if (!first) out += ",";
    first = false;
    json::escape_to(out, kv.first);
    out += ":[";
    for (size_t i = 0; i < kv.second.size(); ++i) {
      if (i) out += ",";
      json::escape_to(out, kv.second[i]);
    }
    out += "]";
  }
  out += "}}";
}

static void apply_options(Oracle& o, const json::Value& v) {
  o.workers = (int)v.get_int("workers", 1);
  o.indexed = v.get_bool("indexed", false);
  o.hardPodAffinityWeight = (int)v.get_int("hardPodAffinityWeight", 1);
  o.lastNodeIndex = (uint64_t)v.get_int("lastNodeIndex", 0);
  if (const json::Value* w = v.find("weights"); w && !w->is_null()) {
    o.weights.least_requested = w->get_int("LeastRequestedPriority", 1);
    o.weights.balanced = w->get_int("BalancedResourceAllocation", 1);
    o.weights.selector_spread = w->get_int("SelectorSpreadPriority", 1);
    o.weights.node_affinity = w->get_int("NodeAffinityPriority", 1);
    o.weights.taint_toleration = w->get_int("TaintTolerationPriority", 1);
    o.weights.inter_pod_affinity = w->get_int("InterPodAffinityPriority", 1);
    o.weights.prefer_avoid = w->get_int("NodePreferAvoidPodsPriority", 10000);
    o.weights.most_requested = w->get_int("MostRequestedPriority", 0);
    o.weights.resource_limits = w->get_int("ResourceLimitsPriority", 0);
    o.weights.image_locality = w->get_int("ImageLocalityPriority", 0);
    o.weights.service_spreading = w->get_int("ServiceSpreadingPriority", 0);
    o.weights.equal = w->get_int("EqualPriority", 0);
  }
  // {"extender": {"keep": [node names] (optional), "scores": {node: weighted score}}}
  if (const json::Value* ev = v.find("extender"); ev && !ev->is_null()) {
    if (const json::Value* k = ev->find("keep"); k && !k->is_null()) {
      o.extender.has_filter = true;
      for (auto& x : *k->a) o.extender.keep.insert(x.s);
    }
    if (const json::Value* sc = ev->find("scores"); sc && sc->is_obj())
      for (auto& kv : *sc->o) o.extender.scores[kv.first] = kv.second.i;
  }
  // {"policy": {"nodeLabelPriorities": [{"label", "presence", "weight"}], "labelsPresence": {"labels": [...], "presence": bool}}}
  if (const json::Value* pv = v.find("policy"); pv && !pv->is_null()) {
    if (const json::Value* nl = pv->find("nodeLabelPriorities"); nl && !nl->is_null())
      for (auto& a : pv->get_arr("nodeLabelPriorities")) {
        NodeLabelPriorityArg x;
        x.label = a.get_str("label");
        x.presence = a.get_bool("presence", true);
        x.weight = a.get_int("weight", 1);
        o.policy.node_label.push_back(x);
      }
    if (const json::Value* lp = pv->find("labelsPresence"); lp && !lp->is_null()) {
      for (auto& l : lp->get_arr("labels")) o.policy.presence_labels.push_back(l.s);
      o.policy.presence = lp->get_bool("presence", true);
    }
  }
}

static std::string run_stream(const json::Value& req) {
  json::Value cv;
  const json::Value* c = req.find("cluster");
  if (!c) {
    cv = json::parse_file(req.get_str("cluster_path"));
    c = &cv;
  }
  ksb::Cluster cl = ksb::cluster_from_json(*c);
  Oracle o;
  apply_options(o, req);
  o.load(cl);
  int64_t limit = req.get_int("limit", -1);
  size_t n = cl.pods.size();
  if (limit >= 0 && (size_t)limit < n) n = (size_t)limit;
  std::string out = "{\"results\":[";
  auto t0 = std::chrono::steady_clock::now();
  std::vector<ScheduleResult> results;
  results.reserve(n);
  for (size_t i = 0; i < n; ++i) results.push_back(o.schedule_one(cl.pods[i]));
  auto t1 = std::chrono::steady_clock::now();
  for (size_t i = 0; i < n; ++i) {
    if (i) out += ",";
    result_to_json(out, results[i]);
  }
  double secs = std::chrono::duration<double>(t1 - t0).count();
  out += "],\"seconds\":" + std::to_string(secs) + ",\"pods\":" + std::to_string(n) +
         ",\"lastNodeIndex\":" + std::to_string(o.lastNodeIndex) + "}";
  return out;
}

// priorityFunction(map, reduce, meta) adapter of priorities/test_util.go:45-61: Map on every
// listed node, then Reduce over all of them (no predicate filtering).
static std::string run_priority(const json::Value& req) {
  ksb::Cluster cl = ksb::cluster_from_json(*req.find("cluster"));
  ksb::Pod pod = ksb::pod_from_json(*req.find("pod"));
  Oracle o;
  apply_options(o, req);
  o.load(cl);
  o.cachedNodeInfoMap = o.cache.nodes;
  std::vector<NodePtr> nodes;
  for (auto& n : cl.nodes) nodes.push_back(o.cachedNodeInfoMap.at(n.name).node);
  std::string name = req.get_str("name");
  HostPriorityList list;
  Oracle::PriorityMeta pm = o.make_priority_meta(pod);
  std::string err;
  if (name == "InterPodAffinityPriority") {
    if (!o.inter_pod_affinity_priority(pod, nodes, &list, &err)) return "{\"error\":\"" + err + "\"}";
  } else {
    for (auto& n : nodes) {
      const NodeInfo& ni = o.cachedNodeInfoMap.at(n->name);
      int64_t s = 0;
      if (name == "LeastRequestedPriority") s = Oracle::least_requested_map(pm, ni);
      else if (name == "BalancedResourceAllocation") s = Oracle::balanced_map(pm, ni);
      else if (name == "SelectorSpreadPriority") s = Oracle::selector_spread_map(pod, pm, ni);
      else if (name == "NodeAffinityPriority") {
        if (!Oracle::node_affinity_map(pod, ni, &s, &err)) return "{\"error\":\"" + err + "\"}";
      } else if (name == "TaintTolerationPriority") s = Oracle::taint_toleration_map(pm, ni);
      else if (name == "NodePreferAvoidPodsPriority") s = Oracle::prefer_avoid_map(pod, ni);
      else if (name == "MostRequestedPriority") s = Oracle::most_requested_map(pm, ni);
      else if (name == "ResourceLimitsPriority") s = Oracle::resource_limits_map(pod, ni);
      else if (name == "ImageLocalityPriority") s = Oracle::image_locality_map(pod, ni);
      else if (name == "NodeLabelPriority") {
        NodeLabelPriorityArg a;
        a.label = req.find("args")->get_str("label");
        a.presence = req.find("args")->get_bool("presence", true);
        s = Oracle::node_label_map(a, ni);
      }
      else return "{\"error\":\"unknown priority\"}";
      list.push_back({n->name, s});
    }
    if (name == "SelectorSpreadPriority") o.selector_spread_reduce(list);
    else if (name == "NodeAffinityPriority") Oracle::normalize_reduce(list, false);
    else if (name == "TaintTolerationPriority") Oracle::normalize_reduce(list, true);
  }
  std::string out = "{\"scores\":[";
  for (size_t i = 0; i < list.size(); ++i) {
    if (i) out += ",";
    out += "{";
    put_kv(out, "host", list[i].host);
    out += "\"score\":" + std::to_string(list[i].score) + "}";
  }
  out += "]}";
  return out;
}

// One predicate (or the whole DefaultProvider set, name "all") on one node.
static std::string run_predicate(const json::Value& req) {
  ksb::Cluster cl = ksb::cluster_from_json(*req.find("cluster"));
  ksb::Pod pod = ksb::pod_from_json(*req.find("pod"));
  Oracle o;
  apply_options(o, req);
  o.load(cl);
  o.cachedNodeInfoMap = o.cache.nodes;
  std::string node_name = req.get_str("node");
  const NodeInfo& ni = o.cachedNodeInfoMap.at(node_name);
  std::string name = req.get_str("name");
  PredicateMeta meta = o.make_predicate_meta(pod);
  std::vector<std::string> fails;
  std::string err;
  bool hard_error = false;
  if (name == "PodFitsResources") {
    Oracle::pod_fits_resources(meta, ni, &fails);
  } else if (name == "GeneralPredicates") {
    // predicates.go:965-1035: noncriticalPredicates (PodFitsResources), then EssentialPredicates (PodFitsHost,
    // PodFitsHostPorts, PodMatchNodeSelector); every failing one adds its reasons
    Oracle::pod_fits_resources(meta, ni, &fails);
    if (!pod.node_name.empty() && pod.node_name != ni.node->name) fails.push_back("HostName");
    if (!meta.podPorts.empty() && ports_conflict(ni.usedPorts, meta.podPorts)) fails.push_back("PodFitsHostPorts");
    if (!Oracle::pod_matches_node_labels(pod, *ni.node)) fails.push_back("MatchNodeSelector");
  } else if (name == "PodFitsHost") {
    if (!pod.node_name.empty() && pod.node_name != ni.node->name) fails.push_back("HostName");
  } else if (name == "PodFitsHostPorts") {
    if (!meta.podPorts.empty() && ports_conflict(ni.usedPorts, meta.podPorts)) fails.push_back("PodFitsHostPorts");
  } else if (name == "PodMatchNodeSelector") {
    if (!Oracle::pod_matches_node_labels(pod, *ni.node)) fails.push_back("MatchNodeSelector");
  } else if (name == "PodToleratesNodeTaints") {
    for (auto& t : ni.taints) {
      if (t.effect != "NoSchedule" && t.effect != "NoExecute") continue;
      if (!tolerations_tolerate_taint(pod.tolerations, t)) {
        fails.push_back("PodToleratesNodeTaints");
        break;
      }
    }
  } else if (name == "CheckNodeMemoryPressure") {
    if (meta.bestEffort && ni.memoryPressure == "True") fails.push_back("NodeUnderMemoryPressure");
  } else if (name == "CheckNodeDiskPressure") {
    if (ni.diskPressure == "True") fails.push_back("NodeUnderDiskPressure");
  } else if (name == "CheckNodeCondition") {
    Oracle::check_node_condition(ni, &fails);
  } else if (name == "CheckNodeLabelPresence") {
    Oracle::check_node_label_presence(o.policy, ni, &fails);
  } else if (name == "MatchInterPodAffinity" || name == "all") {
    std::vector<PodPtr> allPods = o.all_pods();
    std::vector<Oracle::TermProbe> aff, anti;
    if (pod.affinity.present) {
      for (auto& t : required_terms(pod.affinity.pod)) aff.push_back(Oracle::probe(pod, t));
      for (auto& t : required_terms(pod.affinity.anti)) anti.push_back(Oracle::probe(pod, t));
    }
    if (name == "all") {
      if (!o.pod_fits_on_node(pod, meta, ni, allPods, aff, anti, &fails, &err)) hard_error = true;
    } else {
      if (!o.inter_pod_affinity_matches(pod, meta, ni, allPods, aff, anti, &fails, &err)) hard_error = true;
    }
  } else if (name == "HasExtendedResources") {
    std::map<std::string, std::vector<std::string>> binding;
    std::string why;
    bool ok = Oracle::has_extended_resources(pod, ni, &binding, &why);
    if (!ok) fails.push_back(why);
    std::string out = std::string("{\"fits\":") + (ok ? "true" : "false") + ",\"reasons\":[";
    if (!ok) json::escape_to(out, why);
    out += "],\"devices\":{";
    bool first = true;
    for (auto& kv : binding) {
      if (!first) out += ",";
      first = false;
      json::escape_to(out, kv.first);
      out += ":[";
      for (size_t i = 0; i < kv.second.size(); ++i) {
        if (i) out += ",";
        json::escape_to(out, kv.second[i]);
      }
      out += "]";
    }
    out += "}}";
    return out;
  } else {
    return "{\"error\":\"unknown predicate\"}";
  }
  std::string out = std::string("{\"fits\":") + ((fails.empty() && !hard_error) ? "true" : "false") + ",\"reasons\":[";
  for (size_t i = 0; i < fails.size(); ++i) {
    if (i) out += ",";
    json::escape_to(out, fails[i]);
  }
  out += "],\"error\":";
  json::escape_to(out, hard_error ? err : "");
  out += "}";
  return out;
}

// isDeviceAMatch (core/extended_resources.go:152-167) on one device.
static std::string run_device_match(const json::Value& req) {
  ksb::ExtendedDevice dev;
  const json::Value* d = req.find("device");
  dev.id = d->get_str("id");
  dev.attributes = ksb::detail::strmap(d->find("attributes"));
  std::vector<ksb::SelectorRequirement> reqs = ksb::detail::reqs(req.find("required"));
  bool ok = false;
  std::string err;
  bool noerr = Oracle::is_device_a_match(dev, reqs, &ok, &err);
  std::string out = std::string("{\"match\":") + ((noerr && ok) ? "true" : "false") + ",\"error\":";
  json::escape_to(out, noerr ? "" : err);
  out += "}";
  return out;
}

// Dump one node's NodeInfo aggregates after loading the cluster (cache tests).
static std::string node_info_json(const NodeInfo& ni) {
  std::string out = "{";
  out += "\"requested\":{\"cpu\":" + std::to_string(ni.requested.MilliCPU) + ",\"memory\":" +
         std::to_string(ni.requested.Memory) + ",\"gpu\":" + std::to_string(ni.requested.NvidiaGPU) +
         ",\"ephemeral\":" + std::to_string(ni.requested.EphemeralStorage) + ",\"scalars\":{";
  bool first = true;
  for (auto& kv : ni.requested.Scalar) {
    if (kv.second == 0) continue;  // a name whose pods are all gone: the Go map keeps a zero entry, nothing reads it
    if (!first) out += ",";
    first = false;
    json::escape_to(out, kv.first);
    out += ":" + std::to_string(kv.second);
  }
  out += "}},\"nonzero\":{\"cpu\":" + std::to_string(ni.nonzero.MilliCPU) + ",\"memory\":" +
         std::to_string(ni.nonzero.Memory) + "},\"pods\":" + std::to_string(ni.pods.size()) + ",\"usedPorts\":[";
  first = true;
  for (auto& p : ni.usedPorts) {
    if (!first) out += ",";
    first = false;
    json::escape_to(out, p);
  }
  // allocatable as SetNode left it (node_info.go:469-491); RemoveNode empties it (:494-506)
  {
    const Resource none;
    const Resource& al = ni.node ? ni.allocatable : none;
    out += "],\"allocatable\":{\"cpu\":" + std::to_string(al.MilliCPU) + ",\"memory\":" + std::to_string(al.Memory) +
           ",\"pods\":" + std::to_string(al.AllowedPodNumber) + ",\"scalars\":{";
    first = true;
    for (auto& kv : al.Scalar) {
      if (!first) out += ",";
      first = false;
      json::escape_to(out, kv.first);
      out += ":" + std::to_string(kv.second);
    }
    out += "}}";
  }
  out += ",\"available\":{";
  first = true;
  for (auto& kv : ni.ext.available) {
    if (!first) out += ",";
    first = false;
    json::escape_to(out, kv.first);
    out += ":[";
    bool f2 = true;
    for (auto& d : kv.second) {
      if (!f2) out += ",";
      f2 = false;
      json::escape_to(out, d.first);
    }
    out += "]";
  }
  out += "}}";
  return out;
}

static std::string run_node_info(const json::Value& req) {
  ksb::Cluster cl = ksb::cluster_from_json(*req.find("cluster"));
  Oracle o;
  o.load(cl);
  for (auto& p : req.get_arr("forget")) o.cache.ForgetPod(ksb::pod_from_json(p));
  return node_info_json(o.cache.nodes.at(req.get_str("node")));
}

// The cache state machine driven by a list of operations (schedulercache/cache.go:118-312, :426-462):
// {"ttl": seconds, "ops": [{"do": "assume"|"finish"|"forget"|"add"|"update"|"remove"|"cleanup"|"add_node"|
// "remove_node", "pod": {...}, "old": {...}, "node": {...}, "name": "...", "now": seconds}], "report": [node names]}
// -> {"errors": [text per op, "" = nil], "assumed": [per op, for "is_assumed" ops], "nodes": {name: NodeInfo | null}}.
static std::string run_cache_ops(const json::Value& req) {
  SchedulerCache c;
  c.ttl = req.get_num("ttl", 30.0);
  std::string errors = "[";
  bool first = true;
  for (auto& op : req.get_arr("ops")) {
    const std::string what = op.get_str("do");
    std::string err;
    if (what == "assume") err = c.assume(ksb::pod_from_json(*op.find("pod")));
    else if (what == "finish") err = c.finish_binding(ksb::pod_from_json(*op.find("pod")), op.get_num("now", 0.0));
    else if (what == "forget") err = c.forget(ksb::pod_from_json(*op.find("pod")));
    else if (what == "add") err = c.add(ksb::pod_from_json(*op.find("pod")));
    else if (what == "update") err = c.update(ksb::pod_from_json(*op.find("old")), ksb::pod_from_json(*op.find("pod")));
    else if (what == "remove") err = c.remove(ksb::pod_from_json(*op.find("pod")));
    else if (what == "cleanup") c.cleanup(op.get_num("now", 0.0));
    else if (what == "is_assumed") err = c.is_assumed(ksb::pod_from_json(*op.find("pod"))) ? "assumed" : "not assumed";
    else if (what == "add_node" || what == "update_node") c.AddNode(ksb::node_from_json(*op.find("node")));
    else if (what == "remove_node") c.RemoveNode(op.get_str("name"));
    else err = "unknown op " + what;
    if (!first) errors += ",";
    first = false;
    json::escape_to(errors, err);
  }
  errors += "]";
  std::string out = "{\"errors\":" + errors + ",\"nodes\":{";
  first = true;
  for (auto& nv : req.get_arr("report")) {
    const std::string name = nv.s;
    if (!first) out += ",";
    first = false;
    json::escape_to(out, name);
    auto it = c.nodes.find(name);
    out += ":" + (it == c.nodes.end() ? std::string("null") : node_info_json(it->second));
  }
  out += "}}";
  return out;
}

extern "C" {

// genericScheduler.Preempt for one pod on a loaded cluster (core/generic_scheduler.go:199-256), see preempt.hpp.
// {"cluster", "pod", "resources_only": bool, "failed": {node: [reasons]} (optional: the FailedPredicateMap to use instead
// of the one the predicates give)} -> {"potential", "candidates": [[node, [victims]]], "node", "victims", "clear_nomination"}
static std::string run_preempt(const json::Value& req) {
  ksb::Cluster cl = ksb::cluster_from_json(*req.find("cluster"));
  ksb::Pod pod = ksb::pod_from_json(*req.find("pod"));
  Oracle o;
  apply_options(o, req);
  o.load(cl);
  o.cachedNodeInfoMap = o.cache.nodes;
  const bool resources_only = req.get_bool("resources_only", false);
  std::vector<NodePtr> nodes = o.node_list();
  std::map<std::string, std::vector<std::string>> failed;
  if (const json::Value* f = req.find("failed"); f && f->is_obj()) {
    for (auto& kv : *f->o) {
      std::vector<std::string> rs;
      for (auto& r : *kv.second.a) rs.push_back(r.s);
      failed[kv.first] = rs;
    }
  } else {
    PredicateMeta meta = o.make_predicate_meta(pod);
    std::vector<PodPtr> allPods = o.all_pods();
    std::vector<Oracle::TermProbe> none;
    for (auto& n : nodes) {
      const NodeInfo& ni = o.cachedNodeInfoMap.at(n->name);
      std::vector<std::string> fails;
      std::string err;
      if (resources_only) Oracle::pod_fits_resources(meta, ni, &fails);
      else if (!o.pod_fits_on_node(pod, meta, ni, allPods, none, none, &fails, &err)) return "{\"error\":\"" + err + "\"}";
      if (!fails.empty()) failed[n->name] = fails;
    }
  }
  PreemptResult r = preempt(o, pod, nodes, failed, resources_only);
  std::string out = "{";
  put_kv(out, "error", r.error);
  out += "\"potential\":[";
  for (size_t i = 0; i < r.potential.size(); ++i) {
    if (i) out += ",";
    json::escape_to(out, r.potential[i]);
  }
  out += "],\"candidates\":[";
  for (size_t i = 0; i < r.candidates.size(); ++i) {
    if (i) out += ",";
    out += "[";
    json::escape_to(out, r.candidates[i].first);
    out += ",[";
    for (size_t k = 0; k < r.candidates[i].second.pods.size(); ++k) {
      if (k) out += ",";
      json::escape_to(out, r.candidates[i].second.pods[k]->key());
    }
    out += "]]";
  }
  out += "],";
  put_kv(out, "node", r.node);
  out += "\"victims\":[";
  for (size_t i = 0; i < r.victims.size(); ++i) {
    if (i) out += ",";
    json::escape_to(out, r.victims[i]);
  }
  out += std::string("],\"clear_nomination\":") + (r.clear_nomination ? "true" : "false") + "}";
  return out;
}

// The small helpers under algorithm/priorities/util that the path leans on (topologies.go:28-75), one call each:
// {"fn": "same_topology", "nodeA", "nodeB", "key"} | {"fn": "term_namespaces", "pod", "term"} |
// {"fn": "ns_selector_match", "pod", "namespaces": [...], "selector": LabelSelector}
static std::string run_helper(const json::Value& req) {
  const std::string fn = req.get_str("fn");
  if (fn == "same_topology") {
    ksb::Node a = ksb::node_from_json(*req.find("nodeA")), b = ksb::node_from_json(*req.find("nodeB"));
    return std::string("{\"result\":") + (nodes_have_same_topology_key(a, b, req.get_str("key")) ? "true" : "false") + "}";
  }
  if (fn == "term_namespaces") {
    ksb::Pod pod = ksb::pod_from_json(*req.find("pod"));
    ksb::PodAffinityTerm term = ksb::detail::pod_term(*req.find("term"));
    std::string out = "{\"result\":[";
    bool first = true;
    for (auto& n : namespaces_from_term(pod, term)) {
      if (!first) out += ",";
      first = false;
      json::escape_to(out, n);
    }
    return out + "]}";
  }
  if (fn == "ns_selector_match") {
    ksb::Pod pod = ksb::pod_from_json(*req.find("pod"));
    std::set<std::string> nss;
    for (auto& n : req.get_arr("namespaces")) nss.insert(n.s);
    Selector sel;
    std::string err;
    if (!label_selector_as_selector(ksb::detail::label_selector(req.find("selector")), &sel, &err)) return "{\"error\":\"selector\"}";
    return std::string("{\"result\":") + (pod_matches_ns_and_selector(pod, nss, sel) ? "true" : "false") + "}";
  }
  if (fn == "nonzero_requests") {  // non_zero.go:36-50 on one container's requests
    ksb::Pod pod = ksb::pod_from_json(*req.find("pod"));
    int64_t cpu = 0, mem = 0;
    get_nonzero_requests(pod.containers.at(0).requests, &cpu, &mem);
    return "{\"result\":[" + std::to_string(cpu) + "," + std::to_string(mem) + "]}";
  }
  if (fn == "priority_meta") {  // priorities/metadata.go:45-66: non-zero request totals and the PreferNoSchedule tolerations
    Oracle o;
    Oracle::PriorityMeta pm = o.make_priority_meta(ksb::pod_from_json(*req.find("pod")));
    std::string out = "{\"result\":{\"nonzero\":[" + std::to_string(pm.nonZeroCPU) + "," + std::to_string(pm.nonZeroMem) + "],\"tolerations\":[";
    for (size_t i = 0; i < pm.tolerationsPreferNoSchedule.size(); ++i) {
      out += i ? ",[" : "[";
      const ksb::Toleration& t = pm.tolerationsPreferNoSchedule[i];
      json::escape_to(out, t.key);
      out += ",";
      json::escape_to(out, t.op);
      out += ",";
      json::escape_to(out, t.value);
      out += ",";
      json::escape_to(out, t.effect);
      out += "]";
    }
    return out + "]}}";
  }
  if (fn == "used_ports") {  // util/utils.go:33-62 over several pods
    std::set<std::string> all;
    for (auto& pj : req.get_arr("pods"))
      for (auto& x : get_used_ports(ksb::pod_from_json(pj))) all.insert(x);
    std::string out = "{\"result\":[";
    bool first = true;
    for (auto& x : all) {
      if (!first) out += ",";
      first = false;
      json::escape_to(out, x);
    }
    return out + "]}";
  }
  if (fn == "tolerates_taint") {  // api/core/v1/toleration.go:30-56
    const json::Value& tj = *req.find("toleration");
    const json::Value& xj = *req.find("taint");
    ksb::Toleration tol;
    tol.key = tj.get_str("key");
    tol.op = tj.get_str("operator");
    tol.value = tj.get_str("value");
    tol.effect = tj.get_str("effect");
    ksb::Taint taint;
    taint.key = xj.get_str("key");
    taint.value = xj.get_str("value");
    taint.effect = xj.get_str("effect");
    return std::string("{\"result\":") + (tolerates_taint(tol, taint) ? "true" : "false") + "}";
  }
  if (fn == "requirements_match") {  // labels.Requirement.Matches (apimachinery/pkg/labels/selector.go:186-230) over a label set
    Selector sel;
    std::string err;
    if (!requirements_as_selector(ksb::detail::reqs(req.find("reqs")), &sel, &err)) return "{\"error\":\"requirements\"}";
    return std::string("{\"result\":") + (sel_matches(sel, ksb::detail::strmap(req.find("labels"))) ? "true" : "false") + "}";
  }
  if (fn == "requirements_selector" || fn == "label_selector") {
    // NodeSelectorRequirementsAsSelector (pkg/apis/core/v1/helper/helpers.go:212-242) / LabelSelectorAsSelector
    // (apimachinery/pkg/apis/meta/v1/helpers.go:31-67); the selector's requirements sorted by key, as labels.Parse leaves them
    Selector sel;
    std::string err;
    const bool ok = fn == "label_selector" ? label_selector_as_selector(ksb::detail::label_selector(req.find("selector")), &sel, &err)
                                           : requirements_as_selector(ksb::detail::reqs(req.find("reqs")), &sel, &err);
    if (!ok) return "{\"result\":{\"error\":true}}";
    std::sort(sel.reqs.begin(), sel.reqs.end(), [](const Requirement& a, const Requirement& b) { return a.key < b.key; });
    static const char* names[] = {"In", "NotIn", "Equals", "Exists", "DoesNotExist", "Gt", "Lt"};
    std::string out = std::string("{\"result\":{\"nothing\":") + (sel.nothing ? "true" : "false") + ",\"reqs\":[";
    for (size_t i = 0; i < sel.reqs.size(); ++i) {
      out += i ? ",[" : "[";
      json::escape_to(out, sel.reqs[i].key);
      out += ",";
      json::escape_to(out, names[(int)sel.reqs[i].op]);
      out += ",[";
      std::vector<std::string> vals = sel.reqs[i].values;
      std::sort(vals.begin(), vals.end());
      for (size_t k = 0; k < vals.size(); ++k) {
        if (k) out += ",";
        json::escape_to(out, vals[k]);
      }
      out += "]]";
    }
    return out + "]}}";
  }
  if (fn == "pod_priority") {  // util/utils.go:76-87
    return "{\"result\":" + std::to_string(ksb::pod_from_json(*req.find("pod")).priority) + "}";
  }
  if (fn == "sort_by_priority") {  // util.SortableList{CompFunc: HigherPriorityPod}.Sort() = go1.9 sort.Sort (util/utils.go:89-135)
    std::vector<int64_t> prio;
    for (auto& x : req.get_arr("priorities")) prio.push_back(x.i);
    std::vector<int> order(prio.size());
    for (size_t i = 0; i < order.size(); ++i) order[i] = (int)i;
    GoSort gs;
    gs.less = [&](int i, int j) { return prio[order[i]] > prio[order[j]]; };
    gs.swap = [&](int i, int j) { std::swap(order[i], order[j]); };
    gs.sort((int)order.size());
    std::string out = "{\"result\":[";
    for (size_t i = 0; i < order.size(); ++i) out += (i ? "," : "") + std::to_string(order[i]);
    return out + "]}";
  }
  return "{\"error\":\"unknown helper\"}";
}

// Generic entry: {"op": "stream" | "priority" | "predicate" | "device_match" | "node_info" | "preempt", ...}.
// Returns a malloc'd NUL-terminated JSON string; free with oracle_free.
char* oracle_call(const char* json_in) {
  std::string out;
  try {
    json::Value req = json::parse(json_in);
    std::string op = req.get_str("op");
    if (op == "stream") out = run_stream(req);
    else if (op == "priority") out = run_priority(req);
    else if (op == "predicate") out = run_predicate(req);
    else if (op == "device_match") out = run_device_match(req);
    else if (op == "node_info") out = run_node_info(req);
    else if (op == "cache_ops") out = run_cache_ops(req);
    else if (op == "preempt") out = run_preempt(req);
    else if (op == "helper") out = run_helper(req);
    else out = "{\"error\":\"unknown op\"}";
  } catch (const std::exception& e) {
    out = "{\"error\":";
    json::escape_to(out, e.what());
    out += "}";
  }
  char* p = (char*)malloc(out.size() + 1);
  memcpy(p, out.c_str(), out.size() + 1);
  return p;
}

void oracle_free(char* p) { free(p); }

}  // extern "C"
