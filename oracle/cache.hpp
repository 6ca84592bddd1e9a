// ORACLE — TEST INFRASTRUCTURE ONLY. CPU restatement of schedulercache
// (NodeInfo, Resource, the fork's ExtendedResourceCacheManager, and the
// assume/add/remove state the hot path reads). Nothing under kubernetes_b200/
// may include or link this.
//
// Follows (paths relative to /root/reference/plugin/pkg/scheduler):
//   schedulercache/node_info.go:68-111 (Resource, Add), :329-356 (AddPod),
//     :359-419 (RemovePod incl. the missing EphemeralStorage subtraction),
//     :421-433 (calculateResource), :435-466 (updateUsedPorts), :469-491 (SetNode)
//   schedulercache/extended_resources.go:39-207 (fork device cache)
//   schedulercache/cache.go:118-137 (AssumePod), :223-312 (Add/Remove), :347-387 (nodes)
//   algorithm/priorities/util/non_zero.go:29-50 (GetNonzeroRequests)
//   pkg/apis/core/v1/helper/helpers.go:32-84 (IsScalarResourceName)
#pragma once
#include <map>
#include <memory>
#include <set>
#include <string>
#include <vector>

#include "labels.hpp"
#include "model.hpp"

namespace oracle {

using ksb::Node;
using ksb::Pod;
using ksb::ResourceList;

struct Resource {
  int64_t MilliCPU = 0, Memory = 0, NvidiaGPU = 0, EphemeralStorage = 0;
  int AllowedPodNumber = 0;
  std::map<std::string, int64_t> Scalar;
};

inline bool is_default_namespace_resource(const std::string& n) {
  return n.find('/') == std::string::npos || n.find("kubernetes.io/") != std::string::npos;
}
inline bool is_hugepage_resource(const std::string& n) { return n.rfind("hugepages-", 0) == 0; }
inline bool is_scalar_resource_name(const std::string& n) {
  return !is_default_namespace_resource(n) || is_hugepage_resource(n);
}

// node_info.go:90-111
inline void resource_add(Resource& r, const ResourceList& rl) {
  for (auto& kv : rl) {
    const std::string& name = kv.first;
    if (name == ksb::kResCPU) r.MilliCPU += kv.second;
    else if (name == ksb::kResMemory) r.Memory += kv.second;
    else if (name == ksb::kResAlphaGPU) r.NvidiaGPU += kv.second;
    else if (name == ksb::kResPods) r.AllowedPodNumber += (int)kv.second;
    else if (name == ksb::kResEphemeral) r.EphemeralStorage += kv.second;
    else if (is_scalar_resource_name(name)) r.Scalar[name] += kv.second;
  }
}

static const int64_t kDefaultMilliCpuRequest = 100;
static const int64_t kDefaultMemoryRequest = 200LL * 1024 * 1024;

// non_zero.go:36-50
inline void get_nonzero_requests(const ResourceList& req, int64_t* cpu, int64_t* mem) {
  auto c = req.find(ksb::kResCPU);
  *cpu = c == req.end() ? kDefaultMilliCpuRequest : c->second;
  auto m = req.find(ksb::kResMemory);
  *mem = m == req.end() ? kDefaultMemoryRequest : m->second;
}

// v1helper.PodExtendedResourceName (helpers.go:500-510)
inline bool pod_extended_resource_name(const ksb::PodExtendedResource& r, std::string* name, std::string* err) {
  if (r.limits.size() != 1) {
    *err = "unexpected limits length: " + std::to_string(r.limits.size()) + " != 1";
    return false;
  }
  *name = r.limits.begin()->first;
  return true;
}

// resource name -> device id -> device. std::map gives the canonical ascending-ID
// iteration order the harness fixes for the reference's random Go map order
// (SURVEY.md §7 hard part 1.ii, §8c "Harness decisions").
using DeviceMap = std::map<std::string, ksb::ExtendedDevice>;
using ExtendedResourceMap = std::map<std::string, DeviceMap>;

// schedulercache/extended_resources.go:39-207
struct ExtendedResourceManager {
  ExtendedResourceMap allocatable, available, used;

  void AddPod(const Pod& pod) {  // :86-110
    for (auto& pres : pod.extended) {
      std::string rname, err;
      if (!pod_extended_resource_name(pres, &rname, &err)) continue;
      auto av = available.find(rname);
      if (av == available.end()) continue;
      for (auto& id : pres.assigned) {
        auto d = av->second.find(id);
        if (d == av->second.end()) continue;
        used[rname][id] = d->second;
        av->second.erase(d);
      }
    }
  }
  void RemovePod(const Pod& pod) {  // :114-152
    for (auto& pres : pod.extended) {
      std::string rname, err;
      if (!pod_extended_resource_name(pres, &rname, &err)) continue;
      auto u = used.find(rname);
      if (u == used.end()) continue;
      auto al = allocatable.find(rname);
      bool is_allocated = al != allocatable.end();
      for (auto& id : pres.assigned) {
        auto d = u->second.find(id);
        if (d == u->second.end()) continue;
        u->second.erase(d);
        if (!is_allocated) continue;
        auto r = al->second.find(id);
        if (r == al->second.end()) continue;
        available[rname][id] = r->second;
      }
    }
  }
  void SetNode(const Node& node) {  // :154-207
    for (auto& kv : allocatable) {
      const std::string& rname = kv.first;
      auto nres = node.extended.find(rname);
      if (nres == node.extended.end()) {
        available[rname] = DeviceMap();
        continue;
      }
      auto av = available.find(rname);
      if (av == available.end()) continue;
      for (auto& dev : kv.second) {
        bool present = false;
        for (auto& d : nres->second)
          if (d.id == dev.first) present = true;
        if (present) continue;
        av->second.erase(dev.first);
      }
    }
    for (auto& kv : node.extended) {
      const std::string& rname = kv.first;
      if (allocatable.find(rname) == allocatable.end()) {
        DeviceMap m;
        for (auto& d : kv.second) m[d.id] = d;
        available[rname] = m;
        allocatable[rname] = m;
        used[rname] = DeviceMap();
        continue;
      }
      for (auto& d : kv.second) {
        if (used[rname].count(d.id)) continue;
        available[rname][d.id] = d;
        allocatable[rname][d.id] = d;
      }
    }
  }
};

struct NodeInfo {
  std::shared_ptr<const Node> node;  // nil until SetNode
  std::vector<std::shared_ptr<const Pod>> pods;
  std::vector<std::shared_ptr<const Pod>> podsWithAffinity;
  std::set<std::string> usedPorts;
  Resource requested, nonzero, allocatable;
  std::vector<ksb::Taint> taints;
  std::string memoryPressure, diskPressure;  // ConditionStatus, "" until seen
  ExtendedResourceManager ext;
  int64_t generation = 0;  // node_info.go:62: bumped on every mutation

  static bool has_pod_affinity_constraints(const Pod& p) {  // :313-316
    return p.affinity.present && (p.affinity.pod.present || p.affinity.anti.present);
  }

  // :421-433 — containers only, init containers are NOT counted here.
  static void calculate_resource(const Pod& pod, Resource* res, int64_t* n0cpu, int64_t* n0mem) {
    *n0cpu = 0;
    *n0mem = 0;
    for (auto& c : pod.containers) {
      resource_add(*res, c.requests);
      int64_t a, b;
      get_nonzero_requests(c.requests, &a, &b);
      *n0cpu += a;
      *n0mem += b;
    }
  }

  void update_used_ports(const Pod& pod, bool used) {  // :435-466
    for (auto& c : pod.containers)
      for (auto& p : c.ports) {
        if (p.host_port == 0) continue;
        std::string proto = p.protocol.empty() ? "TCP" : p.protocol;
        std::string ip = p.host_ip.empty() ? "0.0.0.0" : p.host_ip;
        std::string s = proto + "/" + ip + "/" + std::to_string(p.host_port);
        if (used) usedPorts.insert(s);
        else usedPorts.erase(s);
      }
  }

  void AddPod(std::shared_ptr<const Pod> pod) {  // :329-356
    Resource res;
    int64_t n0cpu, n0mem;
    calculate_resource(*pod, &res, &n0cpu, &n0mem);
    requested.MilliCPU += res.MilliCPU;
    requested.Memory += res.Memory;
    requested.NvidiaGPU += res.NvidiaGPU;
    requested.EphemeralStorage += res.EphemeralStorage;
    for (auto& kv : res.Scalar) requested.Scalar[kv.first] += kv.second;
    nonzero.MilliCPU += n0cpu;
    nonzero.Memory += n0mem;
    pods.push_back(pod);
    if (has_pod_affinity_constraints(*pod)) podsWithAffinity.push_back(pod);
    ext.AddPod(*pod);
    update_used_ports(*pod, true);
    generation++;
  }

  bool RemovePod(const Pod& pod) {  // :359-419
    std::string k1 = pod.key();
    for (size_t i = 0; i < podsWithAffinity.size(); ++i)
      if (podsWithAffinity[i]->key() == k1) {
        podsWithAffinity[i] = podsWithAffinity.back();
        podsWithAffinity.pop_back();
        break;
      }
    for (size_t i = 0; i < pods.size(); ++i) {
      if (pods[i]->key() != k1) continue;
      pods[i] = pods.back();
      pods.pop_back();
      Resource res;
      int64_t n0cpu, n0mem;
      calculate_resource(pod, &res, &n0cpu, &n0mem);
      requested.MilliCPU -= res.MilliCPU;
      requested.Memory -= res.Memory;
      requested.NvidiaGPU -= res.NvidiaGPU;
      // reference quirk: EphemeralStorage is never subtracted (node_info.go:396-398)
      for (auto& kv : res.Scalar) requested.Scalar[kv.first] -= kv.second;
      nonzero.MilliCPU -= n0cpu;
      nonzero.Memory -= n0mem;
      ext.RemovePod(pod);
      update_used_ports(pod, false);
      generation++;
      return true;
    }
    return false;
  }

  void SetNode(std::shared_ptr<const Node> n) {  // :469-491
    node = n;
    ext.SetNode(*n);
    allocatable = Resource();
    resource_add(allocatable, n->allocatable);
    taints = n->taints;
    for (auto& c : n->conditions) {
      if (c.type == "MemoryPressure") memoryPressure = c.status;
      else if (c.type == "DiskPressure") diskPressure = c.status;
    }
    generation++;
  }
};

// cache.go — only the transitions the in-memory harness drives (no TTL expiry:
// SURVEY.md §8c "all binds succeed, no expiry").
struct SchedulerCache {
  std::map<std::string, NodeInfo> nodes;
  std::map<std::string, std::shared_ptr<const Pod>> podStates;

  void AddNode(const Node& n) { nodes[n.name].SetNode(std::make_shared<const Node>(n)); }  // :347-358
  void UpdateNode(const Node& n) { AddNode(n); }                                             // :360-371
  void RemoveNode(const std::string& name) {  // :373-387: NodeInfo survives while it still has pods
    auto it = nodes.find(name);
    if (it == nodes.end()) return;
    it->second.node.reset();
    it->second.generation++;
    if (it->second.pods.empty()) nodes.erase(it);
  }
  bool AssumePod(const Pod& p) {  // :118-137 (+ addPod :262-270)
    std::string key = p.key();
    if (podStates.count(key)) return false;
    auto sp = std::make_shared<const Pod>(p);
    nodes[p.node_name].AddPod(sp);
    podStates[key] = sp;
    return true;
  }
  bool AddPod(const Pod& p) { return AssumePod(p); }
  bool ForgetPod(const Pod& p) {  // :170-197 (+ removePod :283-293)
    auto st = podStates.find(p.key());
    if (st == podStates.end()) return false;
    auto it = nodes.find(st->second->node_name);
    if (it != nodes.end()) {
      it->second.RemovePod(*st->second);
      if (it->second.pods.empty() && !it->second.node) nodes.erase(it);
    }
    podStates.erase(st);
    return true;
  }
  bool RemovePod(const Pod& p) { return ForgetPod(p); }

  // ---- the full pod state machine (cache.go:118-312, expiry :426-462): Initial -> Assumed -> Added -> Deleted /
  //      Expired. Every transition returns the reference's error text, "" on success. Time is seconds.
  struct PodState {
    std::shared_ptr<const Pod> pod;
    bool bindingFinished = false;
    bool hasDeadline = false;
    double deadline = 0;
  };
  std::map<std::string, PodState> states;
  std::set<std::string> assumedPods;
  double ttl = 30;

  void add_pod_(std::shared_ptr<const Pod> sp) {  // addPod :262-270
    nodes[sp->node_name].AddPod(sp);
    podStates[sp->key()] = sp;
  }
  void remove_pod_(const Pod& p) {  // removePod :283-293 (subtracts the resources of the object it is given)
    auto it = nodes.find(p.node_name);
    if (it == nodes.end()) return;
    it->second.RemovePod(p);
    if (it->second.pods.empty() && !it->second.node) nodes.erase(it);
    podStates.erase(p.key());
  }
  std::string assume(const Pod& p) {  // :118-137
    const std::string key = p.key();
    if (states.count(key)) return "pod " + key + " state wasn't initial but get assumed";
    auto sp = std::make_shared<const Pod>(p);
    add_pod_(sp);
    states[key].pod = sp;
    assumedPods.insert(key);
    return "";
  }
  std::string finish_binding(const Pod& p, double now) {  // :139-160
    auto it = states.find(p.key());
    if (it != states.end() && assumedPods.count(p.key())) {
      it->second.bindingFinished = true;
      it->second.hasDeadline = true;
      it->second.deadline = now + ttl;
    }
    return "";
  }
  std::string forget(const Pod& p) {  // :162-190
    const std::string key = p.key();
    auto it = states.find(key);
    if (it != states.end() && it->second.pod->node_name != p.node_name)
      return "pod " + key + " state was assumed on a different node";
    if (it != states.end() && assumedPods.count(key)) {
      remove_pod_(p);
      assumedPods.erase(key);
      states.erase(key);
      return "";
    }
    return "pod " + key + " state wasn't assumed but get forgotten";
  }
  std::string add(const Pod& p) {  // :223-257
    const std::string key = p.key();
    auto it = states.find(key);
    auto sp = std::make_shared<const Pod>(p);
    if (it != states.end() && assumedPods.count(key)) {
      if (it->second.pod->node_name != p.node_name) {  // added to a different node than it was assumed to
        remove_pod_(*it->second.pod);
        add_pod_(sp);
      }
      assumedPods.erase(key);
      it->second.hasDeadline = false;
      it->second.pod = sp;
      return "";
    }
    if (it == states.end()) {  // expired (or never seen): add it back
      add_pod_(sp);
      states[key].pod = sp;
      return "";
    }
    return "pod was already in added state. Pod key: " + key;
  }
  std::string update(const Pod& oldp, const Pod& newp) {  // :259-285
    const std::string key = oldp.key();
    auto it = states.find(key);
    if (it != states.end() && !assumedPods.count(key)) {
      if (it->second.pod->node_name != newp.node_name)
        return "fatal: pod " + key + " updated on a different node than previously added to";
      remove_pod_(oldp);
      add_pod_(std::make_shared<const Pod>(newp));
      return "";
    }
    return "pod " + key + " state wasn't added but get updated";
  }
  std::string remove(const Pod& p) {  // :287-312
    const std::string key = p.key();
    auto it = states.find(key);
    if (it != states.end() && !assumedPods.count(key)) {
      if (it->second.pod->node_name != p.node_name)
        return "fatal: pod " + key + " removed from a different node than previously added to";
      remove_pod_(*it->second.pod);
      states.erase(key);
      return "";
    }
    return "pod state wasn't added but get removed. Pod key: " + key;
  }
  bool is_assumed(const Pod& p) const { return assumedPods.count(p.key()) != 0; }  // :314-328
  void cleanup(double now) {  // cleanupAssumedPods :431-453, expirePod :455-462
    std::vector<std::string> keys(assumedPods.begin(), assumedPods.end());
    for (auto& key : keys) {
      PodState& ps = states.at(key);
      if (!ps.bindingFinished) continue;
      if (now > ps.deadline) {
        remove_pod_(*ps.pod);
        assumedPods.erase(key);
        states.erase(key);
      }
    }
  }
};

}  // namespace oracle
