// Host-side mirror of the reference's interfaces for this path, written against the C-ABI
// only (include/ksched_b200.h). The reference is Go and this image has no Go toolchain, so
// the shim a maintainer would write in Go is mirrored here in C++ with the
// same names, argument meaning and error behaviour:
//
//   GpuScheduler      <- algorithm.ScheduleAlgorithm   (algorithm/scheduler_interface.go:48-61),
//                        implemented today by genericScheduler (core/generic_scheduler.go:90-161)
//   CacheDecorator    <- schedulercache.Cache           (schedulercache/interface.go:61-121)
//   FakeNodeLister    <- schedulertesting.FakeNodeLister (testing/fake_lister.go:33-38)
//   FitError          <- core.FitError                   (core/generic_scheduler.go:47-88)
//
// Paths are relative to /root/reference/plugin/pkg/scheduler.
#pragma once
#include <map>
#include <stdexcept>
#include <set>
#include <string>
#include <vector>

#include "../../../include/ksb_wire.hpp"
#include "../../../include/ksched_b200.h"
#include "model.hpp"

namespace ksb {

// v1.ExtendedResourceBinding: pod.Spec.ExtendedResources[i].Name -> device ids
using ExtendedResourceBinding = std::map<std::string, std::vector<std::string>>;

struct FitError : std::runtime_error {  // core/generic_scheduler.go:47-51
  int num_all_nodes;
  explicit FitError(const std::string& msg, int n) : std::runtime_error(msg), num_all_nodes(n) {}
};
struct NoNodesAvailable : std::runtime_error {  // ErrNoNodesAvailable, core/generic_scheduler.go:58
  NoNodesAvailable() : std::runtime_error("no nodes available to schedule pods") {}
};

using FakeNodeLister = std::vector<Node>;

class EngineHandle {
 public:
  explicit EngineHandle(const ksb_config& cfg) {
    char err[512] = {0};
    e_ = ksb_create(&cfg, err, sizeof err);
    if (!e_) throw std::runtime_error(std::string("ksb_create: ") + err);
  }
  ~EngineHandle() { ksb_destroy(e_); }
  EngineHandle(const EngineHandle&) = delete;
  EngineHandle& operator=(const EngineHandle&) = delete;
  ksb_engine* get() const { return e_; }
  void check(int32_t rc, const char* what) const {
    if (rc != 0) throw std::runtime_error(std::string(what) + ": " + ksb_last_error(e_));
  }

 private:
  ksb_engine* e_;
};

// schedulercache.Cache decorator: every mutation the Go scheduler applies to its own cache is
// forwarded to the engine. Method names follow interface.go:61-121.
// schedulercache.Cache over the engine: the pod state machine of cache.go:118-312 (Initial -> Assumed -> Added ->
// Deleted / Expired, expiry :426-462) kept on the host, the node aggregates in the engine. Every transition returns
// the reference's error text ("" = nil) instead of throwing, as the Go methods return `error`. Time is in seconds
// and injected, as the reference's tests do with finishBinding / cleanupAssumedPods.
class CacheDecorator {
 public:
  explicit CacheDecorator(EngineHandle& h, double ttl_seconds = 30.0) : h_(h), ttl_(ttl_seconds) {}
  void AddNode(const Node& n) { upsert(n); }
  void UpdateNode(const Node& /*old*/, const Node& n) { upsert(n); }
  void RemoveNode(const Node& n) { h_.check(ksb_node_remove(h_.get(), n.name.data(), n.name.size()), "RemoveNode"); }

  std::string AssumePod(const Pod& p) {  // :118-137
    const std::string key = p.key();
    if (states_.count(key)) return "pod " + key + " state wasn't initial but get assumed";
    add(p, "AssumePod");
    states_[key].pod = p;
    assumed_.insert(key);
    return "";
  }
  std::string FinishBinding(const Pod& p, double now) {  // :139-160
    auto it = states_.find(p.key());
    if (it != states_.end() && assumed_.count(p.key())) {
      it->second.binding_finished = true;
      it->second.has_deadline = true;
      it->second.deadline = now + ttl_;
    }
    return "";
  }
  std::string ForgetPod(const Pod& p) {  // :162-190
    const std::string key = p.key();
    auto it = states_.find(key);
    if (it != states_.end() && it->second.pod.node_name != p.node_name) return "pod " + key + " state was assumed on a different node";
    if (it != states_.end() && assumed_.count(key)) {
      remove(p, "ForgetPod");
      assumed_.erase(key);
      states_.erase(key);
      return "";
    }
    return "pod " + key + " state wasn't assumed but get forgotten";
  }
  std::string AddPod(const Pod& p) {  // :223-257
    const std::string key = p.key();
    auto it = states_.find(key);
    if (it != states_.end() && assumed_.count(key)) {
      if (it->second.pod.node_name != p.node_name) {  // added to a different node than it was assumed to
        remove(it->second.pod, "AddPod");
        add(p, "AddPod");
      }
      assumed_.erase(key);
      it->second.has_deadline = false;
      it->second.pod = p;
      return "";
    }
    if (it == states_.end()) {  // expired, or never seen: add it (back)
      add(p, "AddPod");
      states_[key].pod = p;
      return "";
    }
    return "pod was already in added state. Pod key: " + key;
  }
  std::string UpdatePod(const Pod& oldp, const Pod& newp) {  // :259-285
    const std::string key = oldp.key();
    auto it = states_.find(key);
    if (it != states_.end() && !assumed_.count(key)) {
      if (it->second.pod.node_name != newp.node_name)
        return "fatal: pod " + key + " updated on a different node than previously added to";
      remove(oldp, "UpdatePod");
      add(newp, "UpdatePod");
      return "";
    }
    return "pod " + key + " state wasn't added but get updated";
  }
  std::string RemovePod(const Pod& p) {  // :287-312
    const std::string key = p.key();
    auto it = states_.find(key);
    if (it != states_.end() && !assumed_.count(key)) {
      if (it->second.pod.node_name != p.node_name)
        return "fatal: pod " + key + " removed from a different node than previously added to";
      remove(it->second.pod, "RemovePod");
      states_.erase(key);
      return "";
    }
    return "pod state wasn't added but get removed. Pod key: " + key;
  }
  bool IsAssumedPod(const Pod& p) const { return assumed_.count(p.key()) != 0; }  // :314-328
  void CleanupAssumedPods(double now) {  // :431-462
    std::vector<std::string> keys(assumed_.begin(), assumed_.end());
    for (auto& key : keys) {
      State& st = states_.at(key);
      if (!st.binding_finished) continue;  // binding still in progress
      if (now > st.deadline) {
        remove(st.pod, "expirePod");
        assumed_.erase(key);
        states_.erase(key);
      }
    }
  }
  // Pods the engine placed and assumed itself (ksb_schedule_batch) enter the state machine as assumed.
  void NoteAssumed(const Pod& p) {
    states_[p.key()].pod = p;
    assumed_.insert(p.key());
  }

 private:
  struct State {
    Pod pod;
    bool binding_finished = false;
    bool has_deadline = false;
    double deadline = 0;
  };
  EngineHandle& h_;
  double ttl_;
  std::map<std::string, State> states_;
  std::set<std::string> assumed_;
  void upsert(const Node& n) {
    wire::Writer w;
    wire::put(w, n);
    h_.check(ksb_node_upsert(h_.get(), (const uint8_t*)w.buf.data(), w.buf.size()), "AddNode");
  }
  void add(const Pod& p, const char* what) {
    wire::Writer w;
    wire::put(w, p);
    h_.check(ksb_pod_add(h_.get(), (const uint8_t*)w.buf.data(), w.buf.size()), what);
  }
  void remove(const Pod& p, const char* what) {
    h_.check(ksb_pod_remove(h_.get(), p.ns.data(), p.ns.size(), p.name.data(), p.name.size()), what);
  }
};

struct ScheduleOutcome {
  std::string host;
  ExtendedResourceBinding binding;
  int num_filtered = 0;
  int64_t score = 0;
};

// algorithm.ScheduleAlgorithm over the engine. Schedule returns (selectedMachine, binding) or
// throws FitError / NoNodesAvailable / runtime_error exactly where the reference returns an error.
class GpuScheduler {
 public:
  explicit GpuScheduler(EngineHandle& h) : h_(h) {}

  ScheduleOutcome Schedule(const Pod& pod, const FakeNodeLister& nodes) {
    if (nodes.empty()) throw NoNodesAvailable();  // generic_scheduler.go:116-118
    wire::Writer w;
    wire::put(w, pod);
    ksb_result r;
    h_.check(ksb_schedule(h_.get(), (const uint8_t*)w.buf.data(), w.buf.size(), &r), "Schedule");
    return outcome(pod, r, 0);
  }

  // Turn a raw result into the reference's return values (throws on the error statuses).
  ScheduleOutcome outcome(const Pod& pod, const ksb_result& r, int index_in_batch) {
    if (r.status == KSB_NO_NODES) throw NoNodesAvailable();
    if (r.status == KSB_FIT_ERROR) {
      char buf[4096];
      h_.check(ksb_fit_error_message(h_.get(), index_in_batch, buf, sizeof buf), "FitError");
      throw FitError(buf, r.num_nodes);
    }
    if (r.status != KSB_SCHEDULED) throw std::runtime_error("Schedule: " + std::string(ksb_last_error(h_.get())));
    ScheduleOutcome o;
    const char* name;
    size_t len;
    h_.check(ksb_node_name(h_.get(), r.node_index, &name, &len), "node name");
    o.host.assign(name, len);
    o.num_filtered = r.num_filtered;
    o.score = r.score;
    for (size_t i = 0; i < pod.extended.size() && i < KSB_MAX_EXT_ENTRIES; ++i) {
      const PodExtendedResource& e = pod.extended[i];
      std::vector<std::string> ids;
      if (e.limits.size() == 1) {
        const std::string& rname = e.limits.begin()->first;
        for (int b = 0; b < KSB_MAX_DEVICES; ++b)
          if ((r.device_mask[i] >> b) & 1u) {
            const char* id;
            size_t idl;
            h_.check(ksb_device_id(h_.get(), r.node_index, rname.data(), rname.size(), b, &id, &idl), "device id");
            ids.emplace_back(id, idl);
          }
      }
      o.binding[e.name] = ids;
    }
    return o;
  }

 private:
  EngineHandle& h_;
};

}  // namespace ksb
