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
  // A thin wrapper: the state machine itself is the library's (ksb_cache_*, include/ksched_b200.h), so that the
  // vectors of schedulercache/cache_test.go run against the exported entry points.
  explicit CacheDecorator(EngineHandle& h, double ttl_seconds = 30.0) : h_(h) { ksb_cache_set_ttl(h_.get(), ttl_seconds); }
  void AddNode(const Node& n) { upsert(n); }
  void UpdateNode(const Node& /*old*/, const Node& n) { upsert(n); }
  void RemoveNode(const Node& n) { h_.check(ksb_node_remove(h_.get(), n.name.data(), n.name.size()), "RemoveNode"); }

  std::string AssumePod(const Pod& p) { return call(p, [&](const uint8_t* b, size_t n) { return ksb_cache_assume_pod(h_.get(), b, n); }); }
  std::string FinishBinding(const Pod& p, double now) {
    return call(p, [&](const uint8_t* b, size_t n) { return ksb_cache_finish_binding(h_.get(), b, n, now); });
  }
  std::string ForgetPod(const Pod& p) { return call(p, [&](const uint8_t* b, size_t n) { return ksb_cache_forget_pod(h_.get(), b, n); }); }
  std::string AddPod(const Pod& p) { return call(p, [&](const uint8_t* b, size_t n) { return ksb_cache_add_pod(h_.get(), b, n); }); }
  std::string UpdatePod(const Pod& oldp, const Pod& newp) {
    wire::Writer wo;
    wire::put(wo, oldp);
    return call(newp, [&](const uint8_t* b, size_t n) {
      return ksb_cache_update_pod(h_.get(), (const uint8_t*)wo.buf.data(), wo.buf.size(), b, n);
    });
  }
  std::string RemovePod(const Pod& p) { return call(p, [&](const uint8_t* b, size_t n) { return ksb_cache_remove_pod(h_.get(), b, n); }); }
  bool IsAssumedPod(const Pod& p) const {
    int32_t a = 0;
    wire::Writer w;
    wire::put(w, p);
    ksb_cache_is_assumed_pod(h_.get(), (const uint8_t*)w.buf.data(), w.buf.size(), &a);
    return a != 0;
  }
  void CleanupAssumedPods(double now) { h_.check(ksb_cache_cleanup_assumed(h_.get(), now), "cleanupAssumedPods"); }
  bool SkipPodUpdate(const Pod& p) const {  // factory.go:306-353
    int32_t s = 0;
    wire::Writer w;
    wire::put(w, p);
    ksb_cache_skip_pod_update(h_.get(), (const uint8_t*)w.buf.data(), w.buf.size(), &s);
    return s != 0;
  }

 private:
  EngineHandle& h_;
  template <class F>
  std::string call(const Pod& p, F&& f) {
    wire::Writer w;
    wire::put(w, p);
    if (f((const uint8_t*)w.buf.data(), w.buf.size()) == 0) return "";
    return ksb_last_error(h_.get());
  }
  void upsert(const Node& n) {
    wire::Writer w;
    wire::put(w, n);
    h_.check(ksb_node_upsert(h_.get(), (const uint8_t*)w.buf.data(), w.buf.size()), "AddNode");
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
