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
class CacheDecorator {
 public:
  explicit CacheDecorator(EngineHandle& h) : h_(h) {}
  void AddNode(const Node& n) { upsert(n); }
  void UpdateNode(const Node& /*old*/, const Node& n) { upsert(n); }
  void RemoveNode(const Node& n) { h_.check(ksb_node_remove(h_.get(), n.name.data(), n.name.size()), "RemoveNode"); }
  void AssumePod(const Pod& p) { add(p, "AssumePod"); }
  void AddPod(const Pod& p) { add(p, "AddPod"); }
  void ForgetPod(const Pod& p) { remove(p, "ForgetPod"); }
  void RemovePod(const Pod& p) { remove(p, "RemovePod"); }

 private:
  EngineHandle& h_;
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
