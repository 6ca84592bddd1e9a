// extern "C" surface of include/ksched_b200.h over ksb::Engine.
#include <chrono>
#include <cstdio>
#include <functional>
#include <map>
#include <memory>
#include <mutex>
#include <thread>
#include <cstring>
#include <stdexcept>
#include <string>
#include <vector>

#include "../../../include/ksb_wire.hpp"
#include "../../../include/ksched_b200.h"
#include "engine.hpp"

// Delta queue (the reference's only lock scope is the cache's own short mutex, schedulercache/cache.go:
// 83-97): a mutator that arrives while a scheduling call works on the table does not wait for it. It is queued under a
// lock held for a push_back and applied, in arrival order, when that call ends or before the next one starts, which is
// the reference's snapshot point (UpdateNodeNameToInfoMap, core/generic_scheduler.go:122). With no such call running,
// and nothing queued before it, a mutator is applied on the spot and reports its own failure.
struct ksb_engine {
  ksb::Engine* impl;
  std::mutex err_mu;  // mutators may fail on several threads at once (include/ksched_b200.h "Threading")
  std::string err;
  std::mutex q_mu;
  bool busy = false;  // a scheduling call (or its drain of the queue) owns the engine
  std::vector<std::function<void()>> queue;
  std::vector<std::string> deferred_errors;  // failures of queued mutators, kept for ksb_deferred_errors
  // schedulercache.Cache's pod state machine (schedulercache/cache.go:44-66, :118-462) behind the ksb_cache_* entry
  // points: which pods are assumed / added, on which node, and when an assumed pod expires. It decides at call time
  // (its errors are the call's errors); what it does to the table goes through the delta queue like any mutator.
  struct PodState {
    std::string ns, name, node;
    bool assumed = false, binding_finished = false;
    double deadline = 0;
    std::string record;  // the pod's wire record as it came in: what skipPodUpdate compares, NodeName aside (factory.go:312-353)
  };
  std::mutex cache_mu;
  std::map<std::string, PodState> pod_states;
  double cache_ttl = 30.0;     // factory.go:171: schedulercache.New(30*time.Second, ...)
  bool cache_tracking = false;  // set by the first ksb_cache_* call: from then on ksb_schedule_batch's placements enter as assumed
};

namespace {
template <class F>
int32_t guarded(ksb_engine* e, F&& f) {
  if (!e || !e->impl) return -1;
  try {
    f();
    return 0;
  } catch (const std::exception& ex) {
    std::lock_guard<std::mutex> g(e->err_mu);
    e->err = ex.what();
    return -1;
  }
}

// apply queued mutators in arrival order; a failure is kept (and logged), the rest still apply
void apply_deltas(ksb_engine* e, std::vector<std::function<void()>>& take) {
  for (auto& f : take) {
    try {
      f();
    } catch (const std::exception& ex) {
      fprintf(stderr, "[ksb] queued cache update failed: %s\n", ex.what());
      std::lock_guard<std::mutex> g(e->q_mu);
      if (e->deferred_errors.size() < 256) e->deferred_errors.push_back(ex.what());
    }
  }
  take.clear();
}

// a mutator: on the spot if the engine is free and nothing is queued ahead of it, else queued
template <class F>
int32_t mutate(ksb_engine* e, F&& make_delta) {
  if (!e || !e->impl) return -1;
  std::function<void()> delta;
  try {
    delta = make_delta();  // decodes the wire record on the caller's thread; the closure owns what it needs
  } catch (const std::exception& ex) {
    std::lock_guard<std::mutex> g(e->err_mu);
    e->err = ex.what();
    return -1;
  }
  std::unique_lock<std::mutex> q(e->q_mu);
  if (e->busy || !e->queue.empty()) {
    e->queue.push_back(std::move(delta));
    return 0;
  }
  // free: apply under the queue lock, so that a scheduling call that starts now waits for this one update only
  try {
    delta();
    return 0;
  } catch (const std::exception& ex) {
    q.unlock();
    std::lock_guard<std::mutex> g(e->err_mu);
    e->err = ex.what();
    return -1;
  }
}

// a call that reads or schedules against the table: takes the engine, applies what was queued, runs, and applies what
// arrived meanwhile before giving the engine back
template <class F>
int32_t scheduling(ksb_engine* e, F&& f) {
  if (!e || !e->impl) return -1;
  std::vector<std::function<void()>> take;
  {
    std::unique_lock<std::mutex> q(e->q_mu);
    while (e->busy) {  // another scheduling call: the header asks for one at a time; wait rather than interleave
      q.unlock();
      std::this_thread::yield();
      q.lock();
    }
    e->busy = true;
    take.swap(e->queue);
  }
  apply_deltas(e, take);
  const int32_t rc = guarded(e, f);
  for (;;) {
    {
      std::lock_guard<std::mutex> q(e->q_mu);
      if (e->queue.empty()) {
        e->busy = false;
        break;
      }
      take.swap(e->queue);
    }
    apply_deltas(e, take);
  }
  return rc;
}
}  // namespace

// ---- schedulercache.Cache (schedulercache/interface.go:61-121): the state machine of cache.go, exported ----
namespace {
struct CacheFail : std::runtime_error {
  using std::runtime_error::runtime_error;
};
std::string blank_node_record(ksb::Pod p) {
  p.node_name.clear();
  ksb::wire::Writer w;
  ksb::wire::put(w, p);
  return std::string(w.buf.begin(), w.buf.end());
}
ksb::Pod decode_pod(const uint8_t* buf, size_t len) {
  ksb::wire::Reader r(buf, len);
  return ksb::wire::get_pod(r);
}
// engine side of a transition: through the delta queue; a failure on the spot is the call's failure
void engine_add(ksb_engine* e, const ksb::Pod& p) {
  auto sp = std::make_shared<ksb::Pod>(p);
  if (mutate(e, [&]() -> std::function<void()> { return [e, sp] { e->impl->pod_add(*sp); }; }) != 0)
    throw CacheFail(ksb_last_error(e));
}
void engine_remove(ksb_engine* e, const std::string& ns, const std::string& name) {
  if (mutate(e, [&]() -> std::function<void()> { return [e, ns, name] { e->impl->pod_remove(ns, name); }; }) != 0)
    throw CacheFail(ksb_last_error(e));
}
template <class F>
int32_t cache_call(ksb_engine* e, F&& f) {
  if (!e || !e->impl) return -1;
  std::lock_guard<std::mutex> g(e->cache_mu);
  e->cache_tracking = true;
  try {
    f();
    return 0;
  } catch (const std::exception& ex) {
    std::lock_guard<std::mutex> g2(e->err_mu);
    e->err = ex.what();
    return -1;
  }
}
ksb_engine::PodState make_state(const ksb::Pod& p, bool assumed, const uint8_t* buf, size_t len) {
  ksb_engine::PodState st;
  st.ns = p.ns;
  st.name = p.name;
  st.node = p.node_name;
  st.assumed = assumed;
  st.record.assign((const char*)buf, len);
  return st;
}
}  // namespace

extern "C" {

void ksb_config_default(ksb_config* c) {
  memset(c, 0, sizeof *c);
  c->device = 0;
  c->w_least_requested = 1;
  c->w_balanced = 1;
  c->w_selector_spread = 1;
  c->w_node_affinity = 1;
  c->w_taint_toleration = 1;
  c->w_inter_pod_affinity = 1;
  c->w_prefer_avoid = 10000;
  c->hard_pod_affinity_weight = 1;
  c->label_words = 4;
  c->shard_rank = 0;
  c->shard_count = 1;
  c->group_size = 0;
  c->pipeline = 0;
}

ksb_engine* ksb_create(const ksb_config* cfg, char* err, size_t err_len) {
  try {
    ksb_engine* e = new ksb_engine();
    e->impl = new ksb::Engine(*cfg);
    return e;
  } catch (const std::exception& ex) {
    if (err && err_len) {
      strncpy(err, ex.what(), err_len - 1);
      err[err_len - 1] = 0;
    }
    return nullptr;
  }
}

void ksb_destroy(ksb_engine* e) {
  if (!e) return;
  delete e->impl;
  delete e;
}

const char* ksb_last_error(const ksb_engine* e) {
  if (!e) return "null engine";
  // a per-thread copy: the engine's text may be replaced by a failing call on another thread
  static thread_local std::string mine;
  ksb_engine* m = const_cast<ksb_engine*>(e);
  std::lock_guard<std::mutex> g(m->err_mu);
  mine = m->err;
  return mine.c_str();
}

int32_t ksb_node_upsert(ksb_engine* e, const uint8_t* buf, size_t len) {
  return mutate(e, [&]() -> std::function<void()> {
    ksb::wire::Reader r(buf, len);
    auto nd = std::make_shared<ksb::Node>(ksb::wire::get_node(r));
    return [e, nd] { e->impl->node_upsert(*nd); };
  });
}

int32_t ksb_node_remove(ksb_engine* e, const char* name, size_t name_len) {
  return mutate(e, [&]() -> std::function<void()> {
    std::string nm(name ? name : "", name ? name_len : 0);
    return [e, nm] { e->impl->node_remove(nm); };
  });
}

int32_t ksb_pod_add(ksb_engine* e, const uint8_t* buf, size_t len) {
  return mutate(e, [&]() -> std::function<void()> {
    ksb::wire::Reader r(buf, len);
    auto p = std::make_shared<ksb::Pod>(ksb::wire::get_pod(r));
    return [e, p] { e->impl->pod_add(*p); };
  });
}

int32_t ksb_pod_remove(ksb_engine* e, const char* ns, size_t ns_len, const char* name, size_t name_len) {
  return mutate(e, [&]() -> std::function<void()> {
    std::string a(ns ? ns : "", ns ? ns_len : 0), b(name ? name : "", name ? name_len : 0);
    return [e, a, b] { e->impl->pod_remove(a, b); };
  });
}

int32_t ksb_selector_owner_upsert(ksb_engine* e, int32_t kind, const uint8_t* buf, size_t len) {
  return mutate(e, [&]() -> std::function<void()> {
    ksb::wire::Reader r(buf, len);
    auto o = std::make_shared<ksb::SelectorOwner>(ksb::wire::get_owner(r));
    return [e, kind, o] { e->impl->owner_upsert(kind, *o); };
  });
}

int32_t ksb_selector_owner_remove(ksb_engine* e, int32_t kind, const char* ns, size_t ns_len, const char* name,
                                  size_t name_len) {
  return mutate(e, [&]() -> std::function<void()> {
    std::string a(ns ? ns : "", ns ? ns_len : 0), b(name ? name : "", name ? name_len : 0);
    return [e, kind, a, b] { e->impl->owner_remove(kind, a, b); };
  });
}

int32_t ksb_schedule(ksb_engine* e, const uint8_t* pod, size_t len, ksb_result* out) {
  return scheduling(e, [&] {
    ksb::wire::Reader r(pod, len);
    std::vector<ksb::Pod> batch;
    batch.push_back(ksb::wire::get_pod(r));
    e->impl->schedule_batch(batch, out, false);
  });
}

int32_t ksb_schedule_batch(ksb_engine* e, const uint8_t* pods, const uint64_t* offsets, int32_t n, ksb_result* out) {
  return scheduling(e, [&] {
    auto t0 = std::chrono::steady_clock::now();
    // wire records are independent: decode them on a few host threads when the batch is large
    std::vector<ksb::Pod> batch((size_t)std::max(0, n));
    const unsigned hw = std::max(1u, std::thread::hardware_concurrency());
    const int nthreads = n >= 4096 ? (int)std::min(16u, hw) : 1;  // (nothing else of this engine runs yet)
    std::vector<std::string> errors((size_t)nthreads);
    auto decode_range = [&](int t) {
      try {
        const int32_t lo = (int32_t)((int64_t)n * t / nthreads), hi = (int32_t)((int64_t)n * (t + 1) / nthreads);
        for (int32_t i = lo; i < hi; ++i) {
          ksb::wire::Reader r(pods + offsets[i], (size_t)(offsets[i + 1] - offsets[i]));
          batch[i] = ksb::wire::get_pod(r);
        }
      } catch (const std::exception& ex) {
        errors[t] = ex.what();
        if (errors[t].empty()) errors[t] = "decode error";
      }
    };
    if (nthreads == 1) {
      decode_range(0);
    } else {
      std::vector<std::thread> pool;
      for (int t = 0; t < nthreads; ++t) pool.emplace_back(decode_range, t);
      for (auto& th : pool) th.join();
    }
    for (auto& msg : errors)
      if (!msg.empty()) throw std::runtime_error(msg);
    const double decode_ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
    // callers that drive the cache through ksb_cache_*: the pods this call places and assumes enter its state machine
    // as assumed (scheduler.go:478 calls AssumePod right after Schedule; here both happen inside the batch)
    bool tracking;
    {
      std::lock_guard<std::mutex> g(e->cache_mu);
      tracking = e->cache_tracking;
    }
    std::vector<ksb_engine::PodState> noted;
    if (tracking) {
      noted.resize(batch.size());
      for (size_t i = 0; i < batch.size(); ++i) noted[i] = make_state(batch[i], true, pods + offsets[i], (size_t)(offsets[i + 1] - offsets[i]));
    }
    e->impl->schedule_batch(batch, out, true);
    e->impl->stats.last_batch_decode_ms = decode_ms;
    if (tracking) {
      std::lock_guard<std::mutex> g(e->cache_mu);
      for (size_t i = 0; i < noted.size(); ++i) {
        if (out[i].status != KSB_SCHEDULED) continue;
        const std::string* host = e->impl->node_name_at(out[i].node_index);
        if (!host) continue;
        noted[i].node = *host;
        e->pod_states[noted[i].ns + "/" + noted[i].name] = std::move(noted[i]);
      }
    }
  });
}

int32_t ksb_node_name(const ksb_engine* e, int32_t index, const char** name, size_t* len) {
  if (!e || !e->impl) return -1;
  const std::string* s = e->impl->node_name_at(index);
  if (!s) return -1;
  *name = s->data();
  *len = s->size();
  return 0;
}

int32_t ksb_device_id(const ksb_engine* e, int32_t index, const char* resource, size_t resource_len, int32_t bit,
                      const char** id, size_t* len) {
  if (!e || !e->impl) return -1;
  const std::string* s = e->impl->device_id_at(index, std::string(resource, resource_len), bit);
  if (!s) return -1;
  *id = s->data();
  *len = s->size();
  return 0;
}

int32_t ksb_fit_error_message(ksb_engine* e, int32_t i, char* buf, size_t buf_len) {
  return guarded(e, [&] {
    std::string m = e->impl->fit_error_message(i);
    if (buf && buf_len) {
      strncpy(buf, m.c_str(), buf_len - 1);
      buf[buf_len - 1] = 0;
    }
  });
}

int32_t ksb_failed_predicates(ksb_engine* e, uint32_t* masks, int32_t n) {
  return guarded(e, [&] {
    if (!masks) throw std::runtime_error("masks must not be null");
    e->impl->failed_predicates(masks, n);
  });
}

int32_t ksb_resync_begin(ksb_engine* e) {
  return scheduling(e, [&] { e->impl->resync_begin(); });
}

int32_t ksb_resync_end(ksb_engine* e) {
  return scheduling(e, [&] { e->impl->flush(); });
}

int32_t ksb_scores(ksb_engine* e, int64_t* scores, int32_t n) {
  return guarded(e, [&] {
    if (!scores) throw std::runtime_error("scores must not be null");
    e->impl->scores(scores, n);
  });
}

int32_t ksb_reason_name(ksb_engine* e, int32_t bit, char* buf, size_t buf_len) {
  return guarded(e, [&] {
    std::string t = e->impl->reason_name(bit);
    if (buf && buf_len) {
      strncpy(buf, t.c_str(), buf_len - 1);
      buf[buf_len - 1] = 0;
    }
  });
}

int32_t ksb_node_info(ksb_engine* e, const char* name, size_t name_len, char* buf, size_t buf_len, size_t* len) {
  bool small = false;
  int32_t rc = scheduling(e, [&] {
    std::string t = e->impl->node_info_json(std::string(name ? name : "", name_len));
    if (len) *len = t.size();
    if (!buf || buf_len < t.size() + 1) {
      small = true;
      return;
    }
    memcpy(buf, t.c_str(), t.size() + 1);
  });
  return rc != 0 ? rc : (small ? 1 : 0);
}

// ---- the exported state machine ----
void ksb_cache_set_ttl(ksb_engine* e, double seconds) {
  if (!e) return;
  std::lock_guard<std::mutex> g(e->cache_mu);
  e->cache_ttl = seconds;
  e->cache_tracking = true;
}

int32_t ksb_cache_assume_pod(ksb_engine* e, const uint8_t* buf, size_t len) {  // cache.go:118-137
  return cache_call(e, [&] {
    ksb::Pod p = decode_pod(buf, len);
    const std::string key = p.key();
    if (e->pod_states.count(key)) throw CacheFail("pod " + key + " state wasn't initial but get assumed");
    engine_add(e, p);
    e->pod_states[key] = make_state(p, true, buf, len);
  });
}

int32_t ksb_cache_finish_binding(ksb_engine* e, const uint8_t* buf, size_t len, double now_seconds) {  // cache.go:139-160
  return cache_call(e, [&] {
    ksb::Pod p = decode_pod(buf, len);
    auto it = e->pod_states.find(p.key());
    if (it != e->pod_states.end() && it->second.assumed) {
      it->second.binding_finished = true;
      it->second.deadline = now_seconds + e->cache_ttl;
    }
  });
}

int32_t ksb_cache_forget_pod(ksb_engine* e, const uint8_t* buf, size_t len) {  // cache.go:162-190
  return cache_call(e, [&] {
    ksb::Pod p = decode_pod(buf, len);
    const std::string key = p.key();
    auto it = e->pod_states.find(key);
    if (it != e->pod_states.end() && it->second.node != p.node_name)
      throw CacheFail("pod " + key + " state was assumed on a different node");
    if (it != e->pod_states.end() && it->second.assumed) {
      engine_remove(e, p.ns, p.name);
      e->pod_states.erase(it);
      return;
    }
    throw CacheFail("pod " + key + " state wasn't assumed but get forgotten");
  });
}

int32_t ksb_cache_add_pod(ksb_engine* e, const uint8_t* buf, size_t len) {  // cache.go:223-257
  return cache_call(e, [&] {
    ksb::Pod p = decode_pod(buf, len);
    const std::string key = p.key();
    auto it = e->pod_states.find(key);
    if (it != e->pod_states.end() && it->second.assumed) {
      if (it->second.node != p.node_name) {  // added to a different node than it was assumed to
        engine_remove(e, it->second.ns, it->second.name);
        engine_add(e, p);
      }
      it->second = make_state(p, false, buf, len);
      return;
    }
    if (it == e->pod_states.end()) {  // expired, or never seen: add it (back)
      engine_add(e, p);
      e->pod_states[key] = make_state(p, false, buf, len);
      return;
    }
    throw CacheFail("pod was already in added state. Pod key: " + key);
  });
}

int32_t ksb_cache_update_pod(ksb_engine* e, const uint8_t* old_buf, size_t old_len, const uint8_t* new_buf, size_t new_len) {
  return cache_call(e, [&] {  // cache.go:259-285
    ksb::Pod oldp = decode_pod(old_buf, old_len), newp = decode_pod(new_buf, new_len);
    const std::string key = oldp.key();
    auto it = e->pod_states.find(key);
    if (it != e->pod_states.end() && !it->second.assumed) {
      if (it->second.node != newp.node_name)
        throw CacheFail("fatal: pod " + key + " updated on a different node than previously added to");
      engine_remove(e, oldp.ns, oldp.name);
      engine_add(e, newp);
      // (the reference leaves the cached object stale here, cache.go:269-281; only NodeName is read from it later)
      return;
    }
    throw CacheFail("pod " + key + " state wasn't added but get updated");
  });
}

int32_t ksb_cache_remove_pod(ksb_engine* e, const uint8_t* buf, size_t len) {  // cache.go:287-312
  return cache_call(e, [&] {
    ksb::Pod p = decode_pod(buf, len);
    const std::string key = p.key();
    auto it = e->pod_states.find(key);
    if (it != e->pod_states.end() && !it->second.assumed) {
      if (it->second.node != p.node_name)
        throw CacheFail("fatal: pod " + key + " removed from a different node than previously added to");
      engine_remove(e, it->second.ns, it->second.name);
      e->pod_states.erase(it);
      return;
    }
    throw CacheFail("pod state wasn't added but get removed. Pod key: " + key);
  });
}

int32_t ksb_cache_is_assumed_pod(ksb_engine* e, const uint8_t* buf, size_t len, int32_t* assumed) {  // cache.go:314-328
  return cache_call(e, [&] {
    ksb::Pod p = decode_pod(buf, len);
    auto it = e->pod_states.find(p.key());
    if (assumed) *assumed = it != e->pod_states.end() && it->second.assumed ? 1 : 0;
  });
}

int32_t ksb_cache_cleanup_assumed(ksb_engine* e, double now_seconds) {  // cache.go:431-462
  return cache_call(e, [&] {
    for (auto it = e->pod_states.begin(); it != e->pod_states.end();) {
      ksb_engine::PodState& st = it->second;
      if (st.assumed && st.binding_finished && now_seconds > st.deadline) {  // binding still in progress: keep
        engine_remove(e, st.ns, st.name);
        it = e->pod_states.erase(it);
      } else {
        ++it;
      }
    }
  });
}

int32_t ksb_cache_skip_pod_update(ksb_engine* e, const uint8_t* buf, size_t len, int32_t* skip) {  // factory.go:306-353
  return cache_call(e, [&] {
    ksb::Pod p = decode_pod(buf, len);
    auto it = e->pod_states.find(p.key());
    bool s = false;
    if (it != e->pod_states.end() && it->second.assumed)
      s = blank_node_record(p) == blank_node_record(decode_pod((const uint8_t*)it->second.record.data(), it->second.record.size()));
    if (skip) *skip = s ? 1 : 0;
  });
}

int32_t ksb_schedule_extended(ksb_engine* e, const uint8_t* pod, size_t len, const uint32_t* keep_rows, const int32_t* row_scores,
                              ksb_result* out) {
  return scheduling(e, [&] {
    if (!out) throw std::runtime_error("out must not be null");
    e->impl->schedule_extended(decode_pod(pod, len), keep_rows, row_scores, out);
  });
}

int32_t ksb_preempt(ksb_engine* e, const uint8_t* pod, size_t len, ksb_preempt_result* out) {
  return scheduling(e, [&] {
    if (!out) throw std::runtime_error("out must not be null");
    ksb::PreemptOutcome r = e->impl->preempt(decode_pod(pod, len));
    out->status = r.status;
    out->node_index = r.row;
    out->num_victims = (int32_t)r.victims.size();
    out->num_potential = r.num_potential;
    out->num_candidates = r.num_candidates;
  });
}

int32_t ksb_preempt_victim(ksb_engine* e, int32_t i, const char** key, size_t* len) {
  if (!e || !e->impl || !key || !len) return -1;
  const auto& v = e->impl->last_preempt.victims;
  if (i < 0 || (size_t)i >= v.size()) return -1;
  *key = v[(size_t)i].data();
  *len = v[(size_t)i].size();
  return 0;
}

int32_t ksb_policy_set(ksb_engine* e, int32_t kind, const char* key, size_t key_len, int32_t presence, int32_t weight) {
  return scheduling(e, [&] { e->impl->policy_set(kind, std::string(key ? key : "", key ? key_len : 0), presence != 0, weight); });
}

int32_t ksb_deferred_errors(ksb_engine* e, char* buf, size_t buf_len) {
  if (!e || !e->impl) return -1;
  std::lock_guard<std::mutex> q(e->q_mu);
  const int32_t count = (int32_t)e->deferred_errors.size();
  if (buf && buf_len) {
    std::string all;
    for (auto& t : e->deferred_errors) {
      all += t;
      all.push_back('\n');
    }
    strncpy(buf, all.c_str(), buf_len - 1);
    buf[buf_len - 1] = 0;
  }
  e->deferred_errors.clear();
  return count;
}

int32_t ksb_flush(ksb_engine* e) {
  return scheduling(e, [&] { e->impl->flush(); });
}

uint64_t ksb_get_last_node_index(const ksb_engine* e) { return e && e->impl ? e->impl->last_node_index : 0; }
void ksb_set_last_node_index(ksb_engine* e, uint64_t v) {
  if (e && e->impl) e->impl->last_node_index = v;
}

int32_t ksb_get_stats(const ksb_engine* e, ksb_stats* out) {
  if (!e || !e->impl) return -1;
  *out = e->impl->stats;
  return 0;
}

int32_t ksb_replay_last_batch(ksb_engine* e, int32_t iters, double* avg_ms, ksb_result* out) {
  return scheduling(e, [&] { e->impl->replay_last_batch(iters, avg_ms, out); });
}

int32_t ksb_shard_export(ksb_engine* e, int32_t max_batch_pods, uint8_t* out, size_t cap, size_t* len) {
  return scheduling(e, [&] {
    std::string b = e->impl->shard_export(max_batch_pods);
    if (b.size() > cap) throw std::runtime_error("ksb_shard_export: output buffer too small");
    memcpy(out, b.data(), b.size());
    *len = b.size();
  });
}

int32_t ksb_shard_connect(ksb_engine* e, const uint8_t* blobs, const uint64_t* offsets, int32_t n_ranks) {
  return guarded(e, [&] {
    std::vector<std::string> v;
    for (int32_t i = 0; i < n_ranks; ++i)
      v.emplace_back((const char*)blobs + offsets[i], (size_t)(offsets[i + 1] - offsets[i]));
    e->impl->shard_connect(v);
  });
}

int32_t ksb_replay_last_batch_noalloc(ksb_engine* e, ksb_result* out) {
  return scheduling(e, [&] { e->impl->replay_last_batch_noalloc(out); });
}

}  // extern "C"
