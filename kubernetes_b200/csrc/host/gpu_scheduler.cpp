// Test / bench harness over the host mirror: the scheduleOne loop of the reference
// (scheduler.go:430-497) with in-memory listers and every bind succeeding, driven from JSON
// cluster files (the same files the CPU oracle consumes). Exposed to Python through a few
// extern "C" entry points; everything below them goes through include/ksched_b200.h only.
#include "gpu_scheduler.hpp"

#include <algorithm>
#include <array>
#include <chrono>
#include <cstdlib>
#include <cstring>
#include <memory>

#include "json.hpp"

using namespace ksb;

namespace {

struct Session {
  Cluster cluster;
  // pre-serialised host buffers (what the Go shim would hand to cgo)
  std::vector<std::string> node_bufs, existing_bufs;
  std::vector<std::pair<int, std::string>> owner_bufs;
  std::string pod_stream;
  std::vector<uint64_t> pod_offsets;
};

std::string esc(const std::string& s) {
  std::string o;
  json::escape_to(o, s);
  return o;
}

uint64_t fnv1a(uint64_t h, const void* p, size_t n) {
  const unsigned char* c = (const unsigned char*)p;
  for (size_t i = 0; i < n; ++i) {
    h ^= c[i];
    h *= 1099511628211ull;
  }
  return h;
}

void load_cluster(EngineHandle& h, const Session& s) {
  for (auto& b : s.node_bufs) h.check(ksb_node_upsert(h.get(), (const uint8_t*)b.data(), b.size()), "node_upsert");
  for (auto& kb : s.owner_bufs)
    h.check(ksb_selector_owner_upsert(h.get(), kb.first, (const uint8_t*)kb.second.data(), kb.second.size()),
            "selector_owner_upsert");
  for (auto& b : s.existing_bufs) h.check(ksb_pod_add(h.get(), (const uint8_t*)b.data(), b.size()), "pod_add");
  // the informers' initial sync ends here: the node table goes to the device now, as the reference's cache is
  // already filled when the first pod arrives
  h.check(ksb_flush(h.get()), "flush");
}

ksb_config config_from(const json::Value& opt) {
  ksb_config cfg;
  ksb_config_default(&cfg);
  cfg.device = (int32_t)opt.get_int("device", 0);
  cfg.label_words = (int32_t)opt.get_int("label_words", 4);
  cfg.hard_pod_affinity_weight = (int32_t)opt.get_int("hardPodAffinityWeight", 1);
  cfg.group_size = (int32_t)opt.get_int("group_size", 0);
  cfg.pipeline = (int32_t)opt.get_int("pipeline", 0);
  cfg.shard_rank = (int32_t)opt.get_int("shard_rank", 0);
  cfg.shard_count = (int32_t)opt.get_int("shard_count", 1);
  if (const json::Value* w = opt.find("weights"); w && !w->is_null()) {
    cfg.w_least_requested = (int32_t)w->get_int("LeastRequestedPriority", 1);
    cfg.w_balanced = (int32_t)w->get_int("BalancedResourceAllocation", 1);
    cfg.w_selector_spread = (int32_t)w->get_int("SelectorSpreadPriority", 1);
    cfg.w_node_affinity = (int32_t)w->get_int("NodeAffinityPriority", 1);
    cfg.w_taint_toleration = (int32_t)w->get_int("TaintTolerationPriority", 1);
    cfg.w_inter_pod_affinity = (int32_t)w->get_int("InterPodAffinityPriority", 1);
    cfg.w_prefer_avoid = (int32_t)w->get_int("NodePreferAvoidPodsPriority", 10000);
  }
  return cfg;
}

std::string run_on(EngineHandle& h, Session& s, const json::Value& opt, double load_ms);

// Policy-enabled plugins from the options: {"weights": {"MostRequestedPriority": w, "ResourceLimitsPriority": w, "ImageLocalityPriority": w},
// "policy": {"nodeLabelPriorities": [{"label", "presence", "weight"}], "labelsPresence": {"labels": [...], "presence": b}}}
void apply_policy(EngineHandle& h, const json::Value& opt) {
  if (const json::Value* w = opt.find("weights"); w && !w->is_null()) {
    if (w->get_int("MostRequestedPriority", 0) != 0)
      h.check(ksb_policy_set(h.get(), KSB_POLICY_MOST_REQUESTED, nullptr, 0, 0, (int32_t)w->get_int("MostRequestedPriority", 0)), "policy");
    if (w->get_int("ResourceLimitsPriority", 0) != 0)
      h.check(ksb_policy_set(h.get(), KSB_POLICY_RESOURCE_LIMITS, nullptr, 0, 0, (int32_t)w->get_int("ResourceLimitsPriority", 0)), "policy");
    if (w->get_int("ImageLocalityPriority", 0) != 0)
      h.check(ksb_policy_set(h.get(), KSB_POLICY_IMAGE_LOCALITY, nullptr, 0, 0, (int32_t)w->get_int("ImageLocalityPriority", 0)), "policy");
    if (w->get_int("ServiceSpreadingPriority", 0) != 0)
      h.check(ksb_policy_set(h.get(), KSB_POLICY_SERVICE_SPREADING, nullptr, 0, 0, (int32_t)w->get_int("ServiceSpreadingPriority", 0)), "policy");
    if (w->get_int("EqualPriority", 0) != 0)
      h.check(ksb_policy_set(h.get(), KSB_POLICY_EQUAL, nullptr, 0, 0, (int32_t)w->get_int("EqualPriority", 0)), "policy");
  }
  if (const json::Value* pv = opt.find("policy"); pv && !pv->is_null()) {
    if (pv->has("nodeLabelPriorities"))
      for (auto& a : pv->get_arr("nodeLabelPriorities")) {
        const std::string label = a.get_str("label");
        h.check(ksb_policy_set(h.get(), KSB_POLICY_NODE_LABEL_PRIORITY, label.data(), label.size(), a.get_bool("presence", true) ? 1 : 0,
                               (int32_t)a.get_int("weight", 1)), "policy");
      }
    if (const json::Value* lp = pv->find("labelsPresence"); lp && !lp->is_null())
      for (auto& l : lp->get_arr("labels"))
        h.check(ksb_policy_set(h.get(), KSB_POLICY_LABEL_PRESENCE, l.s.data(), l.s.size(), lp->get_bool("presence", true) ? 1 : 0, 0), "policy");
  }
}

std::string run(Session& s, const json::Value& opt) {
  ksb_config cfg = config_from(opt);
  EngineHandle h(cfg);
  apply_policy(h, opt);
  ksb_set_last_node_index(h.get(), (uint64_t)opt.get_int("lastNodeIndex", 0));
  auto tl0 = std::chrono::steady_clock::now();
  load_cluster(h, s);
  auto tl1 = std::chrono::steady_clock::now();
  return run_on(h, s, opt, std::chrono::duration<double, std::milli>(tl1 - tl0).count());
}

std::string run_on(EngineHandle& h, Session& s, const json::Value& opt, double load_ms) {
  const int64_t batch = opt.get_int("batch", -1);  // -1: whole stream in one call; 0: Schedule + AssumePod per pod
  int64_t limit = opt.get_int("limit", -1);
  const bool want_results = opt.get_bool("results", true);
  const int replay_iters = (int)opt.get_int("replay_iters", 0);
  size_t np = s.cluster.pods.size();
  if (limit >= 0 && (size_t)limit < np) np = (size_t)limit;

  std::vector<ksb_result> results(np);
  std::vector<std::string> hosts(np), messages(np);
  std::vector<ExtendedResourceBinding> bindings(np);
  std::vector<std::string> failed_json(np);  // "failed_predicates": per pod, {node: [reasons]} of the nodes that did not fit
  const bool want_failed = opt.get_bool("failed_predicates", false);
  const bool want_scores = opt.get_bool("scores", false);  // "scores": per pod, {node: total} over the filtered list (batch 0 only)
  std::vector<std::string> scores_json(np);
  GpuScheduler sched(h);
  // one-pod mode drives the exported cache state machine (AssumePod after every Schedule); batch mode leaves it off, so
  // that ksb_schedule_batch does not record every pod it places
  std::unique_ptr<CacheDecorator> cache_holder;
  if (batch == 0) cache_holder = std::make_unique<CacheDecorator>(h);
  double kernel_ms = 0, encode_ms = 0, replay_ms = 0, decode_ms = 0, post_ms = 0, sync_ms = 0;
  std::vector<double> lat_us;
  std::vector<double> call_us;  // batch mode: duration of each ksb_schedule_batch call
  std::vector<std::array<double, 3>> lat_parts;  // encode, table sync, kernel (ms) of each one-pod call
  ksb_stats st0;  // bytes copied before the timed region (cluster load, ksb_flush) are not this run's
  ksb_get_stats(h.get(), &st0);
  auto t0 = std::chrono::steady_clock::now();
  if (batch == 0) {
    // the reference's own call sequence: Schedule, then AssumePod with NodeName and Assigned set
    lat_us.reserve(np);
    for (size_t i = 0; i < np; ++i) {
      const Pod& pod = s.cluster.pods[i];
      auto a0 = std::chrono::steady_clock::now();
      ksb_result r;
      h.check(ksb_schedule(h.get(), (const uint8_t*)s.pod_stream.data() + s.pod_offsets[i],
                           (size_t)(s.pod_offsets[i + 1] - s.pod_offsets[i]), &r),
              "Schedule");
      auto a1 = std::chrono::steady_clock::now();
      lat_us.push_back(std::chrono::duration<double, std::micro>(a1 - a0).count());
      {
        ksb_stats st1;
        ksb_get_stats(h.get(), &st1);
        lat_parts.push_back({st1.last_batch_encode_ms, st1.last_batch_sync_ms, st1.last_batch_kernel_ms});
      }
      results[i] = r;
      if (want_failed && r.num_nodes > 0) {
        // FailedPredicateMap (core/generic_scheduler.go:283-362) of this call, node name -> reason texts
        std::vector<uint32_t> masks((size_t)r.num_nodes);
        h.check(ksb_failed_predicates(h.get(), masks.data(), r.num_nodes), "failed_predicates");
        std::string fj = "{";
        bool first_node = true;
        for (int32_t row = 0; row < r.num_nodes; ++row) {
          if (!masks[row]) continue;
          const char* nm = nullptr;
          size_t nl = 0;
          h.check(ksb_node_name(h.get(), row, &nm, &nl), "node_name");
          fj += std::string(first_node ? "" : ",") + esc(std::string(nm, nl)) + ":[";
          first_node = false;
          bool first_reason = true;
          for (int bit = 0; bit < 32; ++bit) {
            if (!((masks[row] >> bit) & 1u)) continue;
            char buf[160];
            h.check(ksb_reason_name(h.get(), bit, buf, sizeof buf), "reason_name");
            fj += std::string(first_reason ? "" : ",") + esc(buf);
            first_reason = false;
          }
          fj += "]";
        }
        failed_json[i] = fj + "}";
      }
      if (want_scores && r.status == KSB_SCHEDULED && r.num_filtered >= 2) {
        std::vector<int64_t> sc((size_t)r.num_nodes);
        h.check(ksb_scores(h.get(), sc.data(), r.num_nodes), "scores");
        std::string sj = "{";
        bool first_node = true;
        for (int32_t row = 0; row < r.num_nodes; ++row) {
          if (sc[row] == INT64_MIN) continue;
          const char* nm = nullptr;
          size_t nl = 0;
          h.check(ksb_node_name(h.get(), row, &nm, &nl), "node_name");
          sj += std::string(first_node ? "" : ",") + esc(std::string(nm, nl)) + ":" + std::to_string(sc[row]);
          first_node = false;
        }
        scores_json[i] = sj + "}";
      }
      try {
        ScheduleOutcome o = sched.outcome(pod, r, 0);
        hosts[i] = o.host;
        bindings[i] = o.binding;
        Pod assumed = pod;  // scheduler.go:478
        assumed.node_name = o.host;
        for (auto& e : assumed.extended) {
          auto it = o.binding.find(e.name);
          if (it != o.binding.end()) e.assigned = it->second;
        }
        const std::string cache_err = cache_holder->AssumePod(assumed);
        if (!cache_err.empty()) throw std::runtime_error(cache_err);
      } catch (const FitError& fe) {
        messages[i] = fe.what();
      } catch (const std::exception& ex) {
        messages[i] = ex.what();
      }
    }
  } else {
    size_t step = batch < 0 ? np : (size_t)batch;
    for (size_t b0 = 0; b0 < np; b0 += step) {
      size_t b1 = std::min(np, b0 + step);
      auto c0 = std::chrono::steady_clock::now();
      h.check(ksb_schedule_batch(h.get(), (const uint8_t*)s.pod_stream.data(), s.pod_offsets.data() + b0, (int32_t)(b1 - b0),
                                 results.data() + b0),
              "ScheduleBatch");
      call_us.push_back(std::chrono::duration<double, std::micro>(std::chrono::steady_clock::now() - c0).count());
      ksb_stats st;
      ksb_get_stats(h.get(), &st);
      kernel_ms += st.last_batch_kernel_ms;
      encode_ms += st.last_batch_encode_ms;
      decode_ms += st.last_batch_decode_ms;
      post_ms += st.last_batch_post_ms;
      sync_ms += st.last_batch_sync_ms;
      if (want_results || b1 == np) {
        for (size_t i = b0; i < b1; ++i) {
          if (!want_results && results[i].status != KSB_SCHEDULED) continue;
          if (!want_results) continue;
          try {
            ScheduleOutcome o = sched.outcome(s.cluster.pods[i], results[i], (int)(i - b0));
            hosts[i] = o.host;
            bindings[i] = o.binding;
          } catch (const std::exception& ex) {
            messages[i] = ex.what();
          }
        }
      }
    }
  }
  auto t1 = std::chrono::steady_clock::now();
  ksb_stats st_e2e;
  ksb_get_stats(h.get(), &st_e2e);
  if (replay_iters > 0 && batch < 0) {
    std::vector<ksb_result> again(np);
    h.check(ksb_replay_last_batch(h.get(), replay_iters, &replay_ms, again.data()), "replay");
    for (size_t i = 0; i < np; ++i)
      if (again[i].status != results[i].status || again[i].node_index != results[i].node_index ||
          again[i].num_filtered != results[i].num_filtered || again[i].score != results[i].score ||
          memcmp(again[i].device_mask, results[i].device_mask, sizeof again[i].device_mask) != 0)
        throw std::runtime_error("replay diverged from the first run at pod " + std::to_string(i));
  }
  ksb_stats st;
  ksb_get_stats(h.get(), &st);

  uint64_t hash = 1469598103934665603ull, rhash = 1469598103934665603ull;
  size_t scheduled = 0, fit_errors = 0, errors = 0;
  for (size_t i = 0; i < np; ++i) {
    int32_t v[2] = {results[i].status, results[i].node_index};
    hash = fnv1a(hash, v, sizeof v);
    hash = fnv1a(hash, results[i].device_mask, sizeof results[i].device_mask);
    // result_hash: everything Schedule returns per pod (the placement plus filtered count, total score and, below,
    // the FitError text)
    rhash = fnv1a(rhash, v, sizeof v);
    rhash = fnv1a(rhash, results[i].device_mask, sizeof results[i].device_mask);
    int64_t w[2] = {(int64_t)results[i].num_filtered, results[i].score};
    rhash = fnv1a(rhash, w, sizeof w);
    if (want_results && !messages[i].empty()) rhash = fnv1a(rhash, messages[i].data(), messages[i].size());
    if (results[i].status == KSB_SCHEDULED) scheduled++;
    else if (results[i].status == KSB_FIT_ERROR) fit_errors++;
    else errors++;
  }
  std::string out = "{";
  out += "\"pods\":" + std::to_string(np);
  out += ",\"nodes\":" + std::to_string(s.cluster.nodes.size());
  out += ",\"scheduled\":" + std::to_string(scheduled);
  out += ",\"fit_errors\":" + std::to_string(fit_errors);
  out += ",\"errors\":" + std::to_string(errors);
  out += ",\"placement_hash\":\"" + std::to_string(hash) + "\"";
  out += ",\"result_hash\":\"" + std::to_string(rhash) + "\"";
  out += ",\"load_ms\":" + std::to_string(load_ms);
  out += ",\"e2e_ms\":" + std::to_string(std::chrono::duration<double, std::milli>(t1 - t0).count());
  out += ",\"kernel_ms\":" + std::to_string(kernel_ms);
  out += ",\"encode_ms\":" + std::to_string(encode_ms);
  out += ",\"replay_ms\":" + std::to_string(replay_ms);
  out += ",\"decode_ms\":" + std::to_string(decode_ms) + ",\"post_ms\":" + std::to_string(post_ms) + ",\"sync_ms\":" + std::to_string(sync_ms);
  out += ",\"kernel_launches\":" + std::to_string(st.kernel_launches);
  out += ",\"stage_cycles\":[";
  for (int i = 0; i < 13; ++i) out += (i ? "," : "") + std::to_string(st.stage_cycles[i]);
  out += "]";
  out += ",\"groups\":" + std::to_string(st.groups);
  out += ",\"group_restarts\":" + std::to_string(st.group_restarts);
  out += ",\"h2d_bytes\":" + std::to_string(st_e2e.h2d_bytes - st0.h2d_bytes);
  out += ",\"d2h_bytes\":" + std::to_string(st_e2e.d2h_bytes - st0.d2h_bytes);
  out += ",\"lastNodeIndex\":" + std::to_string(ksb_get_last_node_index(h.get()));
  if (!lat_us.empty()) {
    // "lat_skip": leave the first calls (cold dictionaries: first use of a label, selector class, topology key)
    // out of the percentiles; they stay in `slowest`
    const size_t skip = std::min<size_t>((size_t)std::max<int64_t>(0, opt.get_int("lat_skip", 0)), lat_us.size() - 1);
    std::vector<double> sorted(lat_us.begin() + skip, lat_us.end());
    std::sort(sorted.begin(), sorted.end());
    out += ",\"p50_us\":" + std::to_string(sorted[sorted.size() / 2]);
    out += ",\"p99_us\":" + std::to_string(sorted[std::min(sorted.size() - 1, (size_t)(sorted.size() * 0.99))]);
    out += ",\"max_us\":" + std::to_string(sorted.back());
    // the slowest calls and where in the stream they were: first uses of a label bit, selector class or scratch size
    std::vector<size_t> order(lat_us.size());
    for (size_t i = 0; i < order.size(); ++i) order[i] = i;
    const size_t top = std::min<size_t>(8, order.size());
    std::partial_sort(order.begin(), order.begin() + top, order.end(), [&](size_t a, size_t b) { return lat_us[a] > lat_us[b]; });
    out += ",\"slowest\":[";
    for (size_t i = 0; i < top; ++i)
      out += std::string(i ? "," : "") + "[" + std::to_string(order[i]) + "," + std::to_string(lat_us[order[i]]) + "," +
             std::to_string(lat_parts[order[i]][0]) + "," + std::to_string(lat_parts[order[i]][1]) + "," +
             std::to_string(lat_parts[order[i]][2]) + "]";
    out += "]";
  }
  if (call_us.size() > 1) {
    // batch mode: every pod of a call completes when the call returns, so a call's duration is its pods' latency
    std::vector<double> sorted = call_us;
    std::sort(sorted.begin(), sorted.end());
    out += ",\"batch_call_p50_us\":" + std::to_string(sorted[sorted.size() / 2]);
    out += ",\"batch_call_p99_us\":" + std::to_string(sorted[std::min(sorted.size() - 1, (size_t)(sorted.size() * 0.99))]);
    out += ",\"batch_calls\":" + std::to_string(sorted.size());
  }
  if (want_results) {
    out += ",\"results\":[";
    for (size_t i = 0; i < np; ++i) {
      if (i) out += ",";
      out += "{\"host\":" + esc(hosts[i]) + ",\"status\":" + std::to_string(results[i].status) +
             ",\"filtered\":" + std::to_string(results[i].num_filtered) + ",\"score\":" + std::to_string(results[i].score) +
             ",\"message\":" + esc(messages[i]) + (failed_json[i].empty() ? std::string() : ",\"failed\":" + failed_json[i]) +
             (scores_json[i].empty() ? std::string() : ",\"scores\":" + scores_json[i]) +
             ",\"devices\":{";
      bool first = true;
      for (auto& kv : bindings[i]) {
        if (!first) out += ",";
        first = false;
        out += esc(kv.first) + ":[";
        for (size_t k = 0; k < kv.second.size(); ++k) {
          if (k) out += ",";
          out += esc(kv.second[k]);
        }
        out += "]";
      }
      out += "}}";
    }
    out += "]";
  }
  out += "}";
  return out;
}

char* dup(const std::string& s) {
  char* p = (char*)malloc(s.size() + 1);
  memcpy(p, s.c_str(), s.size() + 1);
  return p;
}

}  // namespace

extern "C" {

// Parse a cluster (inline JSON text or a file path) and pre-serialise it into wire records.
void* ksbh_session_create(const char* cluster_json_or_path, int is_path, char** err) {
  try {
    auto s = std::make_unique<Session>();
    json::Value v = is_path ? json::parse_file(cluster_json_or_path) : json::parse(cluster_json_or_path);
    s->cluster = cluster_from_json(v);
    for (auto& n : s->cluster.nodes) {
      wire::Writer w;
      wire::put(w, n);
      s->node_bufs.push_back(std::move(w.buf));
    }
    const std::vector<SelectorOwner>* lists[4] = {&s->cluster.services, &s->cluster.rcs, &s->cluster.rss, &s->cluster.sss};
    for (int k = 0; k < 4; ++k) {
      int idx = 0;
      for (auto o : *lists[k]) {
        if (o.name.empty()) o.name = "owner-" + std::to_string(k) + "-" + std::to_string(idx);
        ++idx;
        wire::Writer w;
        wire::put(w, o);
        s->owner_bufs.push_back({k, std::move(w.buf)});
      }
    }
    for (auto& p : s->cluster.existing_pods) {
      wire::Writer w;
      wire::put(w, p);
      s->existing_bufs.push_back(std::move(w.buf));
    }
    wire::Writer w;
    s->pod_offsets.push_back(0);
    for (auto& p : s->cluster.pods) {
      wire::put(w, p);
      s->pod_offsets.push_back(w.buf.size());
    }
    s->pod_stream = std::move(w.buf);
    return s.release();
  } catch (const std::exception& ex) {
    if (err) *err = dup(ex.what());
    return nullptr;
  }
}

void ksbh_session_destroy(void* s) { delete (Session*)s; }

// The wire record the session built for one object (kind 0 node, 1 pod of the stream, 2 existing pod, 3 selector owner),
// so that a test can hold the C++ encoder against the Python one (kubernetes_b200/wire.py) byte for byte. Returns 0, or 1
// when there is no such record.
int ksbh_session_record(void* session, int kind, int index, const uint8_t** data, size_t* len) {
  Session* s = (Session*)session;
  if (!s || index < 0) return 1;
  const std::string* b = nullptr;
  if (kind == 0 && (size_t)index < s->node_bufs.size()) b = &s->node_bufs[index];
  else if (kind == 2 && (size_t)index < s->existing_bufs.size()) b = &s->existing_bufs[index];
  else if (kind == 3 && (size_t)index < s->owner_bufs.size()) b = &s->owner_bufs[index].second;
  else if (kind == 1 && (size_t)index + 1 < s->pod_offsets.size()) {
    *data = (const uint8_t*)s->pod_stream.data() + s->pod_offsets[index];
    *len = (size_t)(s->pod_offsets[index + 1] - s->pod_offsets[index]);
    return 0;
  }
  if (!b) return 1;
  *data = (const uint8_t*)b->data();
  *len = b->size();
  return 0;
}

// Run the pod stream on a fresh engine. Returns JSON (free with ksbh_free); on failure {"error": ...}.
char* ksbh_session_run(void* session, const char* options_json) {
  try {
    json::Value opt = json::parse(options_json && *options_json ? options_json : "{}");
    return dup(run(*(Session*)session, opt));
  } catch (const std::exception& ex) {
    return dup(std::string("{\"error\":") + esc(ex.what()) + "}");
  }
}

void ksbh_free(char* p) { free(p); }

// The cache state machine driven by a list of operations on a fresh engine; same request and reply as the oracle's
// "cache_ops" (oracle/oracle_capi.cpp): {"ttl", "ops": [{"do", "pod", "old", "node", "name", "now"}], "report": [...]}.
char* ksbh_cache_ops(const char* request_json) {
  try {
    json::Value req = json::parse(request_json);
    ksb_config cfg = config_from(req);
    EngineHandle h(cfg);
    CacheDecorator c(h, req.get_num("ttl", 30.0));
    std::string errors = "[";
    bool first = true;
    for (auto& op : req.get_arr("ops")) {
      const std::string what = op.get_str("do");
      std::string err;
      if (what == "assume") err = c.AssumePod(pod_from_json(*op.find("pod")));
      else if (what == "finish") err = c.FinishBinding(pod_from_json(*op.find("pod")), op.get_num("now", 0.0));
      else if (what == "forget") err = c.ForgetPod(pod_from_json(*op.find("pod")));
      else if (what == "add") err = c.AddPod(pod_from_json(*op.find("pod")));
      else if (what == "update") err = c.UpdatePod(pod_from_json(*op.find("old")), pod_from_json(*op.find("pod")));
      else if (what == "remove") err = c.RemovePod(pod_from_json(*op.find("pod")));
      else if (what == "cleanup") c.CleanupAssumedPods(op.get_num("now", 0.0));
      else if (what == "is_assumed") err = c.IsAssumedPod(pod_from_json(*op.find("pod"))) ? "assumed" : "not assumed";
      else if (what == "add_node" || what == "update_node") c.AddNode(node_from_json(*op.find("node")));
      else if (what == "remove_node") {
        Node n;
        n.name = op.get_str("name");
        c.RemoveNode(n);
      } else err = "unknown op " + what;
      errors += std::string(first ? "" : ",") + esc(err);
      first = false;
    }
    errors += "]";
    std::string out = "{\"errors\":" + errors + ",\"nodes\":{";
    first = true;
    for (auto& nv : req.get_arr("report")) {
      const std::string name = nv.s;
      std::vector<char> buf(1 << 16);
      size_t len = 0;
      int32_t rc = ksb_node_info(h.get(), name.data(), name.size(), buf.data(), buf.size(), &len);
      if (rc == 1) {
        buf.resize(len + 1);
        rc = ksb_node_info(h.get(), name.data(), name.size(), buf.data(), buf.size(), &len);
      }
      h.check(rc, "node_info");
      out += std::string(first ? "" : ",") + esc(name) + ":" + std::string(buf.data(), len);
      first = false;
    }
    out += "}}";
    return dup(out);
  } catch (const std::exception& ex) {
    return dup(std::string("{\"error\":") + esc(ex.what()) + "}");
  }
}

// ---- a loaded engine kept open across calls (sharded runs: load, exchange IPC handles, then run) ----

struct OpenEngine {
  Session* s;
  std::unique_ptr<EngineHandle> h;
  double load_ms;
};

void* ksbh_engine_open(void* session, const char* options_json, char** err) {
  try {
    json::Value opt = json::parse(options_json && *options_json ? options_json : "{}");
    auto oe = std::make_unique<OpenEngine>();
    oe->s = (Session*)session;
    ksb_config cfg = config_from(opt);
    oe->h = std::make_unique<EngineHandle>(cfg);
    apply_policy(*oe->h, opt);
    auto t0 = std::chrono::steady_clock::now();
    load_cluster(*oe->h, *oe->s);
    oe->load_ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
    return oe.release();
  } catch (const std::exception& ex) {
    if (err) *err = dup(ex.what());
    return nullptr;
  }
}

int ksbh_engine_shard_export(void* oe, int max_batch_pods, uint8_t* out, size_t cap, size_t* len, char** err) {
  OpenEngine* e = (OpenEngine*)oe;
  if (ksb_shard_export(e->h->get(), max_batch_pods, out, cap, len) != 0) {
    if (err) *err = dup(ksb_last_error(e->h->get()));
    return -1;
  }
  return 0;
}

int ksbh_engine_shard_connect(void* oe, const uint8_t* blobs, const uint64_t* offsets, int n_ranks, char** err) {
  OpenEngine* e = (OpenEngine*)oe;
  if (ksb_shard_connect(e->h->get(), blobs, offsets, n_ranks) != 0) {
    if (err) *err = dup(ksb_last_error(e->h->get()));
    return -1;
  }
  return 0;
}

char* ksbh_engine_run(void* oe, const char* options_json) {
  OpenEngine* e = (OpenEngine*)oe;
  try {
    json::Value opt = json::parse(options_json && *options_json ? options_json : "{}");
    return dup(run_on(*e->h, *e->s, opt, e->load_ms));
  } catch (const std::exception& ex) {
    return dup(std::string("{\"error\":") + esc(ex.what()) + "}");
  }
}

void ksbh_engine_close(void* oe) { delete (OpenEngine*)oe; }

}  // extern "C"
