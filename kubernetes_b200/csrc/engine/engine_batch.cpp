// Engine, part 3: the batch driver (ksb_schedule / ksb_schedule_batch), replay for measurement,
// FitError text, and the pod mutators that keep device and host mirror in step.
#include <algorithm>
#include <chrono>
#include <cstring>
#include <stdexcept>

#include "engine.hpp"

namespace ksb {

#define CUDA_OK(call)                                                                          \
  do {                                                                                         \
    cudaError_t e_ = (call);                                                                   \
    if (e_ != cudaSuccess)                                                                     \
      throw std::runtime_error(std::string("CUDA error: ") + cudaGetErrorString(e_) + " at " + \
                               __FILE__ + ":" + std::to_string(__LINE__));                     \
  } while (0)

namespace {
struct ScratchLayout {
  size_t norms, zone_sum, raw, score, reasons, seg_max, seg_cnt, barrier, error_flag, fiterr_count, last_index,
      group_done, counters, ghist, fiterr_hist, total;
};
ScratchLayout scratch_layout(int n_pad, int zone_stride, int nseg, int G) {
  ScratchLayout L;
  size_t off = 0;
  auto take = [&](size_t bytes) {
    size_t o = off;
    off += (bytes + 255) & ~(size_t)255;
    return o;
  };
  L.norms = take(2 * (size_t)KSB_MAX_GROUP * sizeof(KsbNorms));
  L.zone_sum = take(2 * (size_t)KSB_MAX_GROUP * zone_stride * 4);
  L.raw = take((size_t)G * n_pad * 16);
  L.score = take((size_t)G * n_pad * 4);
  L.reasons = take((size_t)G * n_pad * 4);
  L.seg_max = take((size_t)G * nseg * 4);
  L.seg_cnt = take((size_t)G * nseg * 4);
  L.barrier = take(4);
  L.error_flag = take(4);
  L.fiterr_count = take(4);
  L.last_index = take(8);
  L.group_done = take(8);
  L.counters = take(128);
  L.ghist = take(2 * (size_t)KSB_MAX_GROUP * KSB_REASON_BITS * 4);
  L.fiterr_hist = take((size_t)KSB_MAX_FITERR * KSB_REASON_BITS * 4);
  L.total = off;
  return L;
}
}  // namespace

KsbBatchArgs Engine::make_args(int n_pods, bool commit) {
  zone_stride = (dT.n_zones + 63) & ~63;
  // group size: as configured, bounded by what the zone-sum arrays may take of shared memory
  group = cfg.group_size > 0 ? std::min(cfg.group_size, KSB_MAX_GROUP) : KSB_MAX_GROUP;
  group = std::max(1, std::min(group, 16384 / zone_stride));
  nseg = (n + 255) / 256;
  if (nseg > KSB_MAX_SEGMENTS) throw std::runtime_error("more than 524,288 node rows in one shard");
  group = std::max(1, std::min(group, KSB_SEG_CACHE / nseg));
  if (zone_stride * 4 * KSB_MAX_GROUP > 160 * 1024) throw std::runtime_error("more than 1,280 distinct zones");
  smem_bytes = ksb_batch_smem_bytes(zone_stride);
  int mg = 0, th = 0;
  int rc = ksb_batch_kernel_limits(cfg.device, smem_bytes, &mg, &th);
  if (rc != 0) throw std::runtime_error(std::string("kernel attribute query failed: ") + cudaGetErrorString((cudaError_t)rc));
  max_grid = mg;
  threads = th;
  if (max_grid < 1) throw std::runtime_error("batch kernel does not fit on this device");
  // enough blocks to cover every segment; when segments are few, extra blocks split the group's pods
  grid = nseg >= max_grid ? max_grid : std::max(1, std::min(max_grid, nseg * group));
  ScratchLayout L = scratch_layout(n_pad, zone_stride, nseg, group);
  b_scratch.ensure(L.total);
  char* base = (char*)b_scratch.d;
  KsbBatchArgs a;
  memset(&a, 0, sizeof a);
  a.table = dT;
  a.headers = (const KsbPodHeader*)b_headers.d;
  a.pool = (const uint64_t*)b_pool.d;
  a.dep = (const int32_t*)b_dep.d;
  a.n_pods = n_pods;
  a.zone_stride = zone_stride;
  a.commit = commit ? 1 : 0;
  a.group_size = group;
  a.nseg = nseg;
  a.results = (KsbResult*)b_results.d;
  a.norms = (KsbNorms*)(base + L.norms);
  a.zone_sum = (int32_t*)(base + L.zone_sum);
  a.raw = (int4*)(base + L.raw);
  a.score = (int32_t*)(base + L.score);
  a.reasons = (uint32_t*)(base + L.reasons);
  a.seg_max = (int32_t*)(base + L.seg_max);
  a.seg_cnt = (int32_t*)(base + L.seg_cnt);
  a.barrier = (uint32_t*)(base + L.barrier);
  a.error_flag = (int32_t*)(base + L.error_flag);
  a.fiterr_count = (int32_t*)(base + L.fiterr_count);
  a.fiterr_hist = (int32_t*)(base + L.fiterr_hist);
  a.last_node_index = (uint64_t*)(base + L.last_index);
  a.group_done = (int32_t*)(base + L.group_done);
  a.counters = (uint64_t*)(base + L.counters);
  a.ghist = (int32_t*)(base + L.ghist);
  return a;
}

// Zero what must start at zero, seed the round-robin counter, launch. Works on whatever scratch the
// arguments point at (the engine's own buffer or the static no-allocation array).
void Engine::launch(const KsbBatchArgs& a) {
  static KsbNorms init[2 * KSB_MAX_GROUP];
  memset(init, 0, sizeof init);
  for (auto& x : init) x.one_row = -1;
  CUDA_OK(cudaMemcpyAsync(a.norms, init, sizeof init, cudaMemcpyHostToDevice, stream));
  CUDA_OK(cudaMemsetAsync(a.zone_sum, 0, 2 * (size_t)KSB_MAX_GROUP * a.zone_stride * 4, stream));
  CUDA_OK(cudaMemsetAsync(a.barrier, 0, 4, stream));
  CUDA_OK(cudaMemsetAsync(a.error_flag, 0, 4, stream));
  CUDA_OK(cudaMemsetAsync(a.fiterr_count, 0, 4, stream));
  CUDA_OK(cudaMemsetAsync(a.group_done, 0, 8, stream));
  CUDA_OK(cudaMemsetAsync(a.counters, 0, 128, stream));
  CUDA_OK(cudaMemsetAsync(a.ghist, 0, 2 * (size_t)KSB_MAX_GROUP * KSB_REASON_BITS * 4, stream));
  CUDA_OK(cudaMemcpyAsync(a.last_node_index, &last_node_index, 8, cudaMemcpyHostToDevice, stream));
  CUDA_OK(cudaStreamSynchronize(stream));  // last_node_index may change right after this call
  int rc = ksb_launch_batch(&a, grid, smem_bytes, stream);
  if (rc != 0) throw std::runtime_error(std::string("kernel launch failed: ") + cudaGetErrorString((cudaError_t)rc));
  stats.kernel_launches += 1;
}

void Engine::snapshot_dynamic() {
  // everything the kernel mutates: int64 family, u32 family, u64 family (ports), spread, tables
  size_t s0 = h_i64.size() * 8, s1 = h_u32.size() * 4, s2 = h_u64.size() * 8;
  size_t s3 = h_spread.size() * (size_t)n_pad * 4, s4 = (size_t)std::max<uint32_t>(arena_words, 1) * 4;
  b_snapshot.ensure(s0 + s1 + s2 + s3 + s4);
  char* d = (char*)b_snapshot.d;
  CUDA_OK(cudaMemcpyAsync(d, b_i64.d, s0, cudaMemcpyDeviceToDevice, stream));
  CUDA_OK(cudaMemcpyAsync(d + s0, b_u32.d, s1, cudaMemcpyDeviceToDevice, stream));
  CUDA_OK(cudaMemcpyAsync(d + s0 + s1, b_u64.d, s2, cudaMemcpyDeviceToDevice, stream));
  if (s3) CUDA_OK(cudaMemcpyAsync(d + s0 + s1 + s2, b_spread.d, s3, cudaMemcpyDeviceToDevice, stream));
  CUDA_OK(cudaMemcpyAsync(d + s0 + s1 + s2 + s3, b_tables.d, s4, cudaMemcpyDeviceToDevice, stream));
  have_snapshot = true;
}

void Engine::restore_dynamic() {
  size_t s0 = h_i64.size() * 8, s1 = h_u32.size() * 4, s2 = h_u64.size() * 8;
  size_t s3 = h_spread.size() * (size_t)n_pad * 4, s4 = (size_t)std::max<uint32_t>(arena_words, 1) * 4;
  char* d = (char*)b_snapshot.d;
  CUDA_OK(cudaMemcpyAsync(b_i64.d, d, s0, cudaMemcpyDeviceToDevice, stream));
  CUDA_OK(cudaMemcpyAsync(b_u32.d, d + s0, s1, cudaMemcpyDeviceToDevice, stream));
  CUDA_OK(cudaMemcpyAsync(b_u64.d, d + s0 + s1, s2, cudaMemcpyDeviceToDevice, stream));
  if (s3) CUDA_OK(cudaMemcpyAsync(b_spread.d, d + s0 + s1 + s2, s3, cudaMemcpyDeviceToDevice, stream));
  CUDA_OK(cudaMemcpyAsync(b_tables.d, d + s0 + s1 + s2 + s3, s4, cudaMemcpyDeviceToDevice, stream));
}

void Engine::schedule_batch(std::vector<Pod>& batch, ksb_result* out, bool assume) {
  std::lock_guard<std::mutex> g(mu);
  CUDA_OK(cudaSetDevice(cfg.device));
  const int np = (int)batch.size();
  auto t0 = std::chrono::steady_clock::now();
  if (layout_dirty) rebuild_layout();
  last_n_nodes = n;
  last_encoded.assign(np, EncodedPod());
  last_results.assign(np, KsbResult());
  last_fiterr_hist.clear();
  have_snapshot = false;
  for (int i = 0; i < np; ++i) {
    out[i].status = KSB_SCHEDULE_ERROR;
    out[i].node_index = -1;
    out[i].num_filtered = 0;
    out[i].num_nodes = n;
    out[i].score = 0;
    for (int j = 0; j < KSB_MAX_EXT_ENTRIES; ++j) out[i].device_mask[j] = 0;
  }
  if (np == 0) return;
  if (n == 0) {  // ErrNoNodesAvailable (generic_scheduler.go:116-118)
    for (int i = 0; i < np; ++i) out[i].status = KSB_NO_NODES;
    return;
  }
  // pass 1: classes; pass 2: requirement blocks
  for (auto& p : batch) register_pod_classes(p);
  layout_arena();
  std::vector<KsbPodHeader> headers(np);
  std::vector<uint64_t> pool;
  for (int i = 0; i < np; ++i) {
    if (batch[i].num_volumes > 0)  // volume predicates are out of scope: refuse loudly, never guess
      throw std::runtime_error("pod " + batch[i].key() + " has volumes: not supported by the GPU engine");
    encode_pod(batch[i], &last_encoded[i]);
    last_encoded[i].h.pool_off = (uint32_t)pool.size();
    headers[i] = last_encoded[i].h;
    pool.insert(pool.end(), last_encoded[i].pool.begin(), last_encoded[i].pool.end());
  }
  if (pool.empty()) pool.push_back(0);
  // dep[j]: the latest earlier pod whose assume step bumps a counter pod j reads (its SelectorSpread column
  // or any inter-pod-affinity table). A group never holds both, so those inputs are constant inside a group.
  std::vector<int32_t> dep(np, -1);
  {
    std::vector<int32_t> last_spread(h_spread.size(), -1), last_table(tables.size(), -1);
    std::unordered_map<uint32_t, int> table_by_off;
    for (size_t t = 0; t < tables.size(); ++t) table_by_off[tables[t].off] = (int)t;
    for (int i = 0; i < np; ++i) {
      const EncodedPod& e = last_encoded[i];
      int32_t d = -1;
      if (e.h.spread_col >= 0) d = std::max(d, last_spread[e.h.spread_col]);
      const uint64_t* pw = e.pool.data() + (size_t)e.h.n_exprs * W + ((e.h.flags & PF_HAS_PORTS) ? 6 * KSB_PORT_WORDS : 0) +
                           (size_t)e.h.n_ext * KSB_CLASS_WORDS;
      for (int c = 0; c < e.h.n_checks; ++c) {
        if (((pw[c] >> 40) & 0xff) == CM_FAIL_ALWAYS) continue;
        d = std::max(d, last_table[table_by_off.at((uint32_t)pw[c])]);
      }
      pw += e.h.n_checks;
      for (int c = 0; c < e.h.n_score; ++c) d = std::max(d, last_table[table_by_off.at((uint32_t)pw[2 * c])]);
      dep[i] = d;
      for (int c : e.spread_inc) last_spread[c] = i;
      for (auto& c : e.commits) last_table[c.table] = i;
    }
  }
  auto t1 = std::chrono::steady_clock::now();
  stats.last_batch_encode_ms = std::chrono::duration<double, std::milli>(t1 - t0).count();

  sync_device();
  b_headers.ensure(headers.size() * sizeof(KsbPodHeader));
  b_pool.ensure(pool.size() * 8);
  b_results.ensure((size_t)np * sizeof(KsbResult));
  b_dep.ensure((size_t)np * 4);
  CUDA_OK(cudaMemcpyAsync(b_dep.d, dep.data(), (size_t)np * 4, cudaMemcpyHostToDevice, stream));
  CUDA_OK(cudaMemcpyAsync(b_headers.d, headers.data(), headers.size() * sizeof(KsbPodHeader), cudaMemcpyHostToDevice, stream));
  CUDA_OK(cudaMemcpyAsync(b_pool.d, pool.data(), pool.size() * 8, cudaMemcpyHostToDevice, stream));
  stats.h2d_bytes += headers.size() * sizeof(KsbPodHeader) + pool.size() * 8 + (size_t)np * 4;
  KsbBatchArgs a = make_args(np, assume);
  if (assume) snapshot_dynamic();
  last_index_before = last_node_index;

  cudaEvent_t ev0, ev1;
  CUDA_OK(cudaEventCreate(&ev0));
  CUDA_OK(cudaEventCreate(&ev1));
  CUDA_OK(cudaStreamSynchronize(stream));
  CUDA_OK(cudaEventRecord(ev0, stream));
  launch(a);
  CUDA_OK(cudaEventRecord(ev1, stream));
  CUDA_OK(cudaMemcpyAsync(last_results.data(), b_results.d, (size_t)np * sizeof(KsbResult), cudaMemcpyDeviceToHost, stream));
  int32_t flags[2] = {0, 0};
  CUDA_OK(cudaMemcpyAsync(&flags[0], a.error_flag, 4, cudaMemcpyDeviceToHost, stream));
  CUDA_OK(cudaMemcpyAsync(&flags[1], a.fiterr_count, 4, cudaMemcpyDeviceToHost, stream));
  CUDA_OK(cudaMemcpyAsync(&last_node_index, a.last_node_index, 8, cudaMemcpyDeviceToHost, stream));
  uint64_t counters[15] = {0};
  CUDA_OK(cudaMemcpyAsync(counters, a.counters, sizeof counters, cudaMemcpyDeviceToHost, stream));
  CUDA_OK(cudaStreamSynchronize(stream));
  float ms = 0;
  CUDA_OK(cudaEventElapsedTime(&ms, ev0, ev1));
  cudaEventDestroy(ev0);
  cudaEventDestroy(ev1);
  stats.last_batch_kernel_ms = ms;
  stats.d2h_bytes += (size_t)np * sizeof(KsbResult) + 16;
  if (flags[0] != 0) {
    layout_dirty = true;  // device state is unknown: rebuild it from the host records next time
    throw std::runtime_error("device kernel reported a barrier timeout");
  }
  int nfe = std::min(flags[1], (int32_t)KSB_MAX_FITERR);
  if (nfe > 0) {
    last_fiterr_hist.resize((size_t)nfe * KSB_REASON_BITS);
    CUDA_OK(cudaMemcpy(last_fiterr_hist.data(), a.fiterr_hist, last_fiterr_hist.size() * 4, cudaMemcpyDeviceToHost));
    stats.d2h_bytes += last_fiterr_hist.size() * 4;
  }
  stats.pods_scheduled += np;
  stats.groups += counters[0];
  stats.group_restarts += counters[1];
  for (int i = 0; i < 13; ++i) stats.stage_cycles[i] = counters[2 + i];

  for (int i = 0; i < np; ++i) {
    const KsbResult& r = last_results[i];
    out[i].status = r.status;
    out[i].node_index = r.row;
    out[i].num_filtered = r.num_filtered;
    out[i].score = r.score;
    for (int j = 0; j < KSB_MAX_EXT; ++j) out[i].device_mask[j] = r.device_mask[j];
    if (!assume || r.status != KSB_SCHEDULED) continue;
    // mirror the device's assume step on the host records (scheduler.go:478, storage.go:180-187)
    HostNode& hn = *rows[r.row];
    Pod& p = batch[i];
    p.node_name = hn.node.name;
    for (size_t j = 0; j < p.extended.size(); ++j) {
      p.extended[j].assigned.clear();
      if (p.extended[j].limits.size() != 1) continue;
      auto ds = hn.ext.find(p.extended[j].limits.begin()->first);
      if (ds == hn.ext.end()) continue;
      int bit = 0;
      for (auto& dev : ds->second.allocatable) {
        if ((r.device_mask[j] >> bit) & 1u) p.extended[j].assigned.push_back(dev.first);
        ++bit;
      }
    }
    apply_node_part_host(r.row, last_encoded[i], r.device_mask);
    apply_class_part_host(r.row, last_encoded[i].commits, last_encoded[i].spread_inc, +1);
    host_add_pod(p, hn);
    pods[p.key()] = HostPod{p};
  }
}

void Engine::replay_last_batch(int iters, double* avg_ms, ksb_result* out) {
  std::lock_guard<std::mutex> g(mu);
  CUDA_OK(cudaSetDevice(cfg.device));
  if (!have_snapshot || last_encoded.empty()) throw std::runtime_error("no batch to replay");
  const int np = (int)last_encoded.size();
  KsbBatchArgs a = make_args(np, true);
  cudaEvent_t ev0, ev1;
  CUDA_OK(cudaEventCreate(&ev0));
  CUDA_OK(cudaEventCreate(&ev1));
  double total = 0;
  uint64_t saved_index = last_node_index;
  const size_t flush_bytes = (size_t)256 << 20;  // larger than the 126 MB L2
  b_flush.ensure(flush_bytes);
  for (int it = 0; it < iters; ++it) {
    restore_dynamic();
    CUDA_OK(cudaMemsetAsync(b_flush.d, it & 0xff, flush_bytes, stream));  // flush L2 between timed runs
    last_node_index = last_index_before;
    CUDA_OK(cudaStreamSynchronize(stream));
    CUDA_OK(cudaEventRecord(ev0, stream));
    launch(a);
    CUDA_OK(cudaEventRecord(ev1, stream));
    CUDA_OK(cudaStreamSynchronize(stream));
    float ms = 0;
    CUDA_OK(cudaEventElapsedTime(&ms, ev0, ev1));
    total += ms;
  }
  cudaEventDestroy(ev0);
  cudaEventDestroy(ev1);
  last_node_index = saved_index;
  *avg_ms = total / std::max(1, iters);
  if (out) {
    std::vector<KsbResult> res(np);
    CUDA_OK(cudaMemcpy(res.data(), b_results.d, (size_t)np * sizeof(KsbResult), cudaMemcpyDeviceToHost));
    for (int i = 0; i < np; ++i) {
      out[i].status = res[i].status;
      out[i].node_index = res[i].row;
      out[i].num_filtered = res[i].num_filtered;
      out[i].num_nodes = n;
      out[i].score = res[i].score;
      for (int j = 0; j < KSB_MAX_EXT; ++j) out[i].device_mask[j] = res[i].device_mask[j];
    }
  }
}

// ksb_replay_last_batch_noalloc: the resident batch once more, with the kernel's working memory taken
// from the module's static device array instead of the engine's cudaMalloc'd scratch. Nothing on this
// path allocates: device state is restored with device-to-device copies into buffers that already
// exist, the counters the kernel accumulates into are re-zeroed in place (the shared array still holds
// the previous call's values), and results are copied straight into the caller's array.
void Engine::replay_last_batch_noalloc(ksb_result* out) {
  std::lock_guard<std::mutex> g(mu);
  CUDA_OK(cudaSetDevice(cfg.device));
  if (!have_snapshot || last_encoded.empty()) throw std::runtime_error("no batch to replay");
  if (!out) throw std::runtime_error("out must not be null");
  static_assert(sizeof(ksb_result) == sizeof(KsbResult), "host and device result records must have one layout");
  const int np = (int)last_encoded.size();
  KsbBatchArgs a;
  memset(&a, 0, sizeof a);
  a.table = dT;
  a.headers = (const KsbPodHeader*)b_headers.d;
  a.pool = (const uint64_t*)b_pool.d;
  a.n_pods = np;
  a.commit = 1;
  a.results = (KsbResult*)b_results.d;
  char err[256] = {0};
  a.dep = (const int32_t*)b_dep.d;
  a.zone_stride = zone_stride;
  a.nseg = nseg;
  if (ksb_noalloc_bind(cfg.device, n_pad, zone_stride, grid, &a, err, sizeof err) != 0) throw std::runtime_error(err);
  restore_dynamic();
  const uint64_t saved_index = last_node_index;
  last_node_index = last_index_before;  // launch() seeds the device counter from this member
  launch(a);                            // async copies/memsets + one kernel launch: nothing allocates
  last_node_index = saved_index;
  int32_t flag = 0;
  CUDA_OK(cudaMemcpyAsync(out, b_results.d, (size_t)np * sizeof(KsbResult), cudaMemcpyDeviceToHost, stream));
  CUDA_OK(cudaMemcpyAsync(&flag, a.error_flag, 4, cudaMemcpyDeviceToHost, stream));
  CUDA_OK(cudaStreamSynchronize(stream));
  if (flag != 0) {
    layout_dirty = true;
    throw std::runtime_error("device kernel reported a barrier timeout");
  }
  for (int i = 0; i < np; ++i) out[i].num_nodes = n;  // the device record keeps a histogram slot in this field
}

// FitError.Error() (generic_scheduler.go:70-88)
std::string Engine::fit_error_message(int i) {
  std::lock_guard<std::mutex> g(mu);
  if (i < 0 || i >= (int)last_results.size()) throw std::runtime_error("pod index out of range");
  const KsbResult& r = last_results[i];
  if (r.status != KSB_FIT_ERROR) throw std::runtime_error("pod did not end in a FitError");
  if (r.fiterr_slot < 0 || (size_t)(r.fiterr_slot + 1) * KSB_REASON_BITS > last_fiterr_hist.size())
    throw std::runtime_error("reason histogram not kept (more than KSB_MAX_FITERR no-fit pods in the batch)");
  const int32_t* hist = &last_fiterr_hist[(size_t)r.fiterr_slot * KSB_REASON_BITS];
  static const char* fixed[KSB_REASON_BITS] = {
      "Insufficient pods", "Insufficient cpu", "Insufficient memory", "Insufficient alpha.kubernetes.io/nvidia-gpu",
      "Insufficient ephemeral-storage", nullptr, nullptr, nullptr, nullptr, "HostName", "PodFitsHostPorts",
      "MatchNodeSelector", "PodToleratesNodeTaints", "NodeUnderMemoryPressure", "NodeUnderDiskPressure", "NodeNotReady",
      "NodeOutOfDisk", "NodeNetworkUnavailable", "NodeUnschedulable", "MatchInterPodAffinity",
      "ExistingPodsAntiAffinityRulesNotMatch", "PodAffinityRulesNotMatch", "PodAntiAffinityRulesNotMatch", nullptr, nullptr,
      nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
  std::map<std::string, int> reasons;
  for (int b = 0; b < KSB_REASON_BITS; ++b) {
    if (hist[b] == 0) continue;
    std::string name;
    if (fixed[b]) name = fixed[b];
    else if (b >= RB_SCALAR0 && b <= RB_SCALAR3) name = "Insufficient " + scalar_names[b - RB_SCALAR0];
    else if (b >= RB_EXT0 && b <= RB_EXT3) name = last_encoded[i].ext_reasons[b - RB_EXT0];
    else if (b == RB_EXT_LIMITS) name = last_encoded[i].ext_reasons[KSB_MAX_EXT];
    else name = "reason-" + std::to_string(b);
    reasons[name] += hist[b];
  }
  std::vector<std::string> strs;
  for (auto& kv : reasons) strs.push_back(std::to_string(kv.second) + " " + kv.first);
  std::sort(strs.begin(), strs.end());
  std::string msg = "0/" + std::to_string(last_n_nodes) + " nodes are available: ";
  for (size_t k = 0; k < strs.size(); ++k) {
    if (k) msg += ", ";
    msg += strs[k];
  }
  msg += ".";
  return msg;
}

// Cache.AssumePod / AddPod outside a batch.
void Engine::pod_add(const Pod& p) {
  std::lock_guard<std::mutex> g(mu);
  if (pods.count(p.key())) throw std::runtime_error("pod " + p.key() + " is already in the cache");  // cache.go:127-129
  HostNode& hn = nodes[p.node_name];
  if (layout_dirty || !hn.has_node || hn.row < 0) {
    // rows are about to be rebuilt from the host records (or the node has no Node object yet)
    host_add_pod(p, hn);
    pods[p.key()] = HostPod{p};
    return;
  }
  CUDA_OK(cudaSetDevice(cfg.device));
  register_pod_classes(p);
  layout_arena();
  EncodedPod e;
  encode_pod(p, &e);
  e.h.pool_off = 0;
  uint32_t masks[KSB_MAX_EXT] = {0, 0, 0, 0};
  for (size_t j = 0; j < p.extended.size() && j < KSB_MAX_EXT; ++j) {
    if (p.extended[j].limits.size() != 1) continue;
    auto ds = hn.ext.find(p.extended[j].limits.begin()->first);
    if (ds == hn.ext.end()) continue;
    int bit = 0;
    for (auto& dev : ds->second.allocatable) {
      for (auto& id : p.extended[j].assigned)
        if (id == dev.first && ds->second.available.count(id)) masks[j] |= 1u << bit;
      ++bit;
    }
  }
  sync_device();
  if (e.pool.empty()) e.pool.push_back(0);
  b_one_header.ensure(sizeof(KsbPodHeader));
  b_one_pool.ensure(e.pool.size() * 8);
  CUDA_OK(cudaMemcpyAsync(b_one_header.d, &e.h, sizeof(KsbPodHeader), cudaMemcpyHostToDevice, stream));
  CUDA_OK(cudaMemcpyAsync(b_one_pool.d, e.pool.data(), e.pool.size() * 8, cudaMemcpyHostToDevice, stream));
  int rc = ksb_launch_apply(&dT, (const KsbPodHeader*)b_one_header.d, (const uint64_t*)b_one_pool.d, hn.row, masks, stream);
  if (rc != 0) throw std::runtime_error(std::string("apply kernel failed: ") + cudaGetErrorString((cudaError_t)rc));
  CUDA_OK(cudaStreamSynchronize(stream));
  stats.kernel_launches += 1;
  stats.h2d_bytes += sizeof(KsbPodHeader) + e.pool.size() * 8;
  apply_node_part_host(hn.row, e, masks);
  apply_class_part_host(hn.row, e.commits, e.spread_inc, +1);
  host_add_pod(p, hn);
  pods[p.key()] = HostPod{p};
}

// Cache.ForgetPod / RemovePod: NodeInfo.RemovePod is not a plain inverse (ports are a set, ephemeral
// storage is never given back), so the row is recomputed from the host record and re-uploaded.
void Engine::pod_remove(const std::string& ns, const std::string& name) {
  std::lock_guard<std::mutex> g(mu);
  auto it = pods.find(ns + "/" + name);
  if (it == pods.end()) throw std::runtime_error("pod " + ns + "/" + name + " is not in the cache");
  Pod p = it->second.pod;
  pods.erase(it);
  auto nit = nodes.find(p.node_name);
  if (nit == nodes.end()) return;
  HostNode& hn = nit->second;
  host_remove_pod(p, hn);
  if (hn.pod_keys.empty() && !hn.has_node) {
    nodes.erase(nit);
    return;
  }
  if (layout_dirty || hn.row < 0) return;
  std::vector<CommitEntry> commits;
  std::vector<int> inc;
  commit_lists(p, sig_of(p), &commits, &inc);
  apply_class_part_host(hn.row, commits, inc, -1);
  fill_row(hn.row);
  dirty_rows.insert(hn.row);
  arena_dirty = true;
}

const std::string* Engine::node_name_at(int row) const {
  if (row < 0 || row >= (int)rows.size()) return nullptr;
  return &rows[row]->node.name;
}

const std::string* Engine::device_id_at(int row, const std::string& resource, int bit) const {
  if (row < 0 || row >= (int)rows.size()) return nullptr;
  auto ds = rows[row]->ext.find(resource);
  if (ds == rows[row]->ext.end()) return nullptr;
  int b = 0;
  for (auto& dev : ds->second.allocatable) {
    if (b == bit) return &dev.first;
    ++b;
  }
  return nullptr;
}

}  // namespace ksb
