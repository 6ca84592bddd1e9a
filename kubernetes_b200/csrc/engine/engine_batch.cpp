// Engine, part 3: the batch driver (ksb_schedule / ksb_schedule_batch), replay for measurement,
// FitError text, and the pod mutators that keep device and host mirror in step.
#include <algorithm>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <future>
#include <stdexcept>
#include <thread>
#include <unordered_set>

#include "engine.hpp"

#include <climits>

#include <nvtx3/nvToolsExt.h>

namespace ksb {

#define CUDA_OK(call)                                                                          \
  do {                                                                                         \
    cudaError_t e_ = (call);                                                                   \
    if (e_ != cudaSuccess)                                                                     \
      throw std::runtime_error(std::string("CUDA error: ") + cudaGetErrorString(e_) + " at " + \
                               __FILE__ + ":" + std::to_string(__LINE__));                     \
  } while (0)

namespace {
// NVTX ranges for a profiler's timeline, named after the reference's own trace of a Schedule call
// (utiltrace steps "Computing predicates", "Prioritizing", "Selecting host", core/generic_scheduler.go:110-154):
// on the device those three are phases 1-3 of one kernel, so one range covers the launch; the host work around it
// gets ranges of its own. Without a tool attached the calls are no-ops.
struct NvtxRange {
  explicit NvtxRange(const char* name) { nvtxRangePushA(name); }
  ~NvtxRange() { nvtxRangePop(); }
  NvtxRange(const NvtxRange&) = delete;
  NvtxRange& operator=(const NvtxRange&) = delete;
};
constexpr size_t kPodSlotDep = (sizeof(KsbPodHeader) + 15) & ~(size_t)15;  // offsets inside the one-pod slot
constexpr size_t kPodSlotPool = kPodSlotDep + 16;
struct ScratchLayout {
  size_t norms, zone_sum, raw, score, reasons, seg_max, seg_cnt, seg_mask, barrier, error_flag, fiterr_count, last_index,
      group_done, counters, res_slot, ghist, release, xbar, xstate, peers, pod_slot, fiterr_hist, tot_norms, tot_zone, xflag, xin_norms,
      xin_zone, base0, total;
};
ScratchLayout scratch_layout(int n_pad, int zone_stride, int nseg, int G, int halves = 1) {
  ScratchLayout L;
  size_t off = 0;
  auto take = [&](size_t bytes) {
    size_t o = off;
    off += (bytes + 255) & ~(size_t)255;
    return o;
  };
  // fixed-size control fields first: their offsets must not depend on a shard's row count, because the
  // other ranks address them inside rank 0's buffer with a layout computed from their own geometry
  L.norms = take(2 * (size_t)KSB_MAX_GROUP * sizeof(KsbNorms));
  L.zone_sum = take(2 * (size_t)KSB_MAX_GROUP * zone_stride * 4);
  L.barrier = take(8);
  L.error_flag = take(4);
  L.fiterr_count = take(4);
  L.last_index = take(8);
  L.group_done = take(32);
  L.counters = take(128);
  L.res_slot = take(sizeof(KsbResult));  // one-pod calls: the result record, read back with the control words
  L.ghist = take(2 * (size_t)KSB_MAX_GROUP * KSB_REASON_BITS * 4);
  L.release = take(8);
  L.xbar = take(4);
  L.xstate = take(4);
  L.peers = take(sizeof(KsbPeer) * KSB_MAX_RANKS);
  L.tot_norms = take(2 * (size_t)KSB_MAX_GROUP * sizeof(KsbNorms));  // sharded, ranks > 0: totals fetched from rank 0
  L.tot_zone = take(2 * (size_t)KSB_MAX_GROUP * zone_stride * 4);
  // one-pod calls: header, dependency index and pool words travel inside the control block's single upload
  L.pod_slot = take(kPodSlotPool + (size_t)ksb_max_pool_words() * 8);
  L.fiterr_hist = take((size_t)KSB_MAX_FITERR * KSB_REASON_BITS * 4);
  // sharded: the inbox peers write into (behind the one-pod upload block, still independent of the row count)
  L.xflag = take(256);  // +0: barrier flags [KSB_MAX_RANKS], +64: evaluation-done flags [KSB_MAX_RANKS], +128: commit-done words [2] x 8 bytes
  // two halves x one slot per source rank x one 8-byte word {value, exchange number} per value
  L.xin_norms = take(2 * (size_t)KSB_MAX_RANKS * KSB_MAX_GROUP * 8 * 8);
  L.xin_zone = take(2 * (size_t)KSB_MAX_RANKS * KSB_MAX_GROUP * zone_stride * 8);
  L.raw = take(halves * (size_t)G * n_pad * 16);
  L.score = take(halves * (size_t)G * n_pad * 4);
  L.reasons = take(halves * (size_t)G * n_pad * 4);
  L.seg_max = take(halves * (size_t)G * nseg * 4);
  L.seg_cnt = take(halves * (size_t)G * nseg * 4);
  L.seg_mask = take(halves * (size_t)G * nseg * 32);
  L.base0 = take(halves * (size_t)G * n_pad * 4);  // sharded engines only use it (rank 0's copy is the one read)
  L.total = off;
  return L;
}
}  // namespace

// Geometry of the next launch (group size, segments, grid, shared memory) and the scratch buffer for it.
void Engine::prepare_geometry() {
  const int R = std::max(1, cfg.shard_count);
  zone_stride = (dT.n_zones + 63) & ~63;
  nseg = (ln + 255) / 256;           // segments of this GPU's table (rank 0 of a sharded engine: all of them)
  eval_nseg = (eval_n + 255) / 256;  // segments this GPU evaluates
  segs_per_shard = R == 1 ? std::max(1, nseg) : rows_per_shard / 256;
  nseg_total = R == 1 ? nseg : (n + 255) / 256;
  // group size: as configured, bounded by what the zone sums and the segment statistics may take of shared memory
  group = cfg.group_size > 0 ? std::min(cfg.group_size, KSB_MAX_GROUP) : KSB_MAX_GROUP;
  group = std::max(1, std::min(group, 16384 / zone_stride));
  if (nseg_total > KSB_MAX_SEGMENTS) throw std::runtime_error("more than 524,288 node rows");
  group = std::max(1, std::min(group, KSB_SEG_CACHE / std::max(1, nseg_total)));
  // pipelined launches: the row cache holds the previous group's rows next to the current one's
  pipe = cfg.pipeline >= 0 && (cfg.pipeline > 0 || n >= 16384);  // the same on every rank: decided by the whole table
  if (pipe) group = std::min(group, KSB_MAX_GROUP / 2);
  if (group < 2) pipe = false;
  if (zone_stride * 4 * KSB_MAX_GROUP > 160 * 1024) throw std::runtime_error("more than 1,280 distinct zones");
  smem_bytes = ksb_batch_smem_bytes(zone_stride);
  if (limits_for_smem != smem_bytes) {  // attribute + occupancy queries: once per shared-memory size
    int mg = 0, th = 0;
    int rc = ksb_batch_kernel_limits(cfg.device, smem_bytes, &mg, &th);
    if (rc != 0) throw std::runtime_error(std::string("kernel attribute query failed: ") + cudaGetErrorString((cudaError_t)rc));
    max_grid = mg;
    threads = th;
    limits_for_smem = smem_bytes;
  }
  if (max_grid < 1) throw std::runtime_error("batch kernel does not fit on this device");
  // enough blocks to cover every segment; when segments are few, extra blocks split the group's pods
  grid = eval_nseg >= max_grid ? max_grid : std::max(1, std::min(max_grid, std::max(1, eval_nseg) * group));
  if (R > 1) grid = max_grid;  // every rank runs the same schedule (block roles depend on the block index)
  if (pipe && grid < 2) pipe = false;
  ScratchLayout L = scratch_layout(ln_pad, zone_stride, std::max(1, nseg), group, pipe ? 2 : 1);
  if (shard_connected && L.total > b_scratch.cap)
    throw std::runtime_error("sharded engine: scratch outgrew the shared buffer (re-run ksb_shard_export/connect)");
  b_scratch.ensure(L.total);
}

KsbBatchArgs Engine::make_args(int n_pods, bool commit) {
  const int R = std::max(1, cfg.shard_count);
  prepare_geometry();
  ScratchLayout L = scratch_layout(ln_pad, zone_stride, std::max(1, nseg), group, pipe ? 2 : 1);
  char* base = (char*)b_scratch.d;
  KsbBatchArgs a;
  memset(&a, 0, sizeof a);
  a.table = dT;
  a.table.ext_keep = n_pods == 1 ? call_ext_keep : nullptr;  // extender verdicts belong to one pod
  a.table.ext_score = n_pods == 1 ? call_ext_score : nullptr;
  a.headers = (const KsbPodHeader*)b_headers.d;
  a.pool = (const uint64_t*)b_pool.d;
  a.dep = (const int32_t*)b_dep.d;
  a.n_pods = n_pods;
  a.zone_stride = zone_stride;
  a.commit = commit ? 1 : 0;
  a.group_size = group;
  a.pipeline = (pipe && n_pods > group && (batch_pipe_ok || cfg.pipeline > 0)) ? 1 : 0;  // a batch of one group gains nothing from the extra iteration
  a.nseg = eval_nseg;  // segments this GPU evaluates
  a.eval_n = eval_n;
  a.seg_gbase = eval_base / 256;
  a.results = (KsbResult*)b_results.d;
  a.norms = (KsbNorms*)(base + L.norms);
  a.zone_sum = (int32_t*)(base + L.zone_sum);
  a.raw = (int4*)(base + L.raw);
  a.score = (int32_t*)(base + L.score);
  a.reasons = (uint32_t*)(base + L.reasons);
  a.seg_max = (int32_t*)(base + L.seg_max);
  a.seg_cnt = (int32_t*)(base + L.seg_cnt);
  a.seg_mask = (uint32_t*)(base + L.seg_mask);
  a.barrier = (uint32_t*)(base + L.barrier);
  a.error_flag = (int32_t*)(base + L.error_flag);
  a.fiterr_count = (int32_t*)(base + L.fiterr_count);
  // reason histograms: one per pod of the launch, so that no FitError of a large single launch loses its text
  a.fiterr_cap = std::max(n_pods, (int)KSB_MAX_FITERR);
  if (a.fiterr_cap > (int)KSB_MAX_FITERR) {
    b_fiterr.ensure((size_t)a.fiterr_cap * KSB_REASON_BITS * 4);
    a.fiterr_hist = (int32_t*)b_fiterr.d;
  } else {
    a.fiterr_hist = (int32_t*)(base + L.fiterr_hist);
  }
  a.last_node_index = (uint64_t*)(base + L.last_index);
  a.group_done = (int32_t*)(base + L.group_done);
  a.counters = (uint64_t*)(base + L.counters);
  a.ghist = (int32_t*)(base + L.ghist);
  a.release = (uint32_t*)(base + L.release);
  a.xbar = (uint32_t*)(base + L.xbar);
  a.xstate = (uint32_t*)(base + L.xstate);
  a.peers = (const KsbPeer*)(base + L.peers);
  a.n_ranks = R;
  a.rank = R == 1 ? 0 : cfg.shard_rank;
  a.segs_per_shard = segs_per_shard;
  a.nseg_total = nseg_total;
  // accumulators and totals of the Reduce inputs: the targets themselves, except on ranks > 0 of a sharded engine
  a.acc_norms = a.norms;
  a.tot_norms = a.norms;
  a.acc_zone_sum = a.zone_sum;
  a.tot_zone_sum = a.zone_sum;
  a.acc_ghist = a.ghist;
  a.l_group_done = a.group_done;
  if (R > 1) {
    if (!shard_connected) throw std::runtime_error("sharded engine: call ksb_shard_export / ksb_shard_connect first");
    // Reduce inputs: every rank accumulates its own partials (acc_*), stores them into its slot of every rank's
    // inbox at the barrier after phase 1 and combines the totals (tot_*) from its own inbox
    a.tot_norms = (KsbNorms*)(base + L.tot_norms);
    a.tot_zone_sum = (int32_t*)(base + L.tot_zone);
    for (int q = 0; q < R; ++q) a.xpeer[q] = q == a.rank ? base : (char*)peer_scratch[q];
    a.xo_flag = (uint32_t)L.xflag;
    a.xo_norms = (uint32_t)L.xin_norms;
    a.xo_zone = (uint32_t)L.xin_zone;
    a.xo_gdone = (uint32_t)L.group_done;
    if (a.rank != 0) {  // FitError histograms and the commit's progress words live in rank 0's memory
      char* b0 = (char*)peer0_scratch;
      a.norms = (KsbNorms*)(b0 + L.norms);
      a.zone_sum = (int32_t*)(b0 + L.zone_sum);
      a.ghist = (int32_t*)(b0 + L.ghist);
      a.group_done = (int32_t*)(b0 + L.group_done);
      a.xbar = (uint32_t*)(b0 + L.xbar);
    }
  }
  // where the evaluation's results for phase 3 go: own arrays, or rank 0's (indexed by global row / segment)
  a.gs_seg_max = a.seg_max;
  a.gs_seg_cnt = a.seg_cnt;
  a.gs_seg_mask = a.seg_mask;
  a.gs_np = ln_pad;
  a.gs_nseg = std::max(1, nseg);
  a.gs_score = a.score;
  a.gs_reasons = a.reasons;
  a.gs_base0 = (int32_t*)(base + L.base0);
  a.gs_row0 = eval_base;
  if (R > 1 && a.rank != 0) {
    // rank 0's scratch layout follows from the whole table's geometry, which every rank knows
    ScratchLayout L0 = scratch_layout(n_pad, zone_stride, std::max(1, nseg_total), group, pipe ? 2 : 1);
    char* b0 = (char*)peer0_scratch;
    a.gs_seg_max = (int32_t*)(b0 + L0.seg_max);
    a.gs_seg_cnt = (int32_t*)(b0 + L0.seg_cnt);
    a.gs_seg_mask = (uint32_t*)(b0 + L0.seg_mask);
    a.gs_np = n_pad;
    a.gs_nseg = std::max(1, nseg_total);
    a.gs_score = (int32_t*)(b0 + L0.score);
    a.gs_reasons = (uint32_t*)(b0 + L0.reasons);
    a.gs_base0 = (int32_t*)(b0 + L0.base0);
  }
  return a;
}

// Zero what must start at zero, seed the round-robin counter, launch. Works on whatever scratch the
// arguments point at (the engine's own buffer or the static no-allocation array).
void Engine::launch(const KsbBatchArgs& a_in) {
  KsbBatchArgs a = a_in;
  a.self.table = a.table;  // this GPU's shard as a peer entry, passed in the kernel's parameters
  a.self.raw = a.raw;
  a.self.score = a.score;
  a.self.reasons = a.reasons;
  a.self.seg_max = a.seg_max;
  a.self.seg_cnt = a.seg_cnt;
  a.self.seg_mask = a.seg_mask;
  a.self.results = a.results;
  a.self.last_node_index = a.last_node_index;
  a.self.nseg = a.gs_nseg;  // phase 3 (rank 0 only) reads the arrays the results were gathered into
  a.self.seg_base = 0;
  a.self.raw_np = a.table.n_pad;
  a.self.raw_row0 = 0;
  if (a.n_ranks == 1 && b_scratch.d && (char*)a.norms == (char*)b_scratch.d) {
    // the engine's own scratch on one GPU: every control field lies in one contiguous block at the start of the
    // buffer, so its initial image goes up in a single copy from page-locked memory
    ScratchLayout L = scratch_layout(ln_pad, a.zone_stride, std::max(1, a.nseg), a.group_size, a.pipeline ? 2 : 1);
    // the one-pod slot is the tail of the block: uploaded only when a one-pod call staged its pod there
    const size_t bytes = L.pod_slot + (one_pod_bytes ? one_pod_bytes : 0);
    pin_ctl.ensure(L.fiterr_hist);
    char* img = (char*)pin_ctl.p;
    memset(img, 0, L.pod_slot);
    if (one_pod_bytes) memcpy(img + L.pod_slot, one_pod_src, one_pod_bytes);
    KsbNorms* nm = (KsbNorms*)(img + L.norms);
    for (int i = 0; i < 2 * KSB_MAX_GROUP; ++i) nm[i].one_row = -1;
    memcpy(img + L.last_index, &last_node_index, 8);
    KsbPeer* self = (KsbPeer*)(img + L.peers);
    self->table = a.table;
    self->raw = a.raw;
    self->score = a.score;
    self->reasons = a.reasons;
    self->seg_max = a.seg_max;
    self->seg_cnt = a.seg_cnt;
    self->seg_mask = a.seg_mask;
    self->results = a.results;
    self->last_node_index = a.last_node_index;
    self->nseg = a.gs_nseg;
    self->seg_base = 0;
    self->raw_np = a.table.n_pad;
    self->raw_row0 = 0;
    CUDA_OK(cudaMemcpyAsync(b_scratch.d, img, bytes, cudaMemcpyHostToDevice, stream));
    int rc = ksb_launch_batch(&a, grid, smem_bytes, stream);
    if (rc != 0) throw std::runtime_error(std::string("kernel launch failed: ") + cudaGetErrorString((cudaError_t)rc));
    stats.kernel_launches += 1;
    return;
  }
  // host staging of this launch lives in the engine (launch_norms, launch_self), never in function statics:
  // several engines of one process may launch from different threads
  launch_norms.assign(2 * KSB_MAX_GROUP, KsbNorms());
  for (auto& x : launch_norms) {
    memset(&x, 0, sizeof x);
    x.one_row = -1;
  }
  if (a.n_ranks == 1) {  // sharded: rank 0's kernel clears these itself, behind the opening barrier
    CUDA_OK(cudaMemcpyAsync(a.norms, launch_norms.data(), launch_norms.size() * sizeof(KsbNorms), cudaMemcpyHostToDevice, stream));
    CUDA_OK(cudaMemsetAsync(a.zone_sum, 0, 2 * (size_t)KSB_MAX_GROUP * a.zone_stride * 4, stream));
    CUDA_OK(cudaMemsetAsync(a.ghist, 0, 2 * (size_t)KSB_MAX_GROUP * KSB_REASON_BITS * 4, stream));
    CUDA_OK(cudaMemsetAsync(a.group_done, 0, 32, stream));
  }
  CUDA_OK(cudaMemsetAsync(a.barrier, 0, 8, stream));  // two counters: slot 1 is the barrier between phases 1 and 2
  CUDA_OK(cudaMemsetAsync(a.error_flag, 0, 4, stream));
  CUDA_OK(cudaMemsetAsync(a.fiterr_count, 0, 4, stream));
  CUDA_OK(cudaMemsetAsync(a.counters, 0, 128, stream));
  CUDA_OK(cudaMemcpyAsync(a.last_node_index, &last_node_index, 8, cudaMemcpyHostToDevice, stream));
  CUDA_OK(cudaMemsetAsync(a.release, 0, 8, stream));
  KsbPeer& self = launch_self;  // what rank 0 sees of this (only) shard: its own buffers
  if (a.n_ranks == 1) {
    CUDA_OK(cudaMemsetAsync(a.xbar, 0, 4, stream));
    CUDA_OK(cudaMemsetAsync(a.xstate, 0, 4, stream));
    self.table = a.table;
    self.raw = a.raw;
    self.score = a.score;
    self.reasons = a.reasons;
    self.seg_max = a.seg_max;
    self.seg_cnt = a.seg_cnt;
    self.seg_mask = a.seg_mask;
    self.results = a.results;
    self.last_node_index = a.last_node_index;
    self.nseg = a.gs_nseg;
    self.seg_base = 0;
    self.raw_np = a.table.n_pad;
    self.raw_row0 = 0;
    CUDA_OK(cudaMemcpyAsync((void*)a.peers, &self, sizeof self, cudaMemcpyHostToDevice, stream));
  } else if (a.rank == 0) {
    // the peer table: this rank from its own buffers, the others from their mapped buffers and geometry
    std::vector<KsbPeer> pt(a.n_ranks);
    pt[0].table = a.table;
    pt[0].raw = a.raw;
    pt[0].score = a.score;
    pt[0].reasons = a.reasons;
    pt[0].seg_max = a.seg_max;
    pt[0].seg_cnt = a.seg_cnt;
    pt[0].seg_mask = a.seg_mask;
    pt[0].results = a.results;
    pt[0].last_node_index = a.last_node_index;
    pt[0].nseg = a.gs_nseg;
    pt[0].seg_base = 0;
    pt[0].raw_np = a.table.n_pad;
    pt[0].raw_row0 = 0;
    for (int q = 1; q < a.n_ranks; ++q) {
      const PeerMap& m = peer_maps[q];
      KsbPeer& P = pt[q];
      P.table = a.table;  // weights, W, zone count
      P.table.n = m.ln;
      P.table.n_pad = m.ln_pad;
      P.table.row_base = m.lbase;
      fill_table_pointers(&P.table, m.p[0], m.p[1], m.p[2], m.p[3], m.p[4], m.p[5], m.p[6], (size_t)m.ln_pad);
      const int nseg_q = (m.ln + 255) / 256;
      ScratchLayout Lq = scratch_layout(m.ln_pad, a.zone_stride, std::max(1, nseg_q), a.group_size, pipe ? 2 : 1);
      char* sb = (char*)m.p[7];
      P.raw = (int4*)(sb + Lq.raw);
      P.score = (int32_t*)(sb + Lq.score);
      P.reasons = (uint32_t*)(sb + Lq.reasons);
      P.seg_max = (int32_t*)(sb + Lq.seg_max);
      P.seg_cnt = (int32_t*)(sb + Lq.seg_cnt);
      P.seg_mask = (uint32_t*)(sb + Lq.seg_mask);
      P.last_node_index = (uint64_t*)(sb + Lq.last_index);
      P.results = (KsbResult*)m.p[8];
      P.nseg = nseg_q;
      P.seg_base = q * a.segs_per_shard;
      P.raw_np = m.ln_pad;
      P.raw_row0 = m.lbase;
    }
    CUDA_OK(cudaMemcpy((void*)a.peers, pt.data(), sizeof(KsbPeer) * pt.size(), cudaMemcpyHostToDevice));
  }
  CUDA_OK(cudaStreamSynchronize(stream));  // last_node_index may change right after this call
  int rc = ksb_launch_batch(&a, grid, smem_bytes, stream);
  if (rc != 0) throw std::runtime_error(std::string("kernel launch failed: ") + cudaGetErrorString((cudaError_t)rc));
  stats.kernel_launches += 1;
}

void Engine::snapshot_dynamic() {
  // everything the kernel mutates: int64 family, u32 family, u64 family (ports), spread, tables
  size_t s0 = (size_t)I64_COLS * ln_pad * 8, s1 = (size_t)U32_COLS * ln_pad * 4, s2 = (size_t)u64_cols() * ln_pad * 8;
  size_t s3 = h_spread.size() * (size_t)ln_pad * 4, s4 = (size_t)std::max<uint32_t>(arena_words, 1) * 4;
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
  size_t s0 = (size_t)I64_COLS * ln_pad * 8, s1 = (size_t)U32_COLS * ln_pad * 4, s2 = (size_t)u64_cols() * ln_pad * 8;
  size_t s3 = h_spread.size() * (size_t)ln_pad * 4, s4 = (size_t)std::max<uint32_t>(arena_words, 1) * 4;
  char* d = (char*)b_snapshot.d;
  CUDA_OK(cudaMemcpyAsync(b_i64.d, d, s0, cudaMemcpyDeviceToDevice, stream));
  CUDA_OK(cudaMemcpyAsync(b_u32.d, d + s0, s1, cudaMemcpyDeviceToDevice, stream));
  CUDA_OK(cudaMemcpyAsync(b_u64.d, d + s0 + s1, s2, cudaMemcpyDeviceToDevice, stream));
  if (s3) CUDA_OK(cudaMemcpyAsync(b_spread.d, d + s0 + s1 + s2, s3, cudaMemcpyDeviceToDevice, stream));
  CUDA_OK(cudaMemcpyAsync(b_tables.d, d + s0 + s1 + s2 + s3, s4, cudaMemcpyDeviceToDevice, stream));
}

// ksb_schedule / ksb_schedule_batch. A selector may mention more distinct label pairs than the label columns have bits
// for: the columns are then doubled (up to 16 words) and the call starts over, unless chunks of it are already committed.
void Engine::schedule_batch(std::vector<Pod>& batch, ksb_result* out, bool assume) {
  for (int attempt = 0;; ++attempt) {
    try {
      schedule_batch_once(batch, out, assume);
      return;
    } catch (const GrowLabels&) {
      std::lock_guard<std::mutex> g(mu);
      if (batch_posted > 0)
        throw std::runtime_error("label universe outgrew its words inside a chunked batch: the first " + std::to_string(batch_posted) +
                                 " pods are committed (see the results), resubmit the rest");
      if (attempt >= 3) throw std::runtime_error("label universe keeps outgrowing its words");
      layout_dirty = true;  // rebuild_layout takes pending_W
    }
  }
}

void Engine::schedule_batch_once(std::vector<Pod>& batch, ksb_result* out, bool assume) {
  std::lock_guard<std::mutex> g(mu);
  batch_posted = 0;
  CUDA_OK(cudaSetDevice(cfg.device));
  const int np = (int)batch.size();
  NvtxRange r_call(np == 1 ? "Scheduling (ksb_schedule)" : "Scheduling (ksb_schedule_batch)");
  auto t0 = std::chrono::steady_clock::now();
  static const bool trace = getenv("KSB_TRACE") != nullptr;  // host timing breakdown of large batches on stderr
  std::vector<std::pair<const char*, double>> marks;
  auto mark = [&](const char* what) {
    if (trace) marks.push_back({what, std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count()});
  };
  if (layout_dirty) rebuild_layout();
  last_n_nodes = n;
  last_encoded.resize(np);  // encode_pod resets every field of an entry
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
  for (auto& p : batch)
    if (p.num_volumes > 0)  // volume predicates are out of scope: refuse loudly, never guess, before anything is placed
      throw std::runtime_error("pod " + p.key() + " has volumes: not supported by the GPU engine");
  if (assume) {
    // ksb_schedule_batch == n x {ksb_schedule; ksb_pod_add}: a key that is already cached, or repeated inside the
    // batch, is the error ksb_pod_add reports (cache.go:127-129); refuse before anything is placed
    // repeated keys inside the batch: 64-bit hashes, exact comparison only where two hashes meet
    auto key_hash = [](const Pod& p) {
      uint64_t h = 1469598103934665603ull;
      for (unsigned char ch : p.ns) h = (h ^ ch) * 1099511628211ull;
      h = (h ^ (unsigned char)'/') * 1099511628211ull;
      for (unsigned char ch : p.name) h = (h ^ ch) * 1099511628211ull;
      return h;
    };
    // (an open-addressing table of the batch's indices under their hashes: a tenth of the time of sorting them)
    size_t cap = 1;
    while (cap < 2 * (size_t)np) cap <<= 1;
    std::vector<int32_t> slot(cap, -1);
    std::vector<uint64_t> hv((size_t)np);
    for (int i = 0; i < np; ++i) {
      const uint64_t h = key_hash(batch[i]);
      hv[i] = h;
      size_t s = (size_t)(h ^ (h >> 29)) & (cap - 1);
      while (slot[s] >= 0) {
        const int j = slot[s];
        if (hv[j] == h && batch[j].ns == batch[i].ns && batch[j].name == batch[i].name)
          throw std::runtime_error("pod " + batch[i].key() + " is already in the cache");
        s = (s + 1) & (cap - 1);
      }
      slot[s] = i;
    }
    if (!pods.empty())
      for (auto& p : batch) {
        const std::string k = p.key();
        if (pods.count(k)) throw std::runtime_error("pod " + k + " is already in the cache");
      }
    pods.reserve(pods.size() + (size_t)np);  // no rehash while the mirror thread inserts
  }
  mark("checks");
  // pass 1: classes (every pod of the batch, so that commit lists name every class a later pod introduces)
  for (auto& p : batch) register_pod_classes(p);
  layout_arena();
  mark("classes");

  // Large batches on one GPU go to the device in chunks: while the kernel of chunk c runs, the host encodes chunk
  // c+1 and mirrors the placements of chunk c-1 into its records. Device work stays strictly ordered on one stream
  // (uploads and the next launch are issued only after the previous chunk's results are back), so the placements
  // are those of one launch over the whole batch.
  constexpr int kChunk = KSB_MAX_FITERR;  // 4,096: every no-fit pod of a chunk keeps its reason histogram
  const bool chunked = assume && cfg.shard_count <= 1 && np >= 2 * kChunk;
  const int chunk_len = chunked ? kChunk : np;
  const int n_chunks = (np + chunk_len - 1) / chunk_len;

  // dep[j]: the latest earlier pod whose assume step bumps a counter pod j reads (its SelectorSpread column
  // or any inter-pod-affinity table). A group never holds both, so those inputs are constant inside a group.
  last_dep.assign(np, -1);
  std::vector<int32_t> last_spread(h_spread.size(), -1), last_table(tables.size(), -1);
  std::unordered_map<uint32_t, int> table_by_off;
  if (np > 1)
    for (size_t t = 0; t < tables.size(); ++t) table_by_off[tables[t].off] = (int)t;
  size_t pool_words = 0;  // words of every pod encoded so far (pool offsets are relative to the whole batch)
  double encode_ms = 0, post_ms = 0, kernel_ms = 0, sync_ms = 0;
  int recent_deps = 0;

  struct Staged {
    size_t hdr_bytes = 0, dep_off = 0, pool_off = 0, pool_words = 0, pool_first = 0;
  };
  Staged staged[2];

  // encode pods [c0, c1) into page-locked staging buffer `which`
  auto encode_chunk = [&](int c0, int c1, int which) {
    NvtxRange r_enc("predicateMetadata / priorityMetadata (encode pods)");
    auto te0 = std::chrono::steady_clock::now();
    const int len = c1 - c0;
    const size_t first_word = pool_words;
    // Large chunks are encoded on several threads in read-only mode: a pod that needs something new interned (a
    // label bit, a signature, a class-mask id ...) is left for the main thread, which encodes those afterwards in
    // stream order. Interned ids are arbitrary, so the order in which they are handed out does not matter.
    std::vector<char> done((size_t)len, 0);
    const unsigned hw = std::max(1u, std::thread::hardware_concurrency());
    const int workers = len >= 1024 ? (int)std::min(6u, hw > 3 ? hw - 2 : 1u) : 1;
    if (workers > 1) {
      std::vector<std::string> errors((size_t)workers);
      auto work = [&](int t) {
        tl_readonly = true;
        const int lo = c0 + (int)((int64_t)len * t / workers), hi = c0 + (int)((int64_t)len * (t + 1) / workers);
        for (int i = lo; i < hi; ++i) {
          try {
            encode_pod(batch[i], &last_encoded[i]);
            done[i - c0] = 1;
          } catch (const NeedsIntern&) {
          } catch (const std::exception&) {
            // limits and malformed pods: the main thread meets the same error in stream order and reports it
          }
        }
        tl_readonly = false;
      };
      std::vector<std::thread> pool;
      for (int t = 0; t < workers; ++t) pool.emplace_back(work, t);
      for (auto& th : pool) th.join();
    }
    if (c0 == 0) mark("first encode: parallel part");
    for (int i = c0; i < c1; ++i) {
      if (!done[i - c0]) encode_pod(batch[i], &last_encoded[i]);
      last_encoded[i].h.pool_off = (uint32_t)pool_words;
      pool_words += last_encoded[i].pool.size();
    }
    if (arena_dirty && c0 > 0) throw std::runtime_error("internal: the counter arena changed in the middle of a batch");
    if (c0 == 0) mark("first encode: serial part");
    Staged& st = staged[which];
    st.hdr_bytes = (size_t)len * sizeof(KsbPodHeader);
    st.dep_off = (st.hdr_bytes + 255) & ~(size_t)255;
    st.pool_off = (st.dep_off + (size_t)len * 4 + 255) & ~(size_t)255;
    st.pool_first = first_word;
    st.pool_words = pool_words - first_word;
    pin_chunk[which].ensure(st.pool_off + std::max<size_t>(1, st.pool_words) * 8);
    if (c0 == 0) mark("first encode: staging buffer");
    KsbPodHeader* headers = (KsbPodHeader*)pin_chunk[which].p;
    int32_t* dep = (int32_t*)((char*)pin_chunk[which].p + st.dep_off);
    uint64_t* pool = (uint64_t*)((char*)pin_chunk[which].p + st.pool_off);
    pool[0] = 0;
    for (int i = c0; i < c1; ++i) {
      const EncodedPod& e = last_encoded[i];
      headers[i - c0] = e.h;
      if (!e.pool.empty()) memcpy(pool + (e.h.pool_off - first_word), e.pool.data(), e.pool.size() * 8);
      int32_t d = -1;
      if (np > 1) {
        if (e.h.spread_col >= 0) d = std::max(d, last_spread[e.h.spread_col]);
        const uint64_t* pw = e.pool.data() + (size_t)e.h.n_exprs * W + ((e.h.flags & PF_HAS_PORTS) ? 6 * KSB_PORT_WORDS : 0) +
                             (size_t)e.h.n_ext * KSB_CLASS_WORDS;
        for (int c = 0; c < e.h.n_checks; ++c) {
          if (((pw[c] >> 40) & 0xff) == CM_FAIL_ALWAYS) continue;
          d = std::max(d, last_table[table_by_off.at((uint32_t)pw[c])]);
        }
        pw += e.h.n_checks;
        for (int c = 0; c < e.h.n_score; ++c) d = std::max(d, last_table[table_by_off.at((uint32_t)pw[2 * c])]);
        for (int c : e.spread_inc) last_spread[c] = i;
        for (auto& c : e.commits) last_table[c.table] = i;
      }
      last_dep[i] = d;
      dep[i - c0] = d >= c0 ? d - c0 : -1;  // relative to the launch; pods of earlier chunks are committed by then
      recent_deps += (d >= 0 && d >= i - KSB_MAX_GROUP) ? 1 : 0;
    }
    encode_ms += std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - te0).count();
  };

  KsbBatchArgs a;
  size_t ctl_lo = 0, ctl_hi = 0, res_off = 0;
  const bool one_call = np == 1 && !assume && cfg.shard_count <= 1;
  // upload staged chunk `which` (pods [c0, c1)) and launch the kernel on it; the stream is idle when this is called
  auto launch_chunk = [&](int c0, int c1, int which) {
    auto ts0 = std::chrono::steady_clock::now();
    const int len = c1 - c0;
    const Staged& st = staged[which];
    // The pipelined schedule evaluates a group next to the commit of the one before it, so a pod may not read a
    // counter written by either. Streams where many pods depend on a very recent pod (dense inter-pod affinity)
    // would stall it at every group: those run the plain schedule.
    batch_pipe_ok = recent_deps * 10 <= c1;
    {
      NvtxRange r_sync("UpdateNodeNameToInfoMap (table sync)");
      sync_device();
    }
    NvtxRange r_launch("Computing predicates + Prioritizing + Selecting host (upload, launch)");
    sync_ms += std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - ts0).count();
    char* src = (char*)pin_chunk[which].p;
    if (one_call) {
      // ksb_schedule: the pod's block rides in the control block's upload and its result record comes back with the
      // control words, so the call costs one copy each way
      a = make_args(len, assume);
      ScratchLayout L = scratch_layout(ln_pad, zone_stride, std::max(1, nseg), group, pipe ? 2 : 1);
      one_pod_buf.assign(kPodSlotPool + std::max<size_t>(1, st.pool_words) * 8, 0);
      memcpy(one_pod_buf.data(), src, sizeof(KsbPodHeader));
      memcpy(one_pod_buf.data() + kPodSlotDep, src + st.dep_off, 4);
      memcpy(one_pod_buf.data() + kPodSlotPool, src + st.pool_off, std::max<size_t>(1, st.pool_words) * 8);
      one_pod_src = one_pod_buf.data();
      one_pod_bytes = one_pod_buf.size();
      char* base = (char*)b_scratch.d;
      a.headers = (const KsbPodHeader*)(base + L.pod_slot);
      a.dep = (const int32_t*)(base + L.pod_slot + kPodSlotDep);
      a.pool = (const uint64_t*)(base + L.pod_slot + kPodSlotPool);
      a.results = (KsbResult*)(base + L.res_slot);
      stats.h2d_bytes += one_pod_bytes;
      last_index_before = last_node_index;
      ctl_lo = (char*)a.barrier - base;
      ctl_hi = (char*)a.results - base + sizeof(KsbResult);
      res_off = (char*)a.results - (char*)a.barrier;
      pin_out.ensure(ctl_hi - ctl_lo);
      launch(a);
      one_pod_bytes = 0;
      CUDA_OK(cudaMemcpyAsync(pin_out.p, base + ctl_lo, ctl_hi - ctl_lo, cudaMemcpyDeviceToHost, stream));
      return;
    }
    one_pod_bytes = 0;
    b_headers.ensure((size_t)np * sizeof(KsbPodHeader));  // whole-batch buffers: the replay entry points run them in one launch
    b_dep.ensure((size_t)np * 4);
    if (shard_connected && (size_t)np * sizeof(KsbResult) > b_results.cap)
      throw std::runtime_error("sharded engine: batch larger than the max_batch_pods given to ksb_shard_export");
    b_results.ensure((size_t)np * sizeof(KsbResult));
    if (c0 == 0) {
      // pool size of the whole batch, estimated from the first chunk; grown with its contents kept if that falls short
      const size_t guess = n_chunks == 1 ? st.pool_words : (st.pool_words * (size_t)np / (size_t)len) * 5 / 4 + 4096;
      b_pool.ensure(std::max<size_t>(1, guess) * 8);
    }
    grow_keep(b_pool, std::max<size_t>(1, st.pool_first + st.pool_words) * 8);
    CUDA_OK(cudaMemcpyAsync((KsbPodHeader*)b_headers.d + c0, src, st.hdr_bytes, cudaMemcpyHostToDevice, stream));
    CUDA_OK(cudaMemcpyAsync((int32_t*)b_dep.d + c0, src + st.dep_off, (size_t)len * 4, cudaMemcpyHostToDevice, stream));
    CUDA_OK(cudaMemcpyAsync((uint64_t*)b_pool.d + st.pool_first, src + st.pool_off, std::max<size_t>(1, st.pool_words) * 8,
                            cudaMemcpyHostToDevice, stream));
    stats.h2d_bytes += st.hdr_bytes + st.pool_words * 8 + (size_t)len * 4;
    a = make_args(len, assume);
    a.headers = (const KsbPodHeader*)b_headers.d + c0;
    a.dep = (const int32_t*)b_dep.d + c0;
    a.results = (KsbResult*)b_results.d + c0;
    if (c0 == 0) {
      if (assume) snapshot_dynamic();
      last_index_before = last_node_index;
    }
    // what comes back: the control words (error flag, FitError count, round-robin counter, counters) in one
    // copy, then the result records
    ctl_lo = (char*)a.barrier - (char*)b_scratch.d;
    ctl_hi = (char*)a.counters - (char*)b_scratch.d + 128;
    res_off = (ctl_hi - ctl_lo + 255) & ~(size_t)255;
    pin_out.ensure(res_off + (size_t)len * sizeof(KsbResult));
    CUDA_OK(cudaEventRecord(ev_start, stream));
    launch(a);
    CUDA_OK(cudaEventRecord(ev_stop, stream));
    char* back = (char*)pin_out.p;
    CUDA_OK(cudaMemcpyAsync(back, (char*)b_scratch.d + ctl_lo, ctl_hi - ctl_lo, cudaMemcpyDeviceToHost, stream));
    CUDA_OK(cudaMemcpyAsync(back + res_off, (KsbResult*)b_results.d + c0, (size_t)len * sizeof(KsbResult), cudaMemcpyDeviceToHost,
                            stream));
  };

  uint64_t counters[16] = {0};
  // wait for the chunk launched last and take its results
  auto wait_chunk = [&](int c0, int c1) {
    const int len = c1 - c0;
    {
      NvtxRange r_wait("Computing predicates + Prioritizing + Selecting host (device)");
      CUDA_OK(cudaStreamSynchronize(stream));
    }
    float ms = 0;
    if (!one_call) CUDA_OK(cudaEventElapsedTime(&ms, ev_start, ev_stop));  // one-pod calls are not timed on the device
    kernel_ms += ms;
    char* back = (char*)pin_out.p;
    memcpy(last_results.data() + c0, back + res_off, (size_t)len * sizeof(KsbResult));
    int32_t flags[2];
    memcpy(&flags[0], back + ((char*)a.error_flag - (char*)a.barrier), 4);
    memcpy(&flags[1], back + ((char*)a.fiterr_count - (char*)a.barrier), 4);
    memcpy(&last_node_index, back + ((char*)a.last_node_index - (char*)a.barrier), 8);
    memcpy(counters, back + ((char*)a.counters - (char*)a.barrier), sizeof counters);
    stats.d2h_bytes += (size_t)len * sizeof(KsbResult) + (ctl_hi - ctl_lo);
    if (flags[0] != 0) {
      layout_dirty = true;  // device state is unknown: rebuild it from the host records next time
      throw std::runtime_error("device kernel stopped with error code " + std::to_string(flags[0]) +
                               " (1-4, 7: barrier wait timed out, 5: a group made no progress, 6: peer error); last stage marker " +
                               std::to_string(counters[14] >> 16) + "/" + std::to_string((counters[14] >> 8) & 0xff) + "/" +
                               std::to_string(counters[14] & 0xff) + "; first timeout at barrier " +
                               std::to_string(counters[15] >> 32) + " block " + std::to_string((counters[15] >> 8) & 0xffffff) +
                               " grid " + std::to_string(grid));
    }
    // reason histograms of the chunk's no-fit pods, appended to the batch's; slots are renumbered accordingly
    const int base = (int)(last_fiterr_hist.size() / KSB_REASON_BITS);
    const int nfe = std::min(flags[1], a.fiterr_cap);
    if (nfe > 0) {
      last_fiterr_hist.resize((size_t)(base + nfe) * KSB_REASON_BITS);
      CUDA_OK(cudaMemcpy(last_fiterr_hist.data() + (size_t)base * KSB_REASON_BITS, a.fiterr_hist, (size_t)nfe * KSB_REASON_BITS * 4,
                         cudaMemcpyDeviceToHost));
      stats.d2h_bytes += (size_t)nfe * KSB_REASON_BITS * 4;
    }
    for (int i = c0; i < c1; ++i) {
      KsbResult& r = last_results[i];
      if (r.fiterr_slot >= 0) r.fiterr_slot = r.fiterr_slot < nfe ? r.fiterr_slot + base : -1;
    }
    stats.pods_scheduled += len;
    stats.groups += counters[0];
    stats.group_restarts += counters[1];
  };

  // results to the caller; the device's assume step mirrored on the host records (scheduler.go:478, storage.go:180-187)
  auto post_chunk = [&](int c0, int c1) {
    NvtxRange r_post("AssumePod (host mirror)");
    auto tp0 = std::chrono::steady_clock::now();
    for (int i = c0; i < c1; ++i) {
      const KsbResult& r = last_results[i];
      out[i].status = r.status;
      out[i].node_index = r.row;
      out[i].num_filtered = r.num_filtered;
      out[i].score = r.score;
      for (int j = 0; j < KSB_MAX_EXT; ++j) out[i].device_mask[j] = r.device_mask[j];
      if (!assume || r.status != KSB_SCHEDULED) continue;
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
      std::string key = p.key();
      host_add_pod(p, hn, &key, last_encoded[i].sig);
      pods.insert_or_assign(std::move(key), HostPod{std::move(p)});  // the batch is the caller's scratch copy: its pods move into the cache
    }
    post_ms += std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - tp0).count();
    if (assume) batch_posted = (size_t)c1;
  };

  auto bounds = [&](int c, int* c0, int* c1) {
    *c0 = c * chunk_len;
    *c1 = std::min(np, *c0 + chunk_len);
  };
  int c0 = 0, c1 = 0, n0 = 0, n1 = 0;
  bounds(0, &c0, &c1);
  encode_chunk(c0, c1, 0);
  mark("first encode");
  launch_chunk(c0, c1, 0);
  mark("first launch");
  // Chunked batches mirror the placements of chunk c on a second host thread while this one encodes chunk c+2 (and
  // the device runs chunk c+1). The two touch disjoint state: the mirror thread writes the request / port columns
  // of the host table, the class counters' values, the per-node records and the pod cache; the encoder reads the
  // dictionaries, the class registry's static fields and node rows, and writes label columns only. Mirror tasks
  // run strictly one after another.
  std::future<void> mirror;
  auto join_mirror = [&] {
    if (mirror.valid()) mirror.get();  // rethrows what the task threw
  };
  try {
    for (int c = 0; c < n_chunks; ++c) {
      bounds(c, &c0, &c1);
      const bool more = c + 1 < n_chunks;
      if (more) {
        bounds(c + 1, &n0, &n1);
        encode_chunk(n0, n1, (c + 1) & 1);  // next to the kernel of chunk c
      }
      wait_chunk(c0, c1);
      if (more) launch_chunk(n0, n1, (c + 1) & 1);
      join_mirror();
      if (more) {
        const int m0 = c0, m1 = c1;
        mirror = std::async(std::launch::async, [&post_chunk, m0, m1] { post_chunk(m0, m1); });  // next to the kernel of chunk c+1
      } else {
        post_chunk(c0, c1);
      }
    }
  } catch (...) {
    try {
      join_mirror();
    } catch (...) {
    }
    // A limit met while encoding a later chunk (label universe, expression or pool caps) or a device error: a kernel
    // may be in flight and the device table may hold commits the host records do not. Drain the stream and have the
    // table rebuilt from the host records, which hold exactly the chunks posted so far (reported in `out`).
    cudaStreamSynchronize(stream);
    layout_dirty = true;
    have_snapshot = false;
    throw;
  }
  mark("chunks done");
  if (n_chunks > 1) {
    // the replay entry points run the whole batch in one launch: dependency indices relative to the batch
    CUDA_OK(cudaMemcpyAsync(b_dep.d, last_dep.data(), (size_t)np * 4, cudaMemcpyHostToDevice, stream));
    CUDA_OK(cudaStreamSynchronize(stream));
    batch_pipe_ok = recent_deps * 10 <= np;
  }
  for (int i = 0; i < 12; ++i) stats.stage_cycles[i] = counters[2 + i];  // of the last launch
  stats.stage_cycles[12] = 0;  // counters[14] is the progress marker of the timeout diagnostics, not a duration
  stats.last_batch_encode_ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count() -
                               kernel_ms - post_ms - sync_ms;
  if (n_chunks > 1) stats.last_batch_encode_ms = encode_ms;  // overlapped with kernels: report the work, not the wait
  stats.last_batch_sync_ms = sync_ms;
  stats.last_batch_kernel_ms = kernel_ms;
  stats.last_batch_post_ms = post_ms;
  mark("end");
  if (trace && np >= 4096) {
    fprintf(stderr, "ksb_schedule_batch %d pods:", np);
    for (auto& m : marks) fprintf(stderr, " %s %.1f ms;", m.first, m.second);
    fprintf(stderr, " kernels %.1f, encode %.1f, mirror %.1f, table sync %.1f\n", kernel_ms, encode_ms, post_ms, sync_ms);
  }
}

// Grow a device buffer and keep what it holds (the stream must be idle: the old block is freed).
void Engine::grow_keep(DevBuf& b, size_t bytes) {
  if (bytes <= b.cap) return;
  DevBuf nb;
  nb.ensure(std::max(bytes, b.cap * 2));
  if (b.d) {
    CUDA_OK(cudaMemcpyAsync(nb.d, b.d, b.cap, cudaMemcpyDeviceToDevice, stream));
    CUDA_OK(cudaStreamSynchronize(stream));
    b.release();
  }
  b = nb;
}

// ksb_resync_begin: drop every cluster record (the caller replays its cache), keep the dictionaries.
void Engine::resync_begin() {
  std::lock_guard<std::mutex> g(mu);
  CUDA_OK(cudaSetDevice(cfg.device));
  cudaStreamSynchronize(stream);  // whatever it returns: a failed launch is what brought the caller here
  cudaGetLastError();             // clear the sticky error of a failed launch, if any
  nodes.clear();
  rows.clear();
  pods.clear();
  for (auto& o : owners) o.clear();
  for (auto& kv : sigs) {
    kv.second->spread_class = -2;  // selector owners are gone: signatures forget their class
    kv.second->placed.clear();     // ... and the pods
  }
  n = 0;
  layout_dirty = true;
  have_snapshot = false;
  label_index_valid = false;
  last_encoded.clear();
  last_results.clear();
  last_fiterr_hist.clear();
}

void Engine::flush() {
  std::lock_guard<std::mutex> g(mu);
  CUDA_OK(cudaSetDevice(cfg.device));
  if (layout_dirty) rebuild_layout();
  if (n == 0) return;
  warm_dictionaries();
  sync_device();
  prepare_geometry();  // the scratch buffer for the first call
  if (cfg.shard_count <= 1) {
    b_fill_rows.ensure((size_t)n_pad * 4);  // row lists of label bits interned later
    pin_col.ensure((size_t)ln_pad * 8);
    // an empty launch of each kernel variant this table can use: module loading and the cooperative-launch set-up
    // otherwise land in the first Schedule call (1.5 - 3 ms)
    for (int p = 0; p < 2; ++p) {
      KsbBatchArgs a = make_args(0, false);
      if (p == 1 && !(pipe && grid >= 2)) break;
      a.pipeline = p;
      one_pod_bytes = 0;
      launch(a);
      CUDA_OK(cudaStreamSynchronize(stream));
    }
  }
  CUDA_OK(cudaStreamSynchronize(stream));
}

// What the first pods of a stream would otherwise pay for inside their Schedule call: the label bits any selector
// over a low-cardinality node label can mention, and the SelectorSpread class of every selector owner on its own
// (a pod matched by several owners still creates its class on first use). Bounded: at most half of the label
// universe and 1,024 classes.
void Engine::warm_dictionaries() {
  if (cfg.shard_count > 1) return;  // shared buffers keep the capacity they were exported with
  if (!label_index_valid) build_label_index();
  std::map<std::string, std::set<std::string>> values;
  for (auto& kv : rows_by_pair) {
    const size_t z = kv.first.find('\0');
    values[kv.first.substr(0, z)].insert(kv.first.substr(z + 1));
  }
  const int budget = W * 64 / 2;  // the rest stays free for what pods actually mention
  for (auto& kv : values) {
    if (kv.second.size() > 16) continue;  // hostname-like keys: one bit per value would not fit
    if ((int)label_defs.size() + (int)kv.second.size() + 1 > budget) break;
    label_bit('K', kv.first, std::string(), 0);
    for (auto& v : kv.second) label_bit('P', kv.first, v, 0);
  }
  for (auto& kv : values)
    if (kv.second.size() > 16 && (int)label_defs.size() < budget) label_bit('K', kv.first, std::string(), 0);
  for (auto& kv : rows_by_key) {  // keys with integer values: the device column Gt / Lt bits are later set from
    bool any = false;
    for (auto& rv : kv.second) any = any || rv.is_num;
    if (any) ensure_numcol(kv.first);
  }
  for (int kind = 0; kind < 4; ++kind)
    for (auto& o : owners[kind]) {
      if (spread_classes.size() >= 1024) return;
      Pod probe;  // a pod this owner (and, as far as known here, no other) selects
      probe.ns = o.ns;
      sel::Selector sx;
      if (kind < 2) {
        if (o.map_selector.empty()) continue;
        probe.labels = o.map_selector;
      } else {
        if (!sel::from_label_selector(o.selector, &sx)) continue;
        if (!o.selector.match_expressions.empty() || o.selector.match_labels.empty()) continue;
        probe.labels = o.selector.match_labels;
      }
      spread_class_of(probe, sig_of(probe));
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
  {  // stage cycles and group counts of the replayed launch (the whole batch in one launch)
    uint64_t counters[16] = {0};
    CUDA_OK(cudaMemcpy(counters, a.counters, sizeof counters, cudaMemcpyDeviceToHost));
    for (int i = 0; i < 12; ++i) stats.stage_cycles[i] = counters[2 + i];
    stats.stage_cycles[12] = 0;
  }
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
  a.n_ranks = 1;
  a.rank = 0;
  a.segs_per_shard = nseg;
  a.nseg_total = nseg;
  if (cfg.shard_count > 1) throw std::runtime_error("ksb_replay_last_batch_noalloc does not support sharded engines");
  if (ksb_noalloc_bind(cfg.device, ln_pad, zone_stride, grid, &a, err, sizeof err) != 0) throw std::runtime_error(err);
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

// ------------------------------------------------------------------ sharding
//
// One engine per GPU, each holding a contiguous range of node rows. Every engine receives every node, pod
// and selector owner (host state is replicated and stays identical); only the device buffers are sharded.
// Phases 1 and 2 run on every GPU over its own rows; the reduction targets live in rank 0's memory and are
// updated with NVLink atomics; phase 3 runs on rank 0, which reaches the other shards' stashes and node
// columns through CUDA-IPC mappings of their buffers.

namespace {
struct ShardBlob {
  uint32_t magic;
  int32_t rank, n_ranks, n, ln, ln_pad, lbase, rows_per_shard;
  uint64_t results_cap;
  cudaIpcMemHandle_t h[9];  // i64, u32, u64, dev_class, topo, spread, tables, scratch, results
};
constexpr uint32_t kShardMagic = 0x4b534231u;
}  // namespace

std::string Engine::shard_export(int max_batch_pods) {
  std::lock_guard<std::mutex> g(mu);
  CUDA_OK(cudaSetDevice(cfg.device));
  if (cfg.shard_count < 2) throw std::runtime_error("engine was not created with shard_count > 1");
  if (cfg.shard_count > KSB_MAX_RANKS) throw std::runtime_error("more than KSB_MAX_RANKS shards");
  if (shard_connected) throw std::runtime_error("shard already connected");
  // fixed capacities: shared buffers may not be reallocated once peers have mapped them
  min_spread_cols = 1024;
  min_table_bytes = (size_t)4 << 20;
  spread_cap_cols = 0;
  b_spread.release();  // re-created at the shared capacity by sync_device (an earlier ksb_flush sized it smaller)
  arena_dirty = true;
  sync_device();
  prepare_geometry();
  CUDA_OK(cudaMemsetAsync(b_scratch.d, 0, b_scratch.cap, stream));
  b_results.ensure((size_t)std::max(1, max_batch_pods) * sizeof(KsbResult));
  CUDA_OK(cudaStreamSynchronize(stream));
  ShardBlob b;
  memset(&b, 0, sizeof b);
  b.magic = kShardMagic;
  b.rank = cfg.shard_rank;
  b.n_ranks = cfg.shard_count;
  b.n = n;
  b.ln = ln;
  b.ln_pad = ln_pad;
  b.lbase = lbase;
  b.rows_per_shard = rows_per_shard;
  b.results_cap = b_results.cap;
  void* ptrs[9] = {b_i64.d, b_u32.d, b_u64.d, b_dev_class.d, b_topo.d, b_spread.d, b_tables.d, b_scratch.d, b_results.d};
  for (int i = 0; i < 9; ++i) CUDA_OK(cudaIpcGetMemHandle(&b.h[i], ptrs[i]));
  shard_results_cap = b_results.cap;
  return std::string((const char*)&b, sizeof b);
}

void Engine::shard_connect(const std::vector<std::string>& blobs) {
  std::lock_guard<std::mutex> g(mu);
  CUDA_OK(cudaSetDevice(cfg.device));
  const int R = cfg.shard_count;
  if ((int)blobs.size() != R) throw std::runtime_error("shard_connect: need one blob per rank");
  std::vector<ShardBlob> bs(R);
  for (int q = 0; q < R; ++q) {
    if (blobs[q].size() != sizeof(ShardBlob)) throw std::runtime_error("shard_connect: bad blob size");
    memcpy(&bs[q], blobs[q].data(), sizeof(ShardBlob));
    if (bs[q].magic != kShardMagic || bs[q].rank != q || bs[q].n_ranks != R || bs[q].n != n ||
        bs[q].rows_per_shard != rows_per_shard)
      throw std::runtime_error("shard_connect: blob " + std::to_string(q) + " does not describe the same cluster");
  }
  peer_maps.assign(R, PeerMap());
  if (cfg.shard_rank == 0) {
    for (int q = 1; q < R; ++q) {
      PeerMap& m = peer_maps[q];
      m.ln = bs[q].ln;
      m.ln_pad = bs[q].ln_pad;
      m.lbase = bs[q].lbase;
      for (int i = 0; i < 9; ++i) {
        CUDA_OK(cudaIpcOpenMemHandle(&m.p[i], bs[q].h[i], cudaIpcMemLazyEnablePeerAccess));
        ipc_mappings.push_back(m.p[i]);
      }
    }
  }
  // every rank maps every other rank's scratch buffer: the barrier flags and the Reduce inboxes are written all-to-all
  peer_scratch.assign(R, nullptr);
  for (int q = 0; q < R; ++q) {
    if (q == cfg.shard_rank) continue;
    if (cfg.shard_rank == 0) {
      peer_scratch[q] = peer_maps[q].p[7];
    } else {
      CUDA_OK(cudaIpcOpenMemHandle(&peer_scratch[q], bs[q].h[7], cudaIpcMemLazyEnablePeerAccess));
      ipc_mappings.push_back(peer_scratch[q]);
    }
  }
  if (cfg.shard_rank != 0) peer0_scratch = peer_scratch[0];
  shard_connected = true;
}

// FitError.Error() (generic_scheduler.go:70-88)
// FailedPredicateMap of the most recent ksb_schedule (core/generic_scheduler.go:283-362): one reason mask per node
// row, bit b = failure reason b (ksb_reason_name). Kept on the device by the launch; copied on demand.
void Engine::failed_predicates(uint32_t* out, int32_t n_out) {
  std::lock_guard<std::mutex> g(mu);
  CUDA_OK(cudaSetDevice(cfg.device));
  if (cfg.shard_count > 1) throw std::runtime_error("ksb_failed_predicates: not available on a sharded engine");
  if (last_results.size() != 1) throw std::runtime_error("ksb_failed_predicates: the last call was not a single-pod ksb_schedule");
  if (n_out < n) throw std::runtime_error("ksb_failed_predicates: buffer holds fewer entries than there are node rows");
  if (n == 0) return;
  ScratchLayout L = scratch_layout(ln_pad, zone_stride, std::max(1, nseg), group, pipe ? 2 : 1);
  CUDA_OK(cudaMemcpyAsync(out, (char*)b_scratch.d + L.reasons, (size_t)n * 4, cudaMemcpyDeviceToHost, stream));
  CUDA_OK(cudaStreamSynchronize(stream));
  stats.d2h_bytes += (size_t)n * 4;
}

// HostPriorityList of the most recent ksb_schedule (core/generic_scheduler.go:509-640): the per-row totals phase 2 left
// in the scratch buffer; rows outside the filtered list are INT64_MIN.
void Engine::scores(int64_t* out, int32_t n_out) {
  std::lock_guard<std::mutex> g(mu);
  CUDA_OK(cudaSetDevice(cfg.device));
  if (cfg.shard_count > 1) throw std::runtime_error("ksb_scores: not available on a sharded engine");
  if (last_results.size() != 1) throw std::runtime_error("ksb_scores: the last call was not a single-pod ksb_schedule");
  if (n_out < n) throw std::runtime_error("ksb_scores: buffer holds fewer entries than there are node rows");
  if (last_results[0].num_filtered < 2)
    throw std::runtime_error("ksb_scores: fewer than two nodes fitted, so nothing was prioritized (generic_scheduler.go:144-149)");
  if (n == 0) return;
  ScratchLayout L = scratch_layout(ln_pad, zone_stride, std::max(1, nseg), group, pipe ? 2 : 1);
  std::vector<int32_t> raw((size_t)n);
  CUDA_OK(cudaMemcpyAsync(raw.data(), (char*)b_scratch.d + L.score, (size_t)n * 4, cudaMemcpyDeviceToHost, stream));
  CUDA_OK(cudaStreamSynchronize(stream));
  stats.d2h_bytes += (size_t)n * 4;
  for (int i = 0; i < n; ++i) out[i] = raw[i] == INT32_MIN ? INT64_MIN : (int64_t)raw[i];
  if (getenv("KSB_SCORES_RAW")) {  // debugging aid: the raw InterPodAffinity sums (Map half) instead of the totals
    std::vector<int32_t> r4((size_t)n * 4);
    CUDA_OK(cudaMemcpy(r4.data(), (char*)b_scratch.d + L.raw, (size_t)n * 16, cudaMemcpyDeviceToHost));
    for (int i = 0; i < n; ++i) out[i] = r4[(size_t)i * 4] == INT32_MIN ? INT64_MIN : (int64_t)r4[(size_t)i * 4 + 3];
  }
}

std::string Engine::reason_name(int bit) {
  std::lock_guard<std::mutex> g(mu);
  static const char* fixed[KSB_REASON_BITS] = {
      "Insufficient pods", "Insufficient cpu", "Insufficient memory", "Insufficient alpha.kubernetes.io/nvidia-gpu",
      "Insufficient ephemeral-storage", nullptr, nullptr, nullptr, nullptr, "HostName", "PodFitsHostPorts",
      "MatchNodeSelector", "PodToleratesNodeTaints", "NodeUnderMemoryPressure", "NodeUnderDiskPressure", "NodeNotReady",
      "NodeOutOfDisk", "NodeNetworkUnavailable", "NodeUnschedulable", "MatchInterPodAffinity",
      "ExistingPodsAntiAffinityRulesNotMatch", "PodAffinityRulesNotMatch", "PodAntiAffinityRulesNotMatch", nullptr, nullptr,
      nullptr, nullptr, nullptr, "CheckNodeLabelPresence", "extender", nullptr, nullptr};
  if (bit < 0 || bit >= KSB_REASON_BITS) throw std::runtime_error("reason bit out of range");
  if (fixed[bit]) return fixed[bit];
  if (bit >= RB_SCALAR0 && bit <= RB_SCALAR3)
    return bit - RB_SCALAR0 < (int)scalar_names.size() ? "Insufficient " + scalar_names[bit - RB_SCALAR0] : std::string();
  if (!last_encoded.empty()) {  // per-pod texts of the fork's extended-resource entries
    if (bit >= RB_EXT0 && bit <= RB_EXT3) return last_encoded[0].ext_reasons[bit - RB_EXT0];
    if (bit == RB_EXT_LIMITS) return last_encoded[0].ext_reasons[KSB_MAX_EXT];
  }
  return std::string();
}

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
      nullptr, nullptr, nullptr, "CheckNodeLabelPresence", "extender", nullptr, nullptr};
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

// Schedule with scheduler extenders (core/generic_scheduler.go:333-351, :608-632). The extenders themselves are HTTP
// services the Go shim talks to (core/extender.go); what reaches the engine is their verdict for this pod: the rows their
// Filter calls kept, applied after the predicates and before the fork's device selection, and the sum over extenders of
// weight * Prioritize score per row, added to the weighted sum of the built-in priorities.
void Engine::schedule_extended(const Pod& p, const uint32_t* keep, const int32_t* scores, ksb_result* out) {
  if (cfg.shard_count > 1) throw std::runtime_error("ksb_schedule_extended: not available on a sharded engine");
  {
    std::lock_guard<std::mutex> g(mu);
    CUDA_OK(cudaSetDevice(cfg.device));
    if (layout_dirty) rebuild_layout();
    if (n == 0) throw std::runtime_error("ksb_schedule_extended: no nodes");
    const size_t words = ((size_t)n_pad + 31) / 32;
    call_ext_keep = nullptr;
    call_ext_score = nullptr;
    if (keep) {
      b_ext_keep.ensure(words * 4);
      CUDA_OK(cudaMemcpyAsync(b_ext_keep.d, keep, (((size_t)n + 31) / 32) * 4, cudaMemcpyHostToDevice, stream));
      call_ext_keep = (const uint32_t*)b_ext_keep.d;
    }
    if (scores) {
      b_ext_score.ensure((size_t)n_pad * 4);
      CUDA_OK(cudaMemcpyAsync(b_ext_score.d, scores, (size_t)n * 4, cudaMemcpyHostToDevice, stream));
      call_ext_score = (const int32_t*)b_ext_score.d;
    }
    CUDA_OK(cudaStreamSynchronize(stream));  // the caller's arrays
    stats.h2d_bytes += (keep ? (((size_t)n + 31) / 32) * 4 : 0) + (scores ? (size_t)n * 4 : 0);
  }
  std::vector<Pod> one{p};
  try {
    schedule_batch(one, out, false);
  } catch (...) {
    call_ext_keep = nullptr;
    call_ext_score = nullptr;
    throw;
  }
  call_ext_keep = nullptr;
  call_ext_score = nullptr;
}

// genericScheduler.Preempt (core/generic_scheduler.go:199-256) for a pod whose Schedule ended in a FitError. The call
// evaluates the pod once more (no assume) to have its FailedPredicateMap on the device, then runs one thread per node:
// nodesWherePreemptionMightHelp, selectVictimsOnNode (ksb_preempt_kernel), and picks the node on the host
// (pickOneNodeForPreemption over the per-row statistics, rows visited in name order where the reference walks a Go map).
// Scope, enforced here: no inter-pod affinity in play (the preemptor has no pod (anti-)affinity terms, no cached pod has
// required anti-affinity terms), no host ports on the preemptor, no PodDisruptionBudgets, no nominated pods.
PreemptOutcome Engine::preempt(const Pod& p) {
  if (cfg.shard_count > 1) throw std::runtime_error("ksb_preempt: not available on a sharded engine");
  if (p.affinity.present && (p.affinity.pod.present || p.affinity.anti.present))
    throw std::runtime_error("preemption for a pod with inter-pod (anti-)affinity terms is not implemented");
  for (auto& c : p.containers)
    for (auto& port : c.ports)
      if (port.host_port != 0) throw std::runtime_error("preemption for a pod with host ports is not implemented");
  {
    std::lock_guard<std::mutex> g(mu);
    for (auto& kv : pods) {
      const Pod& ep = kv.second.pod;
      if (ep.affinity.present && ep.affinity.anti.present && !ep.affinity.anti.required.empty())
        throw std::runtime_error("preemption next to pods with required anti-affinity terms is not implemented");
    }
  }
  // the FailedPredicateMap: one evaluation of the pod, nothing assumed (ksb_schedule)
  std::vector<Pod> one{p};
  ksb_result sr;
  schedule_batch(one, &sr, false);
  std::lock_guard<std::mutex> g(mu);
  CUDA_OK(cudaSetDevice(cfg.device));
  PreemptOutcome out;
  if (sr.status == KSB_NO_NODES) throw std::runtime_error("no nodes available to schedule pods");
  if (sr.status == KSB_SCHEDULE_ERROR) throw std::runtime_error("ksb_preempt: the pod cannot be evaluated");
  ScratchLayout L = scratch_layout(ln_pad, zone_stride, std::max(1, nseg), group, pipe ? 2 : 1);
  // cached pods per node, in NodeInfo.pods order
  std::vector<int32_t> row_start((size_t)n + 1, 0);
  std::vector<KsbVictimPod> vp;
  std::vector<const std::string*> vkey;
  for (int r = 0; r < n; ++r) {
    row_start[r] = (int32_t)vp.size();
    for (auto& key : rows[r]->pod_order) {
      auto it = pods.find(key);
      if (it == pods.end()) continue;
      const Pod& ep = it->second.pod;
      KsbVictimPod v;
      memset(&v, 0, sizeof v);
      int64_t dummy_pods = 0;
      std::map<std::string, int64_t> sc;
      for (auto& c : ep.containers) {  // calculateResource (node_info.go:421-433): containers only
        reslist_to_aggregates_public(c.requests, &v.cpu, &v.mem, &v.agpu, &v.eph, &dummy_pods, &sc);
        auto cpu = c.requests.find(kResCPU);
        auto mem = c.requests.find(kResMemory);
        v.nz_cpu += cpu == c.requests.end() ? 100 : cpu->second;
        v.nz_mem += mem == c.requests.end() ? 200LL * 1024 * 1024 : mem->second;
      }
      for (auto& kv : sc) {
        const int slot = scalar_slot(kv.first);
        v.scalar[slot] += kv.second;
      }
      v.priority = ep.priority;
      vp.push_back(v);
      vkey.push_back(&it->first);
    }
  }
  row_start[n] = (int32_t)vp.size();
  b_pre_pods.ensure(std::max<size_t>(1, vp.size()) * sizeof(KsbVictimPod));
  b_pre_rows.ensure(((size_t)n + 1) * 4);
  // outputs: state, victims, highest [n] int32 each; sum [n] int64; victim list; error flag
  const size_t o_state = 0, o_vict = (size_t)n_pad * 4, o_high = (size_t)n_pad * 8, o_sum = (size_t)n_pad * 12,
               o_list = o_sum + (size_t)n_pad * 8, o_err = o_list + (size_t)KSB_MAX_VICTIMS * 4, o_total = o_err + 16;
  b_pre_out.ensure(o_total);
  if (!vp.empty()) CUDA_OK(cudaMemcpyAsync(b_pre_pods.d, vp.data(), vp.size() * sizeof(KsbVictimPod), cudaMemcpyHostToDevice, stream));
  CUDA_OK(cudaMemcpyAsync(b_pre_rows.d, row_start.data(), row_start.size() * 4, cudaMemcpyHostToDevice, stream));
  CUDA_OK(cudaMemsetAsync((char*)b_pre_out.d + o_err, 0, 16, stream));
  stats.h2d_bytes += vp.size() * sizeof(KsbVictimPod) + row_start.size() * 4;
  const KsbPodHeader& h = last_encoded[0].h;
  KsbPreemptArgs a;
  memset(&a, 0, sizeof a);
  a.table = dT;
  a.reasons = (const uint32_t*)((char*)b_scratch.d + L.reasons);
  a.pods = (const KsbVictimPod*)b_pre_pods.d;
  a.row_start = (const int32_t*)b_pre_rows.d;
  a.req_cpu = h.req_cpu;
  a.req_mem = h.req_mem;
  a.req_agpu = h.req_agpu;
  a.req_eph = h.req_eph;
  for (int s = 0; s < KSB_MAX_SCALARS; ++s) a.scalar_req[s] = h.scalar_req[s];
  a.scalar_mask = h.scalar_mask;
  a.has_request = (h.flags & PF_HAS_REQUEST) ? 1 : 0;
  a.priority = p.priority;
  a.only_row = -1;
  char* ob = (char*)b_pre_out.d;
  a.row_state = (int32_t*)(ob + o_state);
  a.row_victims = (int32_t*)(ob + o_vict);
  a.row_highest = (int32_t*)(ob + o_high);
  a.row_sum = (int64_t*)(ob + o_sum);
  a.victim_list = (int32_t*)(ob + o_list);
  a.error_flag = (int32_t*)(ob + o_err);
  int rc = ksb_launch_preempt(&a, stream);
  if (rc != 0) throw std::runtime_error(std::string("preemption kernel failed: ") + cudaGetErrorString((cudaError_t)rc));
  stats.kernel_launches += 1;
  std::vector<int32_t> st((size_t)n), nv((size_t)n), hi((size_t)n);
  std::vector<int64_t> sum((size_t)n);
  int32_t err = 0;
  CUDA_OK(cudaMemcpyAsync(st.data(), ob + o_state, (size_t)n * 4, cudaMemcpyDeviceToHost, stream));
  CUDA_OK(cudaMemcpyAsync(nv.data(), ob + o_vict, (size_t)n * 4, cudaMemcpyDeviceToHost, stream));
  CUDA_OK(cudaMemcpyAsync(hi.data(), ob + o_high, (size_t)n * 4, cudaMemcpyDeviceToHost, stream));
  CUDA_OK(cudaMemcpyAsync(sum.data(), ob + o_sum, (size_t)n * 8, cudaMemcpyDeviceToHost, stream));
  CUDA_OK(cudaMemcpyAsync(&err, ob + o_err, 4, cudaMemcpyDeviceToHost, stream));
  CUDA_OK(cudaStreamSynchronize(stream));
  stats.d2h_bytes += (size_t)n * 20 + 4;
  if (err != 0) throw std::runtime_error("a node holds more than " + std::to_string(KSB_MAX_VICTIMS) + " lower-priority pods");
  for (int r = 0; r < n; ++r) {
    if (st[r] >= 1) out.num_potential += 1;
    if (st[r] == 2) out.num_candidates += 1;
  }
  if (out.num_potential == 0) {  // generic_scheduler.go:227-231: clean up the pod's nominated node name
    out.status = 2;
    last_preempt = out;
    return out;
  }
  // pickOneNodeForPreemption (:654-758); no PDBs, so every node has zero violations
  int pick = -1;
  for (int r = 0; r < n && pick < 0; ++r)
    if (st[r] == 2 && nv[r] == 0) pick = r;  // a node that needs no preemption
  if (pick < 0) {
    int32_t min_high = INT32_MAX;
    for (int r = 0; r < n; ++r)
      if (st[r] == 2) min_high = std::min(min_high, hi[r]);
    int64_t min_sum = INT64_MAX;
    for (int r = 0; r < n; ++r)
      if (st[r] == 2 && hi[r] == min_high) min_sum = std::min(min_sum, sum[r]);
    int32_t min_n = INT32_MAX;
    for (int r = 0; r < n; ++r)
      if (st[r] == 2 && hi[r] == min_high && sum[r] == min_sum) min_n = std::min(min_n, nv[r]);
    for (int r = 0; r < n && pick < 0; ++r)
      if (st[r] == 2 && hi[r] == min_high && sum[r] == min_sum && nv[r] == min_n) pick = r;
  }
  if (pick < 0) {
    out.status = 1;
    last_preempt = out;
    return out;
  }
  out.status = 0;
  out.row = pick;
  if (nv[pick] > 0) {  // the picked row once more, this time writing its victims
    a.only_row = pick;
    rc = ksb_launch_preempt(&a, stream);
    if (rc != 0) throw std::runtime_error(std::string("preemption kernel failed: ") + cudaGetErrorString((cudaError_t)rc));
    stats.kernel_launches += 1;
    std::vector<int32_t> list((size_t)nv[pick]);
    CUDA_OK(cudaMemcpyAsync(list.data(), ob + o_list, list.size() * 4, cudaMemcpyDeviceToHost, stream));
    CUDA_OK(cudaStreamSynchronize(stream));
    for (int32_t i : list) out.victims.push_back(*vkey[(size_t)i]);
  }
  last_preempt = out;
  return out;
}

// Cache.AssumePod / AddPod outside a batch.
void Engine::pod_add(const Pod& p) {
  std::lock_guard<std::mutex> g(mu);
  have_snapshot = false;
  if (pods.count(p.key())) throw std::runtime_error("pod " + p.key() + " is already in the cache");  // cache.go:127-129
  HostNode& hn = nodes[p.node_name];
  if (layout_dirty || !hn.has_node || hn.row < 0) {
    // rows are about to be rebuilt from the host records (or the node has no Node object yet)
    host_add_pod(p, hn);
    pods[p.key()] = HostPod{p};
    return;
  }
  if (cfg.shard_count > 1) {
    // sharded: keep the host records exact and re-upload this engine's slice before the next batch
    host_add_pod(p, hn);
    pods[p.key()] = HostPod{p};
    layout_dirty = true;
    return;
  }
  CUDA_OK(cudaSetDevice(cfg.device));
  register_pod_classes(p);
  layout_arena();
  EncodedPod e;
  try {
    encode_pod(p, &e);
  } catch (const GrowLabels&) {
    // (a pod's own selectors play no part in AddPod, but the encoder compiles them): wider label columns, then as for
    // any table that is about to be rebuilt
    layout_dirty = true;
    host_add_pod(p, hn);
    pods[p.key()] = HostPod{p};
    return;
  }
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
  have_snapshot = false;
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

// NodeInfo aggregates of one node as JSON (the fields schedulercache's tests compare: cache_test.go:30-44): requested,
// nonzero, pod count, used ports, available device ids. "null" when the cache holds no NodeInfo under that name. For a
// node that has a row on this GPU the device columns are read back and must agree with the host records.
std::string Engine::node_info_json(const std::string& name) {
  std::lock_guard<std::mutex> g(mu);
  CUDA_OK(cudaSetDevice(cfg.device));
  auto it = nodes.find(name);
  if (it == nodes.end()) return "null";
  if (it->second.has_node) {
    if (layout_dirty) rebuild_layout();
    if (n > 0) sync_device();
    CUDA_OK(cudaStreamSynchronize(stream));
  }
  it = nodes.find(name);
  const HostNode& hn = it->second;
  const int lrow = hn.row - lbase;
  if (hn.has_node && hn.row >= 0 && lrow >= 0 && lrow < ln) {
    int64_t dv[6];
    const int cols[6] = {I64_REQ_CPU, I64_REQ_MEM, I64_REQ_AGPU, I64_REQ_EPH, I64_NZ_CPU, I64_NZ_MEM};
    for (int i = 0; i < 6; ++i)
      CUDA_OK(cudaMemcpy(&dv[i], (int64_t*)b_i64.d + (size_t)cols[i] * ln_pad + lrow, 8, cudaMemcpyDeviceToHost));
    int32_t dpods = 0;
    CUDA_OK(cudaMemcpy(&dpods, (int32_t*)b_u32.d + (size_t)U32_NUM_PODS * ln_pad + lrow, 4, cudaMemcpyDeviceToHost));
    const int64_t hv[6] = {hn.req_cpu, hn.req_mem, hn.req_agpu, hn.req_eph, hn.nz_cpu, hn.nz_mem};
    for (int i = 0; i < 6; ++i)
      if (dv[i] != hv[i])
        throw std::runtime_error("node " + name + ": device column " + std::to_string(cols[i]) + " holds " + std::to_string(dv[i]) +
                                 ", host record " + std::to_string(hv[i]));
    if (dpods != (int32_t)hn.pod_keys.size())
      throw std::runtime_error("node " + name + ": device pod count " + std::to_string(dpods) + ", host record " +
                               std::to_string(hn.pod_keys.size()));
  }
  auto esc = [](const std::string& t) {
    std::string o = "\"";
    for (char c : t) {
      if (c == '"' || c == '\\') o.push_back('\\');
      o.push_back(c);
    }
    return o + "\"";
  };
  std::string out = "{\"requested\":{\"cpu\":" + std::to_string(hn.req_cpu) + ",\"memory\":" + std::to_string(hn.req_mem) +
                    ",\"gpu\":" + std::to_string(hn.req_agpu) + ",\"ephemeral\":" + std::to_string(hn.req_eph) + ",\"scalars\":{";
  bool first = true;
  for (auto& kv : hn.req_scalar) {
    if (kv.second == 0) continue;
    out += std::string(first ? "" : ",") + esc(kv.first) + ":" + std::to_string(kv.second);
    first = false;
  }
  out += "}},\"nonzero\":{\"cpu\":" + std::to_string(hn.nz_cpu) + ",\"memory\":" + std::to_string(hn.nz_mem) + "},\"pods\":" +
         std::to_string(hn.pod_keys.size()) + ",\"usedPorts\":[";
  first = true;
  for (auto& p : hn.used_ports) {
    out += std::string(first ? "" : ",") + esc(p);
    first = false;
  }
  {  // allocatable as the row holds it; none once the Node object is gone
    const bool live = hn.has_node && hn.row >= 0;
    out += "],\"allocatable\":{\"cpu\":" + std::to_string(live ? i64(I64_ALLOC_CPU, hn.row) : 0) + ",\"memory\":" +
           std::to_string(live ? i64(I64_ALLOC_MEM, hn.row) : 0) + ",\"pods\":" +
           std::to_string(live ? (int64_t)u32(U32_ALLOC_PODS, hn.row) : 0) + ",\"scalars\":{";
    first = true;
    if (live) {
      std::map<std::string, int64_t> named;
      for (size_t k = 0; k < scalar_names.size() && k < KSB_MAX_SCALARS; ++k) {
        const int64_t v = i64(I64_ALLOC_SCALAR + (int)k, hn.row);
        if (v != 0 || hn.node.allocatable.count(scalar_names[k])) named[scalar_names[k]] = v;
      }
      for (auto& kv : named) {
        out += std::string(first ? "" : ",") + esc(kv.first) + ":" + std::to_string(kv.second);
        first = false;
      }
    }
    out += "}}";
  }
  out += ",\"available\":{";
  first = true;
  for (auto& kv : hn.ext) {
    if (!kv.second.domain_present) continue;
    out += std::string(first ? "" : ",") + esc(kv.first) + ":[";
    first = false;
    bool f2 = true;
    for (auto& id : kv.second.available) {
      out += std::string(f2 ? "" : ",") + esc(id);
      f2 = false;
    }
    out += "]";
  }
  out += "}}";
  return out;
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
