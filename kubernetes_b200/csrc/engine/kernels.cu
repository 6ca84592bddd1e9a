// sm_100a kernels of the scheduling engine.
//
// One persistent cooperative kernel walks a batch of pods in FIFO order and, per pod, does
// what genericScheduler.Schedule does (reference, paths under plugin/pkg/scheduler):
//   phase 1  findNodesThatFit   core/generic_scheduler.go:283-362  -> eval_node (all predicates,
//            the fork's device matching core/extended_resources.go:83-150, and the Map half of
//            every priority) + grid-wide reductions of the Reduce inputs
//   phase 2  PrioritizeNodes    core/generic_scheduler.go:509-640  -> final integer score per node
//            + per-block (max, tie count)
//   phase 3  selectHost         core/generic_scheduler.go:177-192  -> round-robin pick among the
//            max-score ties in descending name (= row) order, then the assume step
//            (scheduler.go:478 -> schedulercache/node_info.go:329-356) applied to the one node
//            column that changed, so the next pod sees it exactly as the reference would.
//
// The path is integer/bit work over a structure-of-arrays node table: one thread per node,
// every column read is a coalesced 4/8-byte load, the pod's requirement block is staged in
// shared memory, reductions are warp shuffles (redux) + one atomic per block. No tensor cores:
// nothing here is a dense contraction. float64 priorities use __d*_rn so no FMA is ever formed
// (Go on amd64 never fuses; the file is also compiled with -fmad=false).
#include <cooperative_groups.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#include "device_types.h"
#include "kernels.h"

namespace {

constexpr int kThreads = 256;
constexpr int kMaxPoolWords = 1024;  // pool words one pod may use (checked on the host)

struct PodView {
  const KsbPodHeader* h;
  const uint64_t* exprs;
  const uint64_t* ports;
  const uint64_t* extcls;
  const uint64_t* checks;
  const uint64_t* score;
  const uint64_t* commits;
  const uint64_t* spread_inc;
};

__device__ __forceinline__ PodView make_view(const KsbPodHeader* h, const uint64_t* pool, int W) {
  PodView v;
  v.h = h;
  const uint64_t* p = pool;
  v.exprs = p;
  p += (int)h->n_exprs * W;
  v.ports = p;
  if (h->flags & PF_HAS_PORTS) p += 6 * KSB_PORT_WORDS;
  v.extcls = p;
  p += (int)h->n_ext * KSB_CLASS_WORDS;
  v.checks = p;
  p += h->n_checks;
  v.score = p;
  p += 2 * (int)h->n_score;
  v.commits = p;
  p += 2 * (int)h->n_commit;
  v.spread_inc = p;
  return v;
}

__device__ __forceinline__ int pool_words(const KsbPodHeader* h, int W) {
  return (int)h->n_exprs * W + ((h->flags & PF_HAS_PORTS) ? 6 * KSB_PORT_WORDS : 0) + (int)h->n_ext * KSB_CLASS_WORDS +
         h->n_checks + 2 * (int)h->n_score + 2 * (int)h->n_commit + h->n_spread_inc;
}

struct EvalOut {
  uint32_t reasons;
  int32_t base;  // w_lr*LeastRequested + w_bra*Balanced + w_avoid*PreferAvoid
  int32_t na;    // NodeAffinity raw count
  int32_t tt;    // TaintToleration raw count
  uint32_t sc;   // SelectorSpread raw count
  int32_t ipa;   // InterPodAffinity raw sum
  uint32_t zone;
};

// least_requested.go:50-61
__device__ __forceinline__ int64_t unused_score(int64_t requested, int64_t capacity) {
  if (capacity == 0) return 0;
  if (requested > capacity) return 0;
  return ((capacity - requested) * 10) / capacity;
}

// The fork's per-node device selection (core/extended_resources.go:83-150), devices visited in
// ascending id order (= bit order). Returns the failing reason bit or 0; fills picks[] on success.
__device__ __forceinline__ uint32_t select_devices(const KsbNodeTable& T, int r, const PodView& pv, uint32_t nflags,
                                                   uint32_t* picks) {
  const KsbPodHeader* h = pv.h;
  uint32_t avail[KSB_MAX_RES];
  bool loaded[KSB_MAX_RES];
#pragma unroll
  for (int s = 0; s < KSB_MAX_RES; ++s) loaded[s] = false;
  for (int i = 0; i < h->n_ext; ++i) {
    if ((h->flags & PF_EXT_LIMITS_ERROR) && i == h->ext_err_entry) return 1u << RB_EXT_LIMITS;
    int rs = h->ext[i].rslot;
    int num = h->ext[i].num;
    if (rs == 255 || !(nflags & (256u << rs))) return 1u << (RB_EXT0 + i);  // available[rName] missing
    if (!loaded[rs]) {
      avail[rs] = T.dev_avail[(size_t)rs * T.n_pad + r];
      loaded[rs] = true;
    }
    uint32_t a = avail[rs];
    uint32_t sel = 0;
    int got = 0;
    if (num > 0) {
      const uint64_t* cm = pv.extcls + i * KSB_CLASS_WORDS;
      const uint4* cls4 = reinterpret_cast<const uint4*>(T.dev_class + ((size_t)rs * T.n_pad + r) * KSB_MAX_DEV);
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        uint4 v = cls4[q];
        uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
        for (int j = 0; j < 4; ++j) {
#pragma unroll
          for (int half = 0; half < 2; ++half) {
            int d = q * 8 + j * 2 + half;
            uint32_t c = (w[j] >> (16 * half)) & 0xffffu;
            bool ok = ((a >> d) & 1u) && ((cm[c >> 6] >> (c & 63)) & 1ull) && got < num;
            if (ok) {
              sel |= 1u << d;
              ++got;
            }
          }
        }
      }
    }
    if (got != num) return 1u << (RB_EXT0 + i);
    picks[i] = sel;
    avail[rs] = a & ~sel;
  }
  return 0;
}

__device__ __forceinline__ bool eval_expr(const uint64_t* lab, const uint64_t* mask, uint32_t kind, int W) {
  bool any = false, all = true;
  for (int w = 0; w < W; ++w) {
    uint64_t m = mask[w];
    uint64_t x = lab[w] & m;
    any |= (x != 0);
    all &= (x == m);
  }
  return kind == EK_ANY ? any : (kind == EK_NONE ? !any : all);
}

// Everything the reference computes for one (pod, node): podFitsOnNode's predicates
// (core/generic_scheduler.go:402-501) and the Map functions of PrioritizeNodes (:557-574).
__device__ __forceinline__ void eval_node(const KsbNodeTable& T, int r, const PodView& pv, EvalOut& o,
                                          uint32_t* picks) {
  const KsbPodHeader* h = pv.h;
  const size_t np = T.n_pad;
  uint32_t reasons = 0;
  const uint32_t pf = h->flags;

  // --- PodFitsResources (predicates.go:630-687)
  const int32_t num_pods = T.num_pods[r];
  const int32_t alloc_pods = T.alloc_pods[r];
  const int64_t acpu = T.alloc_cpu[r], amem = T.alloc_mem[r];
  const int64_t ncpu = T.nz_cpu[r], nmem = T.nz_mem[r];
  const uint32_t nflags = T.flags[r];
  if (num_pods + 1 > alloc_pods) reasons |= 1u << RB_PODS;
  if (pf & PF_HAS_REQUEST) {
    if (acpu < h->req_cpu + T.req_cpu[r]) reasons |= 1u << RB_CPU;
    if (amem < h->req_mem + T.req_mem[r]) reasons |= 1u << RB_MEM;
    if (T.alloc_agpu[r] < h->req_agpu + T.req_agpu[r]) reasons |= 1u << RB_AGPU;
    if (T.alloc_eph[r] < h->req_eph + T.req_eph[r]) reasons |= 1u << RB_EPH;
    uint32_t sm = h->scalar_mask;
#pragma unroll
    for (int s = 0; s < KSB_MAX_SCALARS; ++s)
      if ((sm >> s) & 1u)
        if (T.alloc_scalar[s * np + r] < h->scalar_req[s] + T.req_scalar[s * np + r]) reasons |= 1u << (RB_SCALAR0 + s);
  }
  // --- PodFitsHost (predicates.go:764-776)
  if (h->pinned_row != -1 && h->pinned_row != T.row_base + r) reasons |= 1u << RB_HOSTNAME;
  // --- PodFitsHostPorts (predicates.go:925-945, utils.go:161-202)
  if (pf & PF_HAS_PORTS) {
    uint64_t hit = 0;
#pragma unroll
    for (int w = 0; w < KSB_PORT_WORDS; ++w) {
      hit |= T.port_any[w * np + r] & pv.ports[w];
      hit |= T.port_wild[w * np + r] & pv.ports[KSB_PORT_WORDS + w];
      hit |= T.port_exact[w * np + r] & pv.ports[2 * KSB_PORT_WORDS + w];
    }
    if (hit) reasons |= 1u << RB_PORTS;
  }
  // --- PodMatchNodeSelector (predicates.go:691-761) and NodeAffinityPriority Map (node_affinity.go:34-75)
  int32_t na = 0;
  if (h->n_exprs) {
    uint64_t lab[16];
    const int W = T.W;
    for (int w = 0; w < W; ++w) lab[w] = T.labels[w * np + r];
    int e = 0;
    bool match = true;
    if (pf & PF_NODE_SELECTOR) {
      match = eval_expr(lab, pv.exprs, (h->expr_kind[0]) & 3u, W);
      e = 1;
    }
    if (pf & PF_HAS_REQUIRED) {
      bool any = false;
      for (int t = 0; t < h->n_req_terms; ++t) {
        bool all = true;
        int end = h->term_end[t];
        for (; e < end; ++e) all &= eval_expr(lab, pv.exprs + e * W, (h->expr_kind[e >> 4] >> ((e & 15) * 2)) & 3u, W);
        any |= all;
      }
      match &= any;
    }
    if (!match) reasons |= 1u << RB_SELECTOR;
    for (int t = 0; t < h->n_pref; ++t) {
      bool all = true;
      int end = h->pref_end[t];
      for (; e < end; ++e) all &= eval_expr(lab, pv.exprs + e * W, (h->expr_kind[e >> 4] >> ((e & 15) * 2)) & 3u, W);
      if (all) na += h->pref_weight[t];
    }
  } else if ((pf & PF_HAS_REQUIRED)) {
    reasons |= 1u << RB_SELECTOR;  // required affinity with no usable term matches nothing
  }
  // --- PodToleratesNodeTaints (predicates.go:1349-1373) and TaintTolerationPriority Map (taint_toleration.go:55-73)
  if (T.taint_hard[r] & h->hard_untol) reasons |= 1u << RB_TAINTS;
  const int32_t tt = __popcll(T.taint_soft[r] & h->soft_untol);
  // --- pressure / conditions (predicates.go:1382-1441)
  if ((pf & PF_BEST_EFFORT) && (nflags & NF_MEM_PRESSURE)) reasons |= 1u << RB_MEM_PRESSURE;
  if (nflags & NF_DISK_PRESSURE) reasons |= 1u << RB_DISK_PRESSURE;
  if (nflags & NF_NOT_READY) reasons |= 1u << RB_NOT_READY;
  if (nflags & NF_OUT_OF_DISK) reasons |= 1u << RB_OUT_OF_DISK;
  if (nflags & NF_NET_UNAVAILABLE) reasons |= 1u << RB_NET_UNAVAILABLE;
  if (nflags & NF_UNSCHEDULABLE) reasons |= 1u << RB_UNSCHEDULABLE;
  // --- MatchInterPodAffinity (predicates.go:1047-1346) over per-(term class, topology domain) counts
  for (int i = 0; i < h->n_checks; ++i) {
    uint64_t c = pv.checks[i];
    uint32_t off = (uint32_t)c;
    uint32_t tk = (uint32_t)(c >> 32) & 0xffu;
    uint32_t mode = (uint32_t)(c >> 40) & 0xffu;
    uint32_t self = (uint32_t)(c >> 48) & 1u;
    uint32_t d = mode == CM_FAIL_ALWAYS ? 0u : T.topo[tk * np + r];
    bool in_domain = d != 0 && T.tables[off + d] > 0;
    uint32_t bad = 0;
    if (mode == CM_FORBID_EXISTING) {
      if (in_domain) bad = 1u << RB_IPA_EXISTING_ANTI;
    } else if (mode == CM_REQUIRE) {
      if (!in_domain) {
        bool exists = T.tables[off] > 0;  // slot 0: matching pods anywhere
        if (exists || !self) bad = 1u << RB_IPA_AFFINITY;
      }
    } else if (mode == CM_FORBID_OWN) {
      if (in_domain) bad = 1u << RB_IPA_ANTI;
    } else {
      bad = 1u << RB_IPA_ANTI;
    }
    if (bad) {
      reasons |= bad | (1u << RB_IPA);
      break;  // the reference reports the first failing rule only
    }
  }
  // --- fork: device selection runs only on nodes that passed every predicate (generic_scheduler.go:357)
  if (reasons == 0 && h->n_ext) {
    uint32_t local[KSB_MAX_EXT];
    uint32_t bad = select_devices(T, r, pv, nflags, picks ? picks : local);
    reasons |= bad;
  }

  // --- priorities (Map halves)
  const int64_t cpu = h->nz_cpu + ncpu;
  const int64_t mem = h->nz_mem + nmem;
  const int32_t lr = (int32_t)((unused_score(cpu, acpu) + unused_score(mem, amem)) / 2);  // least_requested.go:63-91
  // balanced_resource_allocation.go:51-100
  const double cf = acpu == 0 ? 1.0 : __ddiv_rn((double)cpu, (double)acpu);
  const double mf = amem == 0 ? 1.0 : __ddiv_rn((double)mem, (double)amem);
  int32_t bra = 0;
  if (!(cf >= 1.0 || mf >= 1.0)) {
    double diff = fabs(__dsub_rn(cf, mf));
    bra = (int32_t)(__dmul_rn(__dsub_rn(1.0, diff), 10.0));
  }
  const int32_t av = (T.avoid[r] & h->avoid_mask) ? 0 : 10;  // node_prefer_avoid_pods.go:29-59
  o.base = T.w_lr * lr + T.w_bra * bra + T.w_avoid * av;
  o.na = na;
  o.tt = tt;
  o.sc = h->spread_col >= 0 ? T.spread[(size_t)h->spread_col * np + r] : 0u;  // selector_spreading.go:64-113
  int64_t ipa = 0;  // interpod_affinity.go:121-216
  for (int i = 0; i < h->n_score; ++i) {
    uint64_t c = pv.score[2 * i];
    uint32_t d = T.topo[((uint32_t)(c >> 32) & 0xffu) * np + r];
    if (d) ipa += (int64_t)pv.score[2 * i + 1] * (int64_t)T.tables[(uint32_t)c + d];
  }
  o.ipa = (int32_t)ipa;
  o.zone = T.zone[r];
  o.reasons = reasons;
}

// PrioritizeNodes' Reduce functions and weighted sum (generic_scheduler.go:575-606).
__device__ __forceinline__ int32_t final_score(const KsbNodeTable& T, const KsbPodHeader* h, const KsbNorms& nm,
                                               const int32_t* zone_sum, int32_t max_zone, int32_t base, int32_t na,
                                               int32_t tt, uint32_t sc, int32_t ipa, uint32_t zone) {
  // selector_spreading.go:119-174
  double f = 10.0;
  if (nm.max_sc > 0) f = __dmul_rn(10.0, __ddiv_rn((double)(nm.max_sc - (int32_t)sc), (double)nm.max_sc));
  if (nm.have_zones && zone != 0) {
    double zs = 10.0;
    if (max_zone > 0) zs = __dmul_rn(10.0, __ddiv_rn((double)(max_zone - zone_sum[zone]), (double)max_zone));
    const double zw = 2.0 / 3.0;
    f = __dadd_rn(__dmul_rn(f, __dsub_rn(1.0, zw)), __dmul_rn(zw, zs));
  }
  int32_t ss = (int32_t)f;
  // reduce.go:29-63
  int32_t nas = nm.max_na > 0 ? (10 * na) / nm.max_na : na;
  int32_t tts = nm.max_tt > 0 ? 10 - (10 * tt) / nm.max_tt : 10;
  // interpod_affinity.go:218-237
  int32_t ipas = 0;
  if (nm.ipa_max - nm.ipa_min > 0)
    ipas = (int32_t)__dmul_rn(10.0, __ddiv_rn((double)(ipa - nm.ipa_min), (double)(nm.ipa_max - nm.ipa_min)));
  return base + T.w_ss * ss + T.w_na * nas + T.w_tt * tts + T.w_ipa * ipas;
}

__device__ __forceinline__ uint32_t ld_acquire(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}

// Grid-wide barrier on a monotonically increasing counter (cooperative launch guarantees that
// all blocks are co-resident). A bounded spin turns a lost block into an error, not a hang.
__device__ __forceinline__ void grid_barrier(KsbBatchArgs& A, uint32_t& epoch) {
  __syncthreads();
  if (threadIdx.x == 0) {
    epoch += gridDim.x;
    __threadfence();
    atomicAdd(A.barrier, 1u);
    uint32_t spins = 0;
    long long t0 = 0;
    while ((int32_t)(ld_acquire(A.barrier) - epoch) < 0) {
      if ((++spins & 1023u) == 0) {
        long long now = clock64();
        if (t0 == 0) t0 = now;
        if (now - t0 > 6000000000ll || ld_acquire((const uint32_t*)A.error_flag) != 0) {  // ~3 s
          atomicExch(A.error_flag, 1);
          break;
        }
      }
    }
  }
  __syncthreads();
}

// The assume step on the chosen node: NodeInfo.AddPod (schedulercache/node_info.go:329-356) plus
// the per-class counters that stand in for the node's pod list. One warp, lane-parallel where it can be.
__device__ __forceinline__ void commit_pod(const KsbNodeTable& T, int r, const PodView& pv, const uint32_t* picks) {
  const KsbPodHeader* h = pv.h;
  const size_t np = T.n_pad;
  const int lane = threadIdx.x & 31;
  if (lane == 0) {
    T.req_cpu[r] += h->add_cpu;
    T.req_mem[r] += h->add_mem;
    T.req_agpu[r] += h->add_agpu;
    T.req_eph[r] += h->add_eph;
    T.nz_cpu[r] += h->nz_cpu;
    T.nz_mem[r] += h->nz_mem;
    T.num_pods[r] += 1;
  }
  if (lane < KSB_MAX_SCALARS && h->scalar_add[lane] != 0) T.req_scalar[lane * np + r] += h->scalar_add[lane];
  if ((h->flags & PF_HAS_PORTS) && lane < KSB_PORT_WORDS) {
    T.port_any[lane * np + r] |= pv.ports[3 * KSB_PORT_WORDS + lane];
    T.port_wild[lane * np + r] |= pv.ports[4 * KSB_PORT_WORDS + lane];
    T.port_exact[lane * np + r] |= pv.ports[5 * KSB_PORT_WORDS + lane];
  }
  if (lane == 1 && h->n_ext) {
    // bind echo (pkg/registry/core/pod/storage/storage.go:180-187): the assigned devices leave `available`
    uint32_t take[KSB_MAX_RES];
#pragma unroll
    for (int s = 0; s < KSB_MAX_RES; ++s) take[s] = 0;
    for (int i = 0; i < h->n_ext; ++i)
      if (h->ext[i].rslot < KSB_MAX_RES) take[h->ext[i].rslot] |= picks[i];
#pragma unroll
    for (int s = 0; s < KSB_MAX_RES; ++s)
      if (take[s]) T.dev_avail[s * np + r] &= ~take[s];
  }
  for (int i = lane; i < h->n_spread_inc; i += 32) T.spread[(size_t)pv.spread_inc[i] * np + r] += 1u;
  for (int i = lane; i < h->n_commit; i += 32) {
    uint64_t c = pv.commits[2 * i];
    int32_t delta = (int32_t)(int64_t)pv.commits[2 * i + 1];
    uint32_t off = (uint32_t)c;
    uint32_t d = T.topo[((uint32_t)(c >> 32) & 0xffu) * np + r];
    if (d) T.tables[off + d] += delta;
    if ((c >> 40) & 1ull) T.tables[off] += delta;  // slot 0: total matching pods anywhere
  }
}

__device__ __forceinline__ int32_t warp_max(int32_t v) { return __reduce_max_sync(0xffffffffu, v); }
__device__ __forceinline__ int32_t warp_min(int32_t v) { return __reduce_min_sync(0xffffffffu, v); }
__device__ __forceinline__ int32_t warp_sum(int32_t v) { return __reduce_add_sync(0xffffffffu, v); }

}  // namespace

// Shared-memory layout (dynamic): pod header | pool words | zone sums.
__global__ void __launch_bounds__(kThreads) ksb_schedule_batch_kernel(KsbBatchArgs A) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  KsbPodHeader* sh = reinterpret_cast<KsbPodHeader*>(smem_raw);
  uint64_t* spool = reinterpret_cast<uint64_t*>(smem_raw + sizeof(KsbPodHeader));
  int32_t* szone = reinterpret_cast<int32_t*>(spool + kMaxPoolWords);
  __shared__ int32_t s_red[8];       // fit, max_na, max_tt, max_sc, ipa_max, ipa_min, have_zones, one_row
  __shared__ int32_t s_bmax, s_bcnt;
  __shared__ int32_t s_sel[4];       // target block, remaining index, chosen local row, scratch
  __shared__ uint32_t s_picks[KSB_MAX_EXT];
  __shared__ int32_t s_warp[kThreads / 32];

  const KsbNodeTable& T = A.table;
  const int tid = threadIdx.x;
  const int lane = tid & 31;
  const int warp = tid >> 5;
  const int nblk = gridDim.x;
  const int b = blockIdx.x;
  // contiguous row range of this block, in ascending name order
  const int rows_per_block = (T.n + nblk - 1) / nblk;
  const int row0 = b * rows_per_block;
  const int row1 = min(T.n, row0 + rows_per_block);
  const int nz = T.n_zones;
  uint32_t epoch = 0;
  uint64_t last_index = *A.last_node_index;

  for (int p = 0; p < A.n_pods; ++p) {
    // ---- stage the pod requirement block
    {
      const uint4* src = reinterpret_cast<const uint4*>(A.headers + p);
      uint4* dst = reinterpret_cast<uint4*>(sh);
      for (int i = tid; i < (int)(sizeof(KsbPodHeader) / 16); i += kThreads) dst[i] = src[i];
      __syncthreads();
      const int pw = pool_words(sh, T.W);
      const uint64_t* psrc = A.pool + sh->pool_off;
      for (int i = tid; i < pw; i += kThreads) spool[i] = psrc[i];
      for (int i = tid; i < nz; i += kThreads) szone[i] = 0;
      if (tid < 8) s_red[tid] = (tid == 5) ? 0 : ((tid == 7) ? -1 : 0);
      __syncthreads();
    }
    const PodView pv = make_view(sh, spool, T.W);
    KsbNorms* norms = A.norms + (p & 1);
    int32_t* zsum = A.zone_sum + (size_t)(p & 1) * A.zone_stride;
    const bool whole_pod_error = (sh->flags & PF_SCHED_ERROR) != 0;

    // ---- phase 1: predicates + Map on every node of this block
    if (!whole_pod_error) {
      int32_t fit = 0, mna = 0, mtt = 0, msc = 0, imax = 0, imin = 0, hz = 0, one = -1;
      for (int r = row0 + tid; r < row1; r += kThreads) {
        EvalOut o;
        eval_node(T, r, pv, o, nullptr);
        int4 raw;
        if (o.reasons == 0) {
          fit += 1;
          mna = max(mna, o.na);
          mtt = max(mtt, o.tt);
          msc = max(msc, (int32_t)o.sc);
          imax = max(imax, o.ipa);
          imin = min(imin, o.ipa);
          one = max(one, T.row_base + r);
          if (o.zone) {
            hz = 1;
            if (o.sc) atomicAdd(&szone[o.zone], (int32_t)o.sc);
          }
          raw = make_int4(o.base, o.na, (o.tt << 24) | (int32_t)(o.sc & 0xffffffu), o.ipa);
        } else {
          raw = make_int4(INT32_MIN, 0, 0, 0);
        }
        A.raw[r] = raw;
        A.reasons[r] = o.reasons;
      }
      fit = warp_sum(fit);
      mna = warp_max(mna);
      mtt = warp_max(mtt);
      msc = warp_max(msc);
      imax = warp_max(imax);
      imin = warp_min(imin);
      hz = warp_max(hz);
      one = warp_max(one);
      if (lane == 0) {
        if (fit) atomicAdd(&s_red[0], fit);
        if (mna) atomicMax(&s_red[1], mna);
        if (mtt) atomicMax(&s_red[2], mtt);
        if (msc) atomicMax(&s_red[3], msc);
        if (imax) atomicMax(&s_red[4], imax);
        if (imin) atomicMin(&s_red[5], imin);
        if (hz) atomicMax(&s_red[6], hz);
        if (one >= 0) atomicMax(&s_red[7], one);
      }
      __syncthreads();
      if (tid == 0) {
        if (s_red[0]) atomicAdd(&norms->fit_count, s_red[0]);
        if (s_red[1]) atomicMax(&norms->max_na, s_red[1]);
        if (s_red[2]) atomicMax(&norms->max_tt, s_red[2]);
        if (s_red[3]) atomicMax(&norms->max_sc, s_red[3]);
        if (s_red[4]) atomicMax(&norms->ipa_max, s_red[4]);
        if (s_red[5]) atomicMin(&norms->ipa_min, s_red[5]);
        if (s_red[6]) atomicMax(&norms->have_zones, 1);
        if (s_red[7] >= 0) atomicMax(&norms->one_row, s_red[7]);
      }
      for (int z = tid; z < nz; z += kThreads)
        if (szone[z]) atomicAdd(&zsum[z], szone[z]);
    }
    grid_barrier(A, epoch);

    // ---- phase 2: Reduce + weighted sum, per-block (max, tie count)
    KsbNorms nm;
    {
      const int32_t* np32 = reinterpret_cast<const int32_t*>(norms);
      nm.fit_count = __ldcg(np32 + 0);
      nm.max_na = __ldcg(np32 + 1);
      nm.max_tt = __ldcg(np32 + 2);
      nm.max_sc = __ldcg(np32 + 3);
      nm.ipa_max = __ldcg(np32 + 4);
      nm.ipa_min = __ldcg(np32 + 5);
      nm.have_zones = __ldcg(np32 + 6);
      nm.one_row = __ldcg(np32 + 7);
    }
    const bool prio_needed = !whole_pod_error && nm.fit_count >= 2 && !(sh->flags & PF_PRIO_ERROR);
    if (prio_needed) {
      int32_t mz = 0;
      for (int z = tid; z < nz; z += kThreads) {
        int32_t v = __ldcg(zsum + z);
        szone[z] = v;
        mz = max(mz, v);
      }
      mz = warp_max(mz);
      if (lane == 0) s_warp[warp] = mz;
      __syncthreads();
      int32_t max_zone = 0;
#pragma unroll
      for (int w = 0; w < kThreads / 32; ++w) max_zone = max(max_zone, s_warp[w]);
      int32_t bmax = INT32_MIN, bcnt = 0;
      for (int r = row0 + tid; r < row1; r += kThreads) {
        int4 raw = A.raw[r];
        int32_t s = INT32_MIN;
        if (raw.x != INT32_MIN) {
          s = final_score(T, sh, nm, szone, max_zone, raw.x, raw.y, (int32_t)((uint32_t)raw.z >> 24),
                          (uint32_t)raw.z & 0xffffffu, raw.w, T.zone[r]);
          if (s > bmax) {
            bmax = s;
            bcnt = 1;
          } else if (s == bmax) {
            bcnt++;
          }
        }
        A.score[r] = s;
      }
      int32_t wm = warp_max(bmax);
      int32_t wc = warp_sum(bmax == wm && bmax != INT32_MIN ? bcnt : 0);
      __syncthreads();
      if (lane == 0) {
        s_warp[warp] = wm;
      }
      __syncthreads();
      int32_t blockmax = INT32_MIN;
#pragma unroll
      for (int w = 0; w < kThreads / 32; ++w) blockmax = max(blockmax, s_warp[w]);
      if (tid == 0) s_bcnt = 0;
      __syncthreads();
      if (lane == 0 && wm == blockmax && wm != INT32_MIN) atomicAdd(&s_bcnt, wc);
      __syncthreads();
      if (tid == 0) {
        A.blk_max[b] = blockmax;
        A.blk_cnt[b] = s_bcnt;
      }
    }
    grid_barrier(A, epoch);

    // ---- phase 3: selectHost + assume, block 0 only (other blocks go straight to the barrier)
    if (b == 0) {
      KsbResult res;
      res.status = 0;
      res.row = -1;
      res.num_filtered = nm.fit_count;
      res.fiterr_slot = -1;
      res.score = 0;
#pragma unroll
      for (int i = 0; i < KSB_MAX_EXT; ++i) res.device_mask[i] = 0;
      int chosen = -1;  // global row
      if (whole_pod_error || (nm.fit_count >= 2 && (sh->flags & PF_PRIO_ERROR))) {
        res.status = 3;
      } else if (nm.fit_count == 0) {
        res.status = 2;
        // FitError: histogram of failure reasons over all nodes (generic_scheduler.go:70-88)
        __shared__ int32_t s_hist[KSB_REASON_BITS];
        __shared__ int32_t s_slot;
        if (tid < KSB_REASON_BITS) s_hist[tid] = 0;
        if (tid == 0) s_slot = atomicAdd(A.fiterr_count, 1);
        __syncthreads();
        if (s_slot < KSB_MAX_FITERR) {
          for (int r = tid; r < T.n; r += kThreads) {
            uint32_t m = __ldcg(A.reasons + r);
            while (m) {
              int bit = __ffs(m) - 1;
              m &= m - 1;
              atomicAdd(&s_hist[bit], 1);
            }
          }
          __syncthreads();
          if (tid < KSB_REASON_BITS) A.fiterr_hist[(size_t)s_slot * KSB_REASON_BITS + tid] = s_hist[tid];
          res.fiterr_slot = s_slot;
        }
      } else if (nm.fit_count == 1) {
        chosen = nm.one_row;  // generic_scheduler.go:144-146: no selectHost, counter untouched
      } else {
        // global max and tie count over blocks; ties are ordered by descending row
        int32_t gmax = INT32_MIN;
        for (int i = tid; i < nblk; i += kThreads) gmax = max(gmax, __ldcg(A.blk_max + i));
        gmax = warp_max(gmax);
        if (lane == 0) s_warp[warp] = gmax;
        __syncthreads();
        gmax = INT32_MIN;
#pragma unroll
        for (int w = 0; w < kThreads / 32; ++w) gmax = max(gmax, s_warp[w]);
        __syncthreads();
        if (tid == 0) {
          int64_t k = 0;
          for (int i = 0; i < nblk; ++i)
            if (__ldcg(A.blk_max + i) == gmax) k += __ldcg(A.blk_cnt + i);
          uint64_t ix = k > 0 ? last_index % (uint64_t)k : 0;
          int tb = -1;
          for (int i = nblk - 1; i >= 0; --i) {
            if (__ldcg(A.blk_max + i) != gmax) continue;
            uint64_t c = (uint64_t)__ldcg(A.blk_cnt + i);
            if (ix < c) {
              tb = i;
              break;
            }
            ix -= c;
          }
          s_sel[0] = tb;
          s_sel[1] = (int32_t)ix;
          s_sel[2] = -1;
        }
        __syncthreads();
        const int tb = s_sel[0];
        int remaining = s_sel[1];
        // rank-select inside the target block, scanning rows downward in chunks of kThreads
        const int trow0 = tb * rows_per_block;
        const int trow1 = min(T.n, trow0 + rows_per_block);
        for (int top = trow1; top > trow0 && s_sel[2] < 0; top -= kThreads) {
          int r = top - 1 - tid;  // thread 0 looks at the highest row
          bool tie = r >= trow0 && __ldcg(A.score + r) == gmax;
          uint32_t bal = __ballot_sync(0xffffffffu, tie);
          if (lane == 0) s_warp[warp] = __popc(bal);
          __syncthreads();
          int before = 0;
          for (int w = 0; w < warp; ++w) before += s_warp[w];
          int total = 0;
#pragma unroll
          for (int w = 0; w < kThreads / 32; ++w) total += s_warp[w];
          int my = before + __popc(bal & ((1u << lane) - 1u));
          if (tie && my == remaining) s_sel[2] = r;
          __syncthreads();
          remaining -= total;
          if (tid == 0) s_sel[1] = remaining;
          __syncthreads();
        }
        if (s_sel[2] >= 0) {
          chosen = T.row_base + s_sel[2];
          res.score = gmax;
        } else {
          res.status = 3;  // only reachable after a barrier timeout
        }
        last_index += 1;
      }
      if (chosen >= 0) {
        const int r = chosen - T.row_base;
        if (warp == 0) {
          if (lane == 0 && sh->n_ext) {
            EvalOut o;
            uint32_t picks[KSB_MAX_EXT] = {0, 0, 0, 0};
            eval_node(T, r, pv, o, picks);
#pragma unroll
            for (int i = 0; i < KSB_MAX_EXT; ++i) s_picks[i] = picks[i];
          }
          __syncwarp();
          if (A.commit) commit_pod(T, r, pv, s_picks);
        }
        res.row = chosen;
        if (sh->n_ext) {
          __syncthreads();
#pragma unroll
          for (int i = 0; i < KSB_MAX_EXT; ++i) res.device_mask[i] = s_picks[i];
        }
      }
      if (tid == 0) {
        A.results[p] = res;
        // reset the other scratch slot for pod p+1
        KsbNorms* nn = A.norms + ((p + 1) & 1);
        nn->fit_count = 0;
        nn->max_na = 0;
        nn->max_tt = 0;
        nn->max_sc = 0;
        nn->ipa_max = 0;
        nn->ipa_min = 0;
        nn->have_zones = 0;
        nn->one_row = -1;
      }
      int32_t* zn = A.zone_sum + (size_t)((p + 1) & 1) * A.zone_stride;
      for (int z = tid; z < nz; z += kThreads) zn[z] = 0;
    } else {
      // every block tracks the round-robin counter so it never has to be re-read
      if (!whole_pod_error && nm.fit_count >= 2 && !(sh->flags & PF_PRIO_ERROR)) last_index += 1;
    }
    grid_barrier(A, epoch);
  }
  if (b == 0 && tid == 0) *A.last_node_index = last_index;
}

// Applies the assume step for a pod placed outside a batch (ksb_pod_add after ksb_schedule).
__global__ void ksb_apply_kernel(KsbNodeTable T, const KsbPodHeader* header, const uint64_t* pool, int row,
                                 uint32_t m0, uint32_t m1, uint32_t m2, uint32_t m3) {
  __shared__ uint32_t picks[KSB_MAX_EXT];
  if (threadIdx.x == 0) {
    picks[0] = m0;
    picks[1] = m1;
    picks[2] = m2;
    picks[3] = m3;
  }
  __syncwarp();
  PodView pv = make_view(header, pool + header->pool_off, T.W);
  commit_pod(T, row, pv, picks);
}

extern "C" int ksb_launch_batch(const KsbBatchArgs* args, int grid, size_t smem_bytes, cudaStream_t stream) {
  void* params[] = {(void*)args};
  cudaError_t e = cudaLaunchCooperativeKernel((const void*)ksb_schedule_batch_kernel, dim3(grid), dim3(kThreads), params,
                                              smem_bytes, stream);
  return (int)e;
}

extern "C" int ksb_batch_kernel_limits(int device, size_t smem_bytes, int* max_grid, int* threads) {
  int per_sm = 0, sms = 0;
  cudaError_t e = cudaFuncSetAttribute(ksb_schedule_batch_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                       (int)smem_bytes);
  if (e != cudaSuccess) return (int)e;
  e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, ksb_schedule_batch_kernel, kThreads, smem_bytes);
  if (e != cudaSuccess) return (int)e;
  e = cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device);
  if (e != cudaSuccess) return (int)e;
  *max_grid = per_sm * sms;
  *threads = kThreads;
  return 0;
}

extern "C" int ksb_launch_apply(const KsbNodeTable* T, const KsbPodHeader* header, const uint64_t* pool, int row,
                                const uint32_t* masks, cudaStream_t stream) {
  ksb_apply_kernel<<<1, 32, 0, stream>>>(*T, header, pool, row, masks[0], masks[1], masks[2], masks[3]);
  return (int)cudaGetLastError();
}

extern "C" size_t ksb_batch_smem_bytes(int n_zones) {
  return sizeof(KsbPodHeader) + (size_t)kMaxPoolWords * 8 + (size_t)(n_zones + 1) * 4;
}
extern "C" int ksb_max_pool_words() { return kMaxPoolWords; }

// ---------------------------------------------------------------------------------------------
// Scratch for ksb_replay_last_batch_noalloc (include/ksched_b200.h).
//
// Some users of this library drive it from places where asking the CUDA runtime for device memory
// is off the table: a control loop whose memory budget was fixed at start-up (a MIG slice or an MPS
// client with a hard limit, where a late cudaMalloc may simply fail), or a callback running on
// somebody else's stream where cudaMalloc's implicit device-wide synchronisation would stall
// unrelated work. For them the batch kernel's working memory lives here, in one array that is part
// of the loaded module: every GPU that loads the library carries these bytes until the process exits.
// The limits below bound what fits; larger jobs must use ksb_replay_last_batch.
//
// All calls share the one array and it keeps whatever the previous call left behind, so the host
// side re-zeroes the counters the kernel accumulates into before each launch, and two calls whose
// kernels overlap on the same GPU would corrupt each other.
#define KSB_NOALLOC_MAX_ROWS 65536    // padded node rows
#define KSB_NOALLOC_MAX_ZONES 1024    // zone-sum slots per ping-pong half
#define KSB_NOALLOC_MAX_GRID 1024     // thread blocks
namespace {
constexpr size_t na_align(size_t x) { return (x + 255) & ~(size_t)255; }
constexpr size_t kNaNorms = 0;
constexpr size_t kNaZone = kNaNorms + na_align(2 * sizeof(KsbNorms));
constexpr size_t kNaRaw = kNaZone + na_align(2 * (size_t)KSB_NOALLOC_MAX_ZONES * 4);
constexpr size_t kNaScore = kNaRaw + na_align((size_t)KSB_NOALLOC_MAX_ROWS * 16);
constexpr size_t kNaReasons = kNaScore + na_align((size_t)KSB_NOALLOC_MAX_ROWS * 4);
constexpr size_t kNaBlkMax = kNaReasons + na_align((size_t)KSB_NOALLOC_MAX_ROWS * 4);
constexpr size_t kNaBlkCnt = kNaBlkMax + na_align((size_t)KSB_NOALLOC_MAX_GRID * 4);
constexpr size_t kNaBarrier = kNaBlkCnt + na_align((size_t)KSB_NOALLOC_MAX_GRID * 4);
constexpr size_t kNaError = kNaBarrier + 256;
constexpr size_t kNaFitCount = kNaError + 256;
constexpr size_t kNaLastIndex = kNaFitCount + 256;
constexpr size_t kNaFitHist = kNaLastIndex + 256;
constexpr size_t kNaTotal = kNaFitHist + na_align((size_t)KSB_MAX_FITERR * KSB_REASON_BITS * 4);
}  // namespace

static __device__ __align__(256) unsigned char ksb_noalloc_scratch[kNaTotal];  // about 2.1 MiB

// In host code `ksb_noalloc_scratch` is only a stand-in for the symbol, not a device address, so the
// address each GPU gave the array is looked up once per device (with that device current) and cached.
extern "C" int ksb_noalloc_bind(int device, int n_pad, int n_zones, int grid, KsbBatchArgs* a, char* err,
                                size_t err_len) {
  static unsigned char* base_of[64] = {nullptr};
  auto fail = [&](const char* msg) {
    if (err && err_len) snprintf(err, err_len, "%s", msg);
    return -1;
  };
  if (device < 0 || device >= 64) return fail("device ordinal out of range");
  if (n_pad > KSB_NOALLOC_MAX_ROWS || n_zones > KSB_NOALLOC_MAX_ZONES || grid > KSB_NOALLOC_MAX_GRID) {
    if (err && err_len)
      snprintf(err, err_len,
               "batch exceeds the no-allocation limits: %d padded node rows (max %d), %d zone slots (max %d), "
               "%d blocks (max %d); use ksb_replay_last_batch",
               n_pad, KSB_NOALLOC_MAX_ROWS, n_zones, KSB_NOALLOC_MAX_ZONES, grid, KSB_NOALLOC_MAX_GRID);
    return -1;
  }
  if (!base_of[device]) {
    int cur = -1;
    if (cudaGetDevice(&cur) != cudaSuccess || cur != device) return fail("the engine's device is not current");
    unsigned char* p = nullptr;
    cudaError_t e = cudaGetSymbolAddress((void**)&p, ksb_noalloc_scratch);
    if (e != cudaSuccess || !p) {
      if (err && err_len) snprintf(err, err_len, "cudaGetSymbolAddress failed: %s", cudaGetErrorString(e));
      return -1;
    }
    base_of[device] = p;
  }
  unsigned char* b = base_of[device];
  a->norms = (KsbNorms*)(b + kNaNorms);
  a->zone_sum = (int32_t*)(b + kNaZone);
  a->zone_stride = KSB_NOALLOC_MAX_ZONES;
  a->raw = (int4*)(b + kNaRaw);
  a->score = (int32_t*)(b + kNaScore);
  a->reasons = (uint32_t*)(b + kNaReasons);
  a->blk_max = (int32_t*)(b + kNaBlkMax);
  a->blk_cnt = (int32_t*)(b + kNaBlkCnt);
  a->barrier = (uint32_t*)(b + kNaBarrier);
  a->error_flag = (int32_t*)(b + kNaError);
  a->fiterr_count = (int32_t*)(b + kNaFitCount);
  a->last_node_index = (uint64_t*)(b + kNaLastIndex);
  a->fiterr_hist = (int32_t*)(b + kNaFitHist);
  return 0;
}
