// sm_100a kernels of the scheduling engine.
//
// One persistent cooperative kernel walks a batch of pods in FIFO order. Pods are taken in GROUPS of
// up to 32: every pod of a group is evaluated against the node table as it stands when the group
// starts (phases 1 and 2, all blocks), then block 0 commits the group's pods one by one, in order,
// re-evaluating only the rows changed by the pods committed before (phase 3). What the reference
// does per pod (paths under /root/reference/plugin/pkg/scheduler):
//   phase 1  findNodesThatFit   core/generic_scheduler.go:283-362  -> eval_node: every predicate, the
//            fork's device matching (core/extended_resources.go:83-150) and the Map half of every
//            priority, + grid-wide reductions of the Reduce inputs
//   phase 2  PrioritizeNodes    core/generic_scheduler.go:509-640  -> final integer score per node,
//            per 256-row segment (max, tie count, holder mask)
//   phase 3  selectHost         core/generic_scheduler.go:177-192  -> round-robin pick among the
//            max-score ties in descending name (= row) order, then the assume step (scheduler.go:478 ->
//            schedulercache/node_info.go:329-356) on the chosen row.
//
// Why grouping is exact. A pod's evaluation can differ from its group-start evaluation only through
//  (a) counters other pods bump (SelectorSpread columns, inter-pod-affinity tables): the host marks, per
//      pod, the latest earlier pod that writes anything it reads (`dep`), and a group never contains
//      both, so inside a group these inputs are constant;
//  (b) the rows of nodes chosen earlier in the group: phase 3 re-evaluates exactly those rows against
//      the current table. Binding only ever adds load, so a row can only turn from fit to unfit. A row
//      that was fit at group start and no longer is leaves the candidate set if it cannot have held any
//      Reduce input (maxima, zone sums, inter-pod extremes); otherwise, or when the maximum lost all its
//      holders, the speculation is dropped and the group ends before that pod, which is then evaluated
//      from scratch as the first pod of the next group. Every unchanged row keeps its group-start
//      score, a changed row's score is its group-start score with the base part exchanged, and the tie
//      counts are patched before the rank-select.
// With group size 1 this degenerates to evaluating every pod against the live table.
//
// Phase 3 is a decider warp and a writer warp: the changed rows' mutable columns live in a shared-memory
// row cache (filled when a row is first chosen, updated by every later commit), so re-evaluating a pod
// on them costs no memory round trip; all warps first bring the group's segment statistics into shared
// memory and reduce them to a maximum, tie count and block sums per pod; the row inside a segment comes
// from the segment's holder mask (phase 2); the writer warp takes the table write-through, the class
// counters and the result records off the decider's chain; FitError histograms are built in phase 2.
//
// Three instantiations: plain; pipelined (one rank: block 0 commits group k-1 while the other blocks
// evaluate group k, see `pipe` in the kernel); sharded (several GPUs, rows of other shards reached over
// NVLink peer mappings).
//
// The work is integer/bit work over a structure-of-arrays node table: one thread per node, coalesced
// 4/8-byte column loads with the common columns of a row kept in registers across the pods of a group,
// requirement blocks staged in shared memory, reductions by warp shuffle (redux) + one atomic per
// block. No tensor cores: nothing here is a dense contraction. float64 priorities use __d*_rn so no FMA
// is ever formed (Go on amd64 never fuses; the file is also compiled with -fmad=false).
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#include "device_types.h"
#include "kernels.h"

// 1: the decider warp also reads the cycle counter between its per-pod steps (ksb_stats.stage_cycles[8..10]);
// costs a few percent of phase 3, so it is a build option (make FINE_TIMING=1)
#ifndef KSB_FINE_TIMING
#define KSB_FINE_TIMING 0
#endif

namespace {

constexpr int kThreads = 256;
constexpr int kWarps = kThreads / 32;
constexpr int kSeg = 256;             // rows per segment (= per block iteration)
constexpr int kMaxG = KSB_MAX_GROUP;  // pods per group
constexpr int kMaxPoolWords = 1024;   // pool words one pod may use (checked on the host)
constexpr int kGroupPoolWords = 4096; // pool words staged per group
constexpr int kSegCache = KSB_SEG_CACHE;  // (pods in group) x (segments) entries of segment stats kept in shared memory
constexpr int kCache = 32;            // row-cache entries (one lane each)
constexpr bool kHoldTableWrites = false;  // pipelined schedule: see the writer warp
constexpr int kDeciders = kWarps - 1; // phase 3: warps that decide; the last warp writes the table through

template <bool CG, typename V>
__device__ __forceinline__ V ldx(const V* p) {
  if (CG) return __ldcg(p);  // L2-coherent: rows block 0 itself just rewrote
  return *p;
}

struct PodView {
  const KsbPodHeader* h;
  const uint64_t* exprs;
  const uint64_t* ports;
  const uint64_t* extcls;
  const uint64_t* checks;
  const uint64_t* score;
  const uint64_t* commits;
  const uint64_t* spread_inc;
  const uint64_t* tw_ext;  // tw > 1: [tw-1] untolerated NoSchedule/NoExecute words, [tw-1] PreferNoSchedule, [tw-1] preferAvoid
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
  v.tw_ext = p + h->n_spread_inc;

  return v;
}

__device__ __forceinline__ int pool_words(const KsbPodHeader* h, int W, int tw) {
  return (int)h->n_exprs * W + ((h->flags & PF_HAS_PORTS) ? 6 * KSB_PORT_WORDS : 0) + (int)h->n_ext * KSB_CLASS_WORDS +
         h->n_checks + 2 * (int)h->n_score + 2 * (int)h->n_commit + h->n_spread_inc + 3 * (tw - 1);
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
// ((capacity - requested) * 10) / capacity, exactly, without the 64-bit integer division routine: the
// quotient is in [0, 10].
__device__ __forceinline__ int64_t unused_score(int64_t requested, int64_t capacity) {
  if (capacity == 0) return 0;
  if (requested > capacity) return 0;
  if (capacity < 0) return ((capacity - requested) * 10) / capacity;  // never for real nodes; keep Go's semantics
  const int64_t num = (capacity - requested) * 10;
  if (num < 0) return num / capacity;  // negative request: not in [0, 10]
  if (capacity < (1ll << 59)) {
    // the quotient is the number of s in 1..10 with s * capacity <= num: ten independent compares, no division
    int32_t cnt = 0;
#pragma unroll
    for (int s = 1; s <= 10; ++s) cnt += (num >= (int64_t)s * capacity) ? 1 : 0;
    return cnt;
  }
  int64_t q = (int64_t)__ddiv_rn((double)num, (double)capacity);
  int64_t r = num - q * capacity;
  while (r < 0) {
    --q;
    r += capacity;
  }
  while (r >= capacity) {
    ++q;
    r -= capacity;
  }
  return q;
}

// 1/a for a in [1, 2^50): hardware estimate + two Newton steps (relative error of a few 2^-53). Only used where an
// exact integer or guard-band check follows.
__device__ __forceinline__ double fast_rcp(double a) {
  double x;
  asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(x) : "d"(a));
  x = __dmul_rn(x, __dsub_rn(2.0, __dmul_rn(a, x)));
  x = __dmul_rn(x, __dsub_rn(2.0, __dmul_rn(a, x)));
  return x;
}

// LeastRequested (least_requested.go:33-91) and BalancedResourceAllocation (balanced_resource_allocation.go:51-100)
// for one (pod, node), with the results of the reference's arithmetic and without its divisions:
//  * ((cap - req) * 10) / cap is estimated as floor(10 - 10 * req * (1/cap)) and corrected with the exact integer
//    remainder, so it is exact whatever the estimate's rounding;
//  * int((1 - |req_c/cap_c - req_m/cap_m|) * 10) is computed with reciprocal multiplies (error below 2e-14 on a value
//    in (0, 10]); only when that lands within 1e-9 of an integer are the reference's two float64 divisions done.
// rcp_c / rcp_m: fast_rcp of the capacities (per row). Outside 0 < capacity < 2^50, 0 <= request the reference's
// own formulas run.
__device__ __forceinline__ void lr_bra(int64_t cpu, int64_t mem, int64_t acpu, int64_t amem, double rcp_c, double rcp_m,
                                       int32_t* lr_out, int32_t* bra_out) {
  const int64_t lim = 1ll << 50;
  if (acpu > 0 && amem > 0 && acpu < lim && amem < lim && cpu >= 0 && mem >= 0 && cpu < lim && mem < lim) {
    const double cfa = __dmul_rn((double)cpu, rcp_c);
    const double mfa = __dmul_rn((double)mem, rcp_m);
    int64_t qc = 0, qm = 0;
    if (cpu <= acpu) {
      qc = (int64_t)__double2int_rd(__dsub_rn(10.0, __dmul_rn(10.0, cfa)));
      const int64_t r = (acpu - cpu) * 10 - qc * acpu;
      qc += (r >= acpu) ? 1 : ((r < 0) ? -1 : 0);
    }
    if (mem <= amem) {
      qm = (int64_t)__double2int_rd(__dsub_rn(10.0, __dmul_rn(10.0, mfa)));
      const int64_t r = (amem - mem) * 10 - qm * amem;
      qm += (r >= amem) ? 1 : ((r < 0) ? -1 : 0);
    }
    *lr_out = (int32_t)((qc + qm) / 2);
    int32_t bra = 0;
    if (cpu < acpu && mem < amem) {  // below 2^50 the float64 fractions reach 1.0 exactly when request >= capacity
      const double v = __dmul_rn(__dsub_rn(1.0, fabs(__dsub_rn(cfa, mfa))), 10.0);
      const double near = rint(v);
      if (fabs(__dsub_rn(v, near)) > 1e-9) {
        bra = (int32_t)v;
      } else {
        const double cf = __ddiv_rn((double)cpu, (double)acpu);
        const double mf = __ddiv_rn((double)mem, (double)amem);
        bra = (int32_t)(__dmul_rn(__dsub_rn(1.0, fabs(__dsub_rn(cf, mf))), 10.0));
      }
    }
    *bra_out = bra;
    return;
  }
  *lr_out = (int32_t)((unused_score(cpu, acpu) + unused_score(mem, amem)) / 2);
  const double cf = acpu == 0 ? 1.0 : __ddiv_rn((double)cpu, (double)acpu);
  const double mf = amem == 0 ? 1.0 : __ddiv_rn((double)mem, (double)amem);
  int32_t bra = 0;
  if (!(cf >= 1.0 || mf >= 1.0)) {
    double diff = fabs(__dsub_rn(cf, mf));
    bra = (int32_t)(__dmul_rn(__dsub_rn(1.0, diff), 10.0));
  }
  *bra_out = bra;
}

// The fork's per-node device selection (core/extended_resources.go:83-150), devices visited in
// ascending id order (= bit order). Returns the failing reason bit or 0; fills picks[] on success.
// `avail_in` holds the node's available-device masks per resource slot.
__device__ __forceinline__ uint32_t select_devices(const KsbNodeTable& T, int r, const PodView& pv, uint32_t nflags,
                                                   const uint32_t* avail_in, uint32_t* picks) {
  const KsbPodHeader* h = pv.h;
  uint32_t avail[KSB_MAX_RES];
#pragma unroll
  for (int s = 0; s < KSB_MAX_RES; ++s) avail[s] = avail_in[s];
  for (int i = 0; i < h->n_ext; ++i) {
    if ((h->flags & PF_EXT_LIMITS_ERROR) && i == h->ext_err_entry) return 1u << RB_EXT_LIMITS;
    int rs = h->ext[i].rslot;
    int num = h->ext[i].num;
    if (rs == 255 || !(nflags & (256u << rs))) return 1u << (RB_EXT0 + i);  // available[rName] missing
    uint32_t a = avail[rs];
    uint32_t sel = 0;
    int got = 0;
    if (num > 0) {
      const uint64_t* cm = pv.extcls + i * KSB_CLASS_WORDS;
      const uint4* cls4 = reinterpret_cast<const uint4*>(T.dev_class + ((size_t)rs * T.n_pad + r) * KSB_MAX_DEV);
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        uint4 v = cls4[q];  // device classes are static: never rewritten by the kernel
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

// Position of the (n+1)-th set bit of w counted from bit 0 upwards (w has more than n set bits).
__device__ __forceinline__ int nth_set_from_bottom(uint32_t w, int n) {
  int p = 0;
#pragma unroll
  for (int width = 16; width >= 1; width >>= 1) {
    const int c = __popc(w & ((1u << width) - 1u));
    if (n >= c) {
      n -= c;
      p += width;
      w >>= width;
    }
  }
  return p;
}

// Devices of row r, resource slot rs, whose attribute class is in the mask cm (isDeviceAMatch,
// core/extended_resources.go:152-167, precomputed per class on the host). One thread.
__device__ __forceinline__ uint32_t device_match_mask(const KsbNodeTable& T, int r, int rs, const uint64_t* cm) {
  const uint4* cls4 = reinterpret_cast<const uint4*>(T.dev_class + ((size_t)rs * T.n_pad + r) * KSB_MAX_DEV);
  uint4 v[4];
#pragma unroll
  for (int q = 0; q < 4; ++q) v[q] = cls4[q];  // device classes are static: never rewritten by the kernel
  uint32_t match = 0;
#pragma unroll
  for (int q = 0; q < 4; ++q) {
    const uint32_t w[4] = {v[q].x, v[q].y, v[q].z, v[q].w};
#pragma unroll
    for (int j = 0; j < 4; ++j) {
#pragma unroll
      for (int half = 0; half < 2; ++half) {
        const uint32_t c = (w[j] >> (16 * half)) & 0xffffu;
        match |= (uint32_t)((cm[c >> 6] >> (c & 63)) & 1ull) << (q * 8 + j * 2 + half);
      }
    }
  }
  return match;
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

// The columns of one row that almost every pod reads. Phase 1 evaluates many pods of a group on the same row, so
// they are loaded once per row and kept in registers across the pods.
constexpr int kLabFast = 4;  // label words held in registers (the default label_words)
struct RowCommon {
  int64_t acpu, amem, rcpu, rmem, ncpu, nmem;
  uint64_t hard, soft, avoid;
  uint64_t lab[kLabFast];
  int32_t num_pods, alloc_pods;
  uint32_t nflags, zone;
  double rcp_cpu, rcp_mem;  // fast_rcp of the allocatable cpu / memory (lr_bra)
  uint32_t over;            // bit 0: allocatable alpha-gpu < requested on the row, bit 1: the same for ephemeral storage
};

template <bool CG>
__device__ __forceinline__ void load_row_common(const KsbNodeTable& T, int r, RowCommon& rc) {
  const size_t np = T.n_pad;
  rc.num_pods = ldx<CG>(T.num_pods + r);
  rc.alloc_pods = T.alloc_pods[r];
  rc.acpu = T.alloc_cpu[r];
  rc.amem = T.alloc_mem[r];
  rc.rcpu = ldx<CG>(T.req_cpu + r);
  rc.rmem = ldx<CG>(T.req_mem + r);
  rc.ncpu = ldx<CG>(T.nz_cpu + r);
  rc.nmem = ldx<CG>(T.nz_mem + r);
  rc.nflags = T.flags[r];
  rc.zone = T.zone[r];
  rc.hard = T.taint_hard[r];
  rc.soft = T.taint_soft[r];
  rc.avoid = T.avoid[r];
#pragma unroll
  for (int w = 0; w < kLabFast; ++w) rc.lab[w] = w < T.W ? T.labels[w * np + r] : 0ull;
  rc.rcp_cpu = fast_rcp((double)rc.acpu);
  rc.rcp_mem = fast_rcp((double)rc.amem);
  // pods that ask for neither (nearly all) then need no load of these four columns
  rc.over = (T.alloc_agpu[r] < ldx<CG>(T.req_agpu + r) ? 1u : 0u) | (T.alloc_eph[r] < ldx<CG>(T.req_eph + r) ? 2u : 0u);
}

// eval_expr for label_words == kLabFast: the row's label words come from registers. `wm` names the words the mask
// touches (host-computed): an untouched word changes neither "any" nor "all", so it is not even loaded.
__device__ __forceinline__ bool eval_expr4(const uint64_t (&lab)[kLabFast], const uint64_t* mask, uint32_t kind, uint32_t wm) {
  bool any = false, all = true;
#pragma unroll
  for (int w = 0; w < kLabFast; ++w) {
    if ((wm >> w) & 1u) {  // uniform across the warp: one pod per iteration
      const uint64_t m = mask[w];
      const uint64_t x = lab[w] & m;
      any |= (x != 0);
      all &= (x == m);
    }
  }
  return kind == EK_ANY ? any : (kind == EK_NONE ? !any : all);
}

// PodMatchNodeSelector (predicates.go:691-761) and the NodeAffinityPriority Map (node_affinity.go:34-75) over the
// pod's compiled expressions; EXPR(e) evaluates expression e on the row's labels.
#define KSB_LABEL_RULES(EXPR)                                                    \
  do {                                                                           \
    int e = 0;                                                                   \
    bool match = true;                                                           \
    if (pf & PF_NODE_SELECTOR) {                                                 \
      match = EXPR(0);                                                           \
      e = 1;                                                                     \
    }                                                                            \
    if (pf & PF_HAS_REQUIRED) {                                                  \
      bool any = false;                                                          \
      for (int t = 0; t < h->n_req_terms; ++t) {                                 \
        bool all = true;                                                         \
        const int end = h->term_end[t];                                          \
        for (; e < end; ++e) all &= EXPR(e);                                     \
        any |= all;                                                              \
      }                                                                          \
      match &= any;                                                              \
    }                                                                            \
    if (!match) reasons |= 1u << RB_SELECTOR;                                    \
    for (int t = 0; t < h->n_pref; ++t) {                                        \
      bool all = true;                                                           \
      const int end = h->pref_end[t];                                            \
      for (; e < end; ++e) all &= EXPR(e);                                       \
      if (all) na += h->pref_weight[t];                                          \
    }                                                                            \
  } while (0)

// The priorities a Policy can enable next to the DefaultProvider set (Map only, no Reduce), as one addend to the
// bind-dependent part of the score: MostRequestedPriority (most_requested.go:32-88) over the same non-zero totals as
// LeastRequested, ResourceLimitsPriority (resource_limits.go:34-94) and the node's NodeLabelPriority sum (node_label.go).
// (a real call, taken only when a Policy enabled one of the two: inlined, its 64-bit divisions cost every evaluation
// registers and spills, 3 % of the headline launch and 15 % of a one-pod call)
__device__ __noinline__ int32_t policy_score_enabled(int32_t w_mr, int32_t w_rl, int64_t lim_cpu, int64_t lim_mem, int64_t cpu,
                                                     int64_t mem, int64_t acpu, int64_t amem) {
  int32_t s = 0;
  if (w_mr) {
    // most_requested.go:61-75: (requested * 10) / capacity is in [0, 10] here: the number of s in 1..10 with
    // s * capacity <= requested * 10, ten independent compares instead of the 64-bit division routine
    auto used = [](int64_t requested, int64_t capacity) -> int64_t {
      if (capacity == 0 || requested > capacity) return 0;
      const int64_t num = requested * 10;
      if (capacity < 0 || num < 0 || capacity >= (1ll << 59)) return num / capacity;  // never for real nodes; Go's own arithmetic
      int32_t cnt = 0;
#pragma unroll
      for (int k = 1; k <= 10; ++k) cnt += (num >= (int64_t)k * capacity) ? 1 : 0;
      return cnt;
    };
    s += w_mr * (int32_t)((used(cpu, acpu) + used(mem, amem)) / 2);
  }
  if (w_rl) {
    const bool c1 = lim_cpu != 0 && acpu != 0 && lim_cpu <= acpu;
    const bool m1 = lim_mem != 0 && amem != 0 && lim_mem <= amem;
    s += w_rl * ((c1 || m1) ? 1 : 0);
  }
  return s;
}
__device__ __forceinline__ int32_t policy_score(const KsbNodeTable& T, const KsbPodHeader* h, int64_t cpu, int64_t mem, int64_t acpu,
                                                int64_t amem, int32_t static_score, int r) {
  // ImageLocalityPriority (image_locality.go:32-62): the host keeps the 0..10 score of the pod's image list per node
  if (T.w_il && h->img_col >= 0) static_score += T.w_il * (int32_t)T.spread[(size_t)h->img_col * T.n_pad + r];
  if ((T.w_mr | T.w_rl) == 0) return static_score;
  return static_score + policy_score_enabled(T.w_mr, T.w_rl, h->lim_cpu, h->lim_mem, cpu, mem, acpu, amem);
}

// Everything the reference computes for one (pod, node): podFitsOnNode's predicates
// (core/generic_scheduler.go:402-501) and the Map functions of PrioritizeNodes (:557-574).
// CG = true reads the columns the kernel mutates straight from L2 (phase 3, rows just rewritten).
template <bool CG, bool POL>
__device__ __forceinline__ void eval_node(const KsbNodeTable& T, int r, const RowCommon& rc, const PodView& pv, EvalOut& o,
                                          uint32_t* picks) {
  const KsbPodHeader* h = pv.h;
  const size_t np = T.n_pad;
  uint32_t reasons = 0;
  const uint32_t pf = h->flags;

  // --- PodFitsResources (predicates.go:630-687)
  const int32_t num_pods = rc.num_pods;
  const int32_t alloc_pods = rc.alloc_pods;
  const int64_t acpu = rc.acpu, amem = rc.amem;
  const int64_t ncpu = rc.ncpu, nmem = rc.nmem;
  const uint32_t nflags = rc.nflags;
  if (num_pods + 1 > alloc_pods) reasons |= 1u << RB_PODS;
  if (pf & PF_HAS_REQUEST) {
    if (acpu < h->req_cpu + rc.rcpu) reasons |= 1u << RB_CPU;
    if (amem < h->req_mem + rc.rmem) reasons |= 1u << RB_MEM;
    if (h->req_agpu == 0) {
      if (rc.over & 1u) reasons |= 1u << RB_AGPU;
    } else if (T.alloc_agpu[r] < h->req_agpu + ldx<CG>(T.req_agpu + r)) {
      reasons |= 1u << RB_AGPU;
    }
    if (h->req_eph == 0) {
      if (rc.over & 2u) reasons |= 1u << RB_EPH;
    } else if (T.alloc_eph[r] < h->req_eph + ldx<CG>(T.req_eph + r)) {
      reasons |= 1u << RB_EPH;
    }
    const uint32_t sm = h->scalar_mask;
    if (sm) {
#pragma unroll
      for (int s = 0; s < KSB_MAX_SCALARS; ++s)
        if ((sm >> s) & 1u)
          if (T.alloc_scalar[s * np + r] < h->scalar_req[s] + ldx<CG>(T.req_scalar + s * np + r))
            reasons |= 1u << (RB_SCALAR0 + s);
    }
  }
  // --- PodFitsHost (predicates.go:764-776)
  if (h->pinned_row != -1 && h->pinned_row != T.row_base + r) reasons |= 1u << RB_HOSTNAME;
  // --- PodFitsHostPorts (predicates.go:925-945, utils.go:161-202)
  if (pf & PF_HAS_PORTS) {
    uint64_t hit = 0;
#pragma unroll
    for (int w = 0; w < KSB_PORT_WORDS; ++w) {
      hit |= ldx<CG>(T.port_any + w * np + r) & pv.ports[w];
      hit |= ldx<CG>(T.port_wild + w * np + r) & pv.ports[KSB_PORT_WORDS + w];
      hit |= ldx<CG>(T.port_exact + w * np + r) & pv.ports[2 * KSB_PORT_WORDS + w];
    }
    if (hit) reasons |= 1u << RB_PORTS;
  }
  // --- PodMatchNodeSelector (predicates.go:691-761) and NodeAffinityPriority Map (node_affinity.go:34-75)
  int32_t na = 0;
  if (h->n_exprs) {
    const int W = T.W;
#define KSB_KIND(e) ((h->expr_kind[(e) >> 4] >> (((e) & 15) * 2)) & 3u)
    if (W == kLabFast) {
#define KSB_EXPR_FAST(e) eval_expr4(rc.lab, pv.exprs + (e) * kLabFast, KSB_KIND(e), h->expr_wm[(e)])
      KSB_LABEL_RULES(KSB_EXPR_FAST);
#undef KSB_EXPR_FAST
    } else {
      uint64_t lab[16];
      for (int w = 0; w < W; ++w) lab[w] = T.labels[w * np + r];
#define KSB_EXPR_GEN(e) eval_expr(lab, pv.exprs + (e) * W, KSB_KIND(e), W)
      KSB_LABEL_RULES(KSB_EXPR_GEN);
#undef KSB_EXPR_GEN
    }
#undef KSB_KIND
  } else if ((pf & PF_HAS_REQUIRED)) {
    reasons |= 1u << RB_SELECTOR;  // required affinity with no usable term matches nothing
  }
  // --- PodToleratesNodeTaints (predicates.go:1349-1373) and TaintTolerationPriority Map (taint_toleration.go:55-73)
  if (rc.hard & h->hard_untol) reasons |= 1u << RB_TAINTS;
  int32_t tt = __popcll(rc.soft & h->soft_untol);
  uint64_t avoid_hit = rc.avoid & h->avoid_mask;
  if (POL && T.tw > 1) {  // more than 64 distinct taints / references: the words above 0 (static columns)
    const int x = T.tw - 1;
    uint64_t hard_hit = 0;
    for (int w = 0; w < x; ++w) {
      hard_hit |= T.taint_hard[(size_t)(w + 1) * np + r] & pv.tw_ext[w];
      tt += __popcll(T.taint_soft[(size_t)(w + 1) * np + r] & pv.tw_ext[x + w]);
      avoid_hit |= T.avoid[(size_t)(w + 1) * np + r] & pv.tw_ext[2 * x + w];
    }
    if (hard_hit) reasons |= 1u << RB_TAINTS;
  }
  // --- pressure / conditions (predicates.go:1382-1441)
  if ((pf & PF_BEST_EFFORT) && (nflags & NF_MEM_PRESSURE)) reasons |= 1u << RB_MEM_PRESSURE;
  if (nflags & NF_DISK_PRESSURE) reasons |= 1u << RB_DISK_PRESSURE;
  if (nflags & NF_NOT_READY) reasons |= 1u << RB_NOT_READY;
  if (nflags & NF_OUT_OF_DISK) reasons |= 1u << RB_OUT_OF_DISK;
  if (nflags & NF_NET_UNAVAILABLE) reasons |= 1u << RB_NET_UNAVAILABLE;
  if (nflags & NF_UNSCHEDULABLE) reasons |= 1u << RB_UNSCHEDULABLE;
  if (POL && (nflags & NF_LABEL_PRESENCE)) reasons |= 1u << RB_LABEL_PRESENCE;  // CheckNodeLabelPresence (predicates.go:800-826), per node on the host
  // --- MatchInterPodAffinity (predicates.go:1047-1346) over per-(term class, topology domain) counts.
  //     Inside a group no pod reads a table an earlier pod of the group writes, so plain loads do.
  for (int i = 0; i < h->n_checks; ++i) {
    uint64_t c = pv.checks[i];
    uint32_t off = (uint32_t)c;
    uint32_t tk = (uint32_t)(c >> 32) & 0xffu;
    uint32_t mode = (uint32_t)(c >> 40) & 0xffu;
    uint32_t self = (uint32_t)(c >> 48) & 1u;
    uint32_t d = mode == CM_FAIL_ALWAYS ? 0u : T.topo[tk * np + r];
    bool in_domain = d != 0 && ldx<CG>(T.tables + off + d) > 0;
    uint32_t bad = 0;
    if (mode == CM_FORBID_EXISTING) {
      if (in_domain) bad = 1u << RB_IPA_EXISTING_ANTI;
    } else if (mode == CM_REQUIRE) {
      if (!in_domain) {
        bool exists = ldx<CG>(T.tables + off) > 0;  // slot 0: matching pods anywhere
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
  // --- scheduler extenders filter after the predicates and before the fork's device selection (generic_scheduler.go:333-351)
  if constexpr (POL)
    if (reasons == 0 && T.ext_keep && !((T.ext_keep[r >> 5] >> (r & 31)) & 1u)) reasons |= 1u << RB_EXTENDER;  // only nodes that passed are shown to the extenders
  // --- fork: device selection runs only on nodes that passed every predicate (generic_scheduler.go:357)
  if (reasons == 0 && h->n_ext) {
    uint32_t local[KSB_MAX_EXT];
    uint32_t avail[KSB_MAX_RES];
#pragma unroll
    for (int s = 0; s < KSB_MAX_RES; ++s) avail[s] = ldx<CG>(T.dev_avail + (size_t)s * np + r);
    uint32_t bad = select_devices(T, r, pv, nflags, avail, picks ? picks : local);
    reasons |= bad;
  }

  // --- priorities (Map halves)
  const int64_t cpu = h->nz_cpu + ncpu;
  const int64_t mem = h->nz_mem + nmem;
  int32_t lr, bra;  // least_requested.go:63-91, balanced_resource_allocation.go:51-100
  lr_bra(cpu, mem, acpu, amem, rc.rcp_cpu, rc.rcp_mem, &lr, &bra);
  const int32_t av = avoid_hit ? 0 : 10;  // node_prefer_avoid_pods.go:29-59
  // (the node's NodeLabelPriority sum is read here, and only when a Policy configured an instance: carried with the
  //  row's common columns it cost every evaluation a register)
  o.base = T.w_lr * lr + T.w_bra * bra + T.w_avoid * av;
  if constexpr (POL) o.base += policy_score(T, h, cpu, mem, acpu, amem, T.has_static ? T.static_score[r] : 0, r);
  o.na = na;
  o.tt = tt;
  o.sc = h->spread_col >= 0 ? ldx<CG>(T.spread + (size_t)h->spread_col * np + r) : 0u;  // selector_spreading.go:64-113
  int64_t ipa = 0;  // interpod_affinity.go:121-216
  for (int i = 0; i < h->n_score; ++i) {
    uint64_t c = pv.score[2 * i];
    uint32_t d = T.topo[((uint32_t)(c >> 32) & 0xffu) * np + r];
    if (d) ipa += (int64_t)pv.score[2 * i + 1] * (int64_t)ldx<CG>(T.tables + (uint32_t)c + d);
  }
  o.ipa = (int32_t)ipa;
  o.zone = rc.zone;
  o.reasons = reasons;
}

// PrioritizeNodes' Reduce functions and weighted sum (generic_scheduler.go:575-606). The float64 pieces are
// functions of a handful of values per pod (the node's matching-pod count, its zone), so phase 2 tabulates them per
// pod with exactly these operations (ScoreLut) and a (pod, node) evaluation looks them up; values beyond the tables
// are computed in place by the same functions.
// selector_spreading.go:141-147: 10 * (maxCount - count) / maxCount
__device__ __forceinline__ double spread_node_f(int32_t max_sc, int32_t sc) {
  return max_sc > 0 ? __dmul_rn(10.0, __ddiv_rn((double)(max_sc - sc), (double)max_sc)) : 10.0;
}
// selector_spreading.go:151-164: zoneWeighting * zoneScore
__device__ __forceinline__ double spread_zone_term(int32_t max_zone, int32_t zone_sum) {
  const double zw = 2.0 / 3.0;
  const double zs = max_zone > 0 ? __dmul_rn(10.0, __ddiv_rn((double)(max_zone - zone_sum), (double)max_zone)) : 10.0;
  return __dmul_rn(zw, zs);
}
__device__ __forceinline__ double spread_node_part(double f) {  // fScore * (1.0 - zoneWeighting)
  const double zw = 2.0 / 3.0;
  return __dmul_rn(f, __dsub_rn(1.0, zw));
}
// reduce.go:29-63, reverse = true
__device__ __forceinline__ int32_t taint_score(int32_t max_tt, int32_t tt) { return max_tt > 0 ? 10 - (10 * tt) / max_tt : 10; }

constexpr int kLutSc = 8, kLutZone = 16, kLutTt = 8;
struct ScoreLut {
  double fs[kLutSc];     // spread_node_f(max_sc, count)
  double fsz[kLutSc];    // spread_node_part(fs[count])
  double fz[kLutZone];   // spread_zone_term(max_zone, zone_sum[zone])
  int32_t tts[kLutTt];   // taint_score(max_tt, count)
};

__device__ __forceinline__ int32_t final_score(const KsbNodeTable& T, const KsbNorms& nm, const ScoreLut& L, const int32_t* zone_sum,
                                               int32_t max_zone, int32_t base, int32_t na, int32_t tt, uint32_t sc,
                                               int32_t ipa, uint32_t zone) {
  // selector_spreading.go:119-174
  const bool small = sc < (uint32_t)kLutSc;
  double f;
  if (nm.have_zones && zone != 0) {
    const double part = small ? L.fsz[sc] : spread_node_part(spread_node_f(nm.max_sc, (int32_t)sc));
    const double zt = zone < (uint32_t)kLutZone ? L.fz[zone] : spread_zone_term(max_zone, zone_sum[zone]);
    f = __dadd_rn(part, zt);
  } else {
    f = small ? L.fs[sc] : spread_node_f(nm.max_sc, (int32_t)sc);
  }
  const int32_t ss = (int32_t)f;
  // reduce.go:29-63
  const int32_t nas = na == 0 ? 0 : (nm.max_na > 0 ? (10 * na) / nm.max_na : na);
  const int32_t tts = tt < kLutTt ? L.tts[tt] : taint_score(nm.max_tt, tt);
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

__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}

// Spin until *p (read with `sys`) reaches `target`; a bounded wait turns a lost participant into an error.
template <bool SYS>
__device__ __forceinline__ void spin_until(KsbBatchArgs& A, const uint32_t* p, uint32_t target, int code) {
  uint32_t spins = 0;
  long long t0 = 0;
  while ((int32_t)((SYS ? ld_acquire_sys(p) : ld_acquire(p)) - target) < 0) {
    ++spins;
    if (spins > 64u) __nanosleep(256);  // a long wait (the serial phase): stop hammering the counter's L2 slice
    if ((spins & 1023u) == 0) {
      long long now = clock64();
      if (t0 == 0) t0 = now;
      if (now - t0 > 6000000000ll || ld_acquire((const uint32_t*)A.error_flag) != 0) {  // ~3 s
        if (atomicCAS(A.error_flag, 0, code) == 0)  // first failure: remember which barrier and who
          A.counters[15] = ((uint64_t)(target / gridDim.x) << 32) | ((uint64_t)blockIdx.x << 8) | (uint64_t)code;
        break;
      }
    }
  }
}

// Barrier over every block of every GPU holding a shard. Locally: a monotonically increasing arrival
// counter (cooperative launch guarantees co-residency). Across GPUs: one block of each rank (`xb`), once its own
// grid has arrived, bumps a counter in rank 0's memory over NVLink and waits for all ranks, then releases
// its grid. With one rank it is a plain grid barrier.
// `wait` false: arrive without waiting; the block contributes nothing to the phase that ends here.
// `push` runs on the crossing block after its grid has arrived and before the cross-GPU leg (the rank's partial
// reductions go to rank 0 in one go instead of one NVLink atomic per block and value); `fetch` after it, before the
// grid is released (totals and progress words come back once per rank instead of once per block).
#if KSB_FINE_TIMING == 3
__device__ long long g_xdbg[8];  // crossing block: cycles in local wait, push, cross-GPU leg, fetch
#define KSB_XT(i)                                   \
  do {                                              \
    if (threadIdx.x == 0) {                         \
      const long long n_ = clock64();               \
      g_xdbg[i] += n_ - xt_;                        \
      xt_ = n_;                                     \
    }                                               \
  } while (0)
#else
#define KSB_XT(i) ((void)0)
#endif
struct NoXfer {
  __device__ __forceinline__ void operator()() const {}
};
// `far`: this block stored into another GPU's memory since the last barrier (results for phase 3 on a rank other
// than 0, the writer warp's copies on rank 0): its arrival is preceded by a system-scope fence instead of a GPU-scope
// one. `early` (rank 0 at the barrier that ends an iteration): the grid is released as soon as every rank has
// arrived; what this rank sends (progress words, its own flag) leaves afterwards, off the committing block's path.
template <class Push, class Fetch>
__device__ __forceinline__ void grid_barrier_x(KsbBatchArgs& A, uint32_t& epoch, uint32_t& xepoch, bool wait, int xb, Push push,
                                               Fetch fetch, bool far = false, bool early = false) {
  __syncthreads();
  if (A.n_ranks == 1) {
    if (threadIdx.x == 0) {
      epoch += gridDim.x;
      if (wait) __threadfence();  // a block that does not wait contributed nothing to the phase that ends here
      atomicAdd(A.barrier, 1u);
      if (wait) spin_until<false>(A, A.barrier, epoch, 1);
    }
    __syncthreads();
    return;
  }
  const bool crossing = (int)blockIdx.x == xb;
#if KSB_FINE_TIMING == 3
  long long xt_ = clock64();
#endif
  if (threadIdx.x == 0) {
    epoch += gridDim.x;
    if (far) __threadfence_system();
    else if (wait) __threadfence();
    if (crossing) KSB_XT(0);
    atomicAdd(A.barrier, 1u);
    if (crossing) spin_until<false>(A, A.barrier, epoch, 2);
    else if (wait) spin_until<false>(A, A.release, epoch, 4);
  }
  if (crossing) {
    __syncthreads();
    KSB_XT(1);
    // arrival number of this barrier (every thread of the block keeps count; the other blocks never read theirs)
    xepoch += (uint32_t)A.n_ranks;
    const uint32_t* my_flags = reinterpret_cast<const uint32_t*>(A.xpeer[A.rank] + A.xo_flag);
    if (early) {
      // every other rank's flag (its results for the next commit came before it), then the grid goes on
      if ((int)threadIdx.x < A.n_ranks && (int)threadIdx.x != A.rank) spin_until<true>(A, my_flags + threadIdx.x, xepoch, 3);
      __syncthreads();
      KSB_XT(4);
      fetch();
      __syncthreads();
      if (threadIdx.x == 0) {
        __threadfence();
        atomicExch(A.release, epoch);
      }
    }
    push();
    __syncthreads();
    KSB_XT(2);
    if (threadIdx.x == 0) {
      __threadfence_system();  // this block's pushes, and through the local arrivals every block's writes, before the flags
      KSB_XT(3);
      for (int q = 0; q < A.n_ranks; ++q)
        *reinterpret_cast<volatile uint32_t*>(A.xpeer[q] + A.xo_flag + 4 * A.rank) = xepoch;
    }
    if (!early) {
      // one lane per rank polls this GPU's own flag array: no NVLink read on the path
      if ((int)threadIdx.x < A.n_ranks) spin_until<true>(A, my_flags + threadIdx.x, xepoch, 3);
      __syncthreads();
      KSB_XT(4);
      fetch();
      __syncthreads();
      KSB_XT(5);
      if (threadIdx.x == 0) {
        __threadfence();
        atomicExch(A.release, epoch);
      }
    }
  }
  if (threadIdx.x == 0 && wait) __threadfence();
  __syncthreads();
}
// The barrier between phases 1 and 2 on a sharded table: each GPU's own grid barrier, with the exchange of the Reduce
// inputs (`exch`, which needs no other cross-GPU synchronisation: see there) run by the crossing block between the
// arrival of its grid and its release. `x1` numbers the exchanges (carried from launch to launch).
template <class Exch>
__device__ __forceinline__ void reduce_barrier_x(KsbBatchArgs& A, uint32_t& epoch, uint32_t& x1, bool wait, int xb, Exch exch) {
  __syncthreads();
  x1 += 1u;
  const bool crossing = (int)blockIdx.x == xb;
  if (threadIdx.x == 0) {
    epoch += gridDim.x;
    if (wait) __threadfence();
    atomicAdd(A.barrier + 1, 1u);
    if (crossing) spin_until<false>(A, A.barrier + 1, epoch, 2);
    else if (wait) spin_until<false>(A, A.release + 1, epoch, 4);
  }
  if (crossing) {
    __syncthreads();
    exch(x1);
    __syncthreads();
    if (threadIdx.x == 0) {
      __threadfence();
      atomicExch(A.release + 1, epoch);
    }
  }
  if (threadIdx.x == 0 && wait) __threadfence();
  __syncthreads();
}

// The same barrier on one GPU (no exchange).
__device__ __forceinline__ void grid_barrier_b1(KsbBatchArgs& A, uint32_t& epoch, bool wait) {
  __syncthreads();
  if (threadIdx.x == 0) {
    epoch += gridDim.x;
    if (wait) __threadfence();
    atomicAdd(A.barrier + 1, 1u);
    if (wait) spin_until<false>(A, A.barrier + 1, epoch, 1);
  }
  __syncthreads();
}

// The barrier that ends a pipelined iteration on a sharded table. Nobody needs the full cross-GPU rendezvous here:
//  * a rank other than 0 may start the next evaluation once the commit it ran next to is complete ("commit done":
//    one 8-byte word {pods consumed, number} that block 0 of rank 0 stores into every rank the moment it finishes,
//    after a system-scope fence for the writer warp's copies of the committed rows), and
//  * the committing block may start the next commit once every rank's evaluation results have reached rank 0
//    ("evaluation done": one word per rank, stored by its crossing block after the rank's own grid has arrived).
// Everything else is each GPU's own grid barrier. `x2` numbers these barriers (carried from launch to launch).
template <class Push>
__device__ __forceinline__ void end_barrier_x(KsbBatchArgs& A, uint32_t& epoch, uint32_t& x2, int xb, Push push, bool committer,
                                              int done_idx) {
  __syncthreads();
  x2 += 1u;
  const bool crossing = (int)blockIdx.x == xb;
  char* mine = A.xpeer[A.rank] + A.xo_flag;
  if (threadIdx.x == 0) {
    epoch += gridDim.x;
    if (committer) {
      __threadfence_system();
      const unsigned long long word = ((unsigned long long)(uint32_t)__ldcv(A.group_done + done_idx) << 32) | (unsigned long long)x2;
      for (int q = 1; q < A.n_ranks; ++q)
        *reinterpret_cast<volatile unsigned long long*>(A.xpeer[q] + A.xo_flag + 128 + 8 * (x2 & 1u)) = word;
    } else if (A.rank != 0) {
      __threadfence_system();  // this block's results went into rank 0's memory
    } else {
      __threadfence();
    }
#ifdef KSB_NO_LAST_ARRIVER
    if (false) {
#else
    if (A.rank == 0) {
#endif
      // nothing crosses from here: a plain grid barrier in which the LAST block to arrive (as a rule the committing one)
      // releases the others itself instead of waiting for the crossing block to notice and do it
      const uint32_t before = atomicAdd(A.barrier, 1u);
      if (before + 1u == epoch) {
        __threadfence();
        atomicExch(A.release, epoch);
      } else {
        spin_until<false>(A, A.release, epoch, 4);
      }
    } else {
      atomicAdd(A.barrier, 1u);
      if (crossing) spin_until<false>(A, A.barrier, epoch, 2);
      else spin_until<false>(A, A.release, epoch, 4);
    }
  }
#ifdef KSB_NO_LAST_ARRIVER
  if (crossing) {
    __syncthreads();
    if (A.rank != 0) {
#else
  if (crossing && A.rank != 0) {
    __syncthreads();
    {
#endif
      push();
      __syncthreads();
      if (threadIdx.x == 0) {
        __threadfence_system();
        *reinterpret_cast<volatile uint32_t*>(A.xpeer[0] + A.xo_flag + 64 + 4 * A.rank) = x2;
        const unsigned long long* w = reinterpret_cast<const unsigned long long*>(mine + 128 + 8 * (x2 & 1u));
        unsigned long long v;
        uint32_t spins = 0;
        long long t0 = 0;
        for (;;) {
          asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(w) : "memory");
          if ((uint32_t)v == x2) break;
          if ((++spins & 1023u) == 0) {
            const long long now = clock64();
            if (t0 == 0) t0 = now;
            if (now - t0 > 6000000000ll || ld_acquire((const uint32_t*)A.error_flag) != 0) {
              atomicCAS(A.error_flag, 0, 7);
              break;
            }
          }
        }
        A.l_group_done[done_idx] = (int32_t)(v >> 32);
      }
      __syncthreads();
    }
    if (threadIdx.x == 0) {
      __threadfence();
      atomicExch(A.release, epoch);
    }
  }
  if (threadIdx.x == 0) __threadfence();
  __syncthreads();
  if (committer) {  // the other ranks' results for the next commit
    if (threadIdx.x >= 1 && (int)threadIdx.x < A.n_ranks)
      spin_until<true>(A, reinterpret_cast<const uint32_t*>(mine + 64) + threadIdx.x, x2, 3);
    __syncthreads();
  }
}
__device__ __forceinline__ void grid_barrier(KsbBatchArgs& A, uint32_t& epoch, uint32_t& xepoch, bool wait = true, int xb = 0) {
  grid_barrier_x(A, epoch, xepoch, wait, xb, NoXfer(), NoXfer());
}

// The assume step on the chosen node: NodeInfo.AddPod (schedulercache/node_info.go:329-356) plus
// the per-class counters that stand in for the node's pod list. One warp, lane-parallel where it can
// be. Reads bypass L1: the same row may be committed twice within one group.
__device__ __forceinline__ void commit_pod(const KsbNodeTable& T, int r, const PodView& pv, const uint32_t* picks) {
  const KsbPodHeader* h = pv.h;
  const size_t np = T.n_pad;
  const int lane = threadIdx.x & 31;
  if (lane == 0) {
    T.req_cpu[r] = __ldcg(T.req_cpu + r) + h->add_cpu;
    T.req_mem[r] = __ldcg(T.req_mem + r) + h->add_mem;
    T.req_agpu[r] = __ldcg(T.req_agpu + r) + h->add_agpu;
    T.req_eph[r] = __ldcg(T.req_eph + r) + h->add_eph;
    T.nz_cpu[r] = __ldcg(T.nz_cpu + r) + h->nz_cpu;
    T.nz_mem[r] = __ldcg(T.nz_mem + r) + h->nz_mem;
    T.num_pods[r] = __ldcg(T.num_pods + r) + 1;
  }
  if (lane < KSB_MAX_SCALARS && h->scalar_add[lane] != 0)
    T.req_scalar[lane * np + r] = __ldcg(T.req_scalar + lane * np + r) + h->scalar_add[lane];
  if ((h->flags & PF_HAS_PORTS) && lane < KSB_PORT_WORDS) {
    T.port_any[lane * np + r] = __ldcg(T.port_any + lane * np + r) | pv.ports[3 * KSB_PORT_WORDS + lane];
    T.port_wild[lane * np + r] = __ldcg(T.port_wild + lane * np + r) | pv.ports[4 * KSB_PORT_WORDS + lane];
    T.port_exact[lane * np + r] = __ldcg(T.port_exact + lane * np + r) | pv.ports[5 * KSB_PORT_WORDS + lane];
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
      if (take[s]) T.dev_avail[s * np + r] = __ldcg(T.dev_avail + s * np + r) & ~take[s];
  }
  for (int i = lane; i < h->n_spread_inc; i += 32) {
    uint32_t* p = T.spread + (size_t)pv.spread_inc[i] * np + r;
    *p = __ldcg(p) + 1u;
  }
  for (int i = lane; i < h->n_commit; i += 32) {
    uint64_t c = pv.commits[2 * i];
    int32_t delta = (int32_t)(int64_t)pv.commits[2 * i + 1];
    uint32_t off = (uint32_t)c;
    uint32_t d = T.topo[((uint32_t)(c >> 32) & 0xffu) * np + r];
    if (d) T.tables[off + d] = __ldcg(T.tables + off + d) + delta;
    if ((c >> 40) & 1ull) T.tables[off] = __ldcg(T.tables + off) + delta;  // slot 0: total matching pods anywhere
  }
}

__device__ __forceinline__ int32_t warp_max(int32_t v) { return __reduce_max_sync(0xffffffffu, v); }
__device__ __forceinline__ int32_t warp_min(int32_t v) { return __reduce_min_sync(0xffffffffu, v); }

// Mutable (and a few static) columns of one node row, kept in shared memory by phase 3 for every row
// chosen earlier in the current group.
//   q[0..3] alloc cpu/mem/alpha-gpu/ephemeral   q[4..7] requested cpu/mem/alpha-gpu/ephemeral
//   q[8..9] nonzero cpu/mem   q[10..13] scalar alloc   q[14..17] scalar requested
//   q[18..21] port_any   q[22..25] port_wild   q[26..29] port_exact   q[30] preferAvoid set
struct RowCache {
  int64_t q[31];
  int32_t apods, npods;
  uint32_t flags, zone;
  uint32_t avail[KSB_MAX_RES];
  int32_t sscore;  // the node's NodeLabelPriority sum (static; the u32 column after dev_avail)
  int32_t row, pad;
  // per extended-resource entry slot: which devices of the row match the class mask interned as `mid` (0: none kept)
  uint32_t mm[KSB_MAX_EXT];
  uint16_t mid[KSB_MAX_EXT];
  double rcp[2];  // fast_rcp of q[0], q[1] (allocatable cpu / memory), set when the entry is filled
};

// reason bits that depend on columns a bind changes; every other bit of a row is constant inside a group
constexpr uint32_t kDynMask = 0x1ffu | (1u << RB_PORTS) | (0xfu << RB_EXT0) | (1u << RB_EXT_LIMITS);

// select_devices on a row-cache entry: the per-device class test is taken from the entry when an earlier pod of the
// group with the same interned class mask already ran it on this row. Same results as select_devices.
// `keep` false: the entry is read by several lanes at once (one pod per lane), so nothing is written back to it.
__device__ __forceinline__ uint32_t select_devices_cached(const KsbNodeTable& T, RowCache* c, const PodView& pv, uint32_t* picks,
                                                          bool keep = true) {
  const KsbPodHeader* h = pv.h;
  uint32_t avail[KSB_MAX_RES];
#pragma unroll
  for (int s = 0; s < KSB_MAX_RES; ++s) avail[s] = c->avail[s];
  for (int i = 0; i < h->n_ext; ++i) {
    if ((h->flags & PF_EXT_LIMITS_ERROR) && i == h->ext_err_entry) return 1u << RB_EXT_LIMITS;
    const int rs = h->ext[i].rslot;
    const int num = h->ext[i].num;
    if (rs == 255 || !(c->flags & (256u << rs))) return 1u << (RB_EXT0 + i);  // available[rName] missing
    uint32_t sel = 0;
    if (num > 0) {
      const uint16_t id = h->ext[i].cls_id;
      uint32_t match;
      if (id != 0 && c->mid[i] == id) {
        match = c->mm[i];
      } else {
        match = device_match_mask(T, c->row, rs, pv.extcls + i * KSB_CLASS_WORDS);
        if (keep) {
          c->mm[i] = match;
          c->mid[i] = id;
        }
      }
      const uint32_t m = avail[rs] & match;
      if (__popc(m) < num) return 1u << (RB_EXT0 + i);
      // the first `num` matching available devices in ascending id order
      const int last = nth_set_from_bottom(m, num - 1);
      sel = m & (0xffffffffu >> (31 - last));
    } else if (num != 0) {
      return 1u << (RB_EXT0 + i);  // a negative count never equals the number taken
    }
    picks[i] = sel;
    avail[rs] &= ~sel;
  }
  return 0;
}

// The part of eval_node that can change while a group is committed, computed from the row cache.
// Must stay in step with eval_node (the group-size tests compare the two paths bit for bit).
template <bool POL>
__device__ __forceinline__ void eval_cached(const KsbNodeTable& T, RowCache& c, const PodView& pv,
                                            uint32_t static_reasons, uint32_t* reasons_out, int32_t* base_out,
                                            uint32_t* picks, bool want_score, bool keep = true) {
  const KsbPodHeader* h = pv.h;
  const uint32_t pf = h->flags;
  uint32_t reasons = static_reasons;
  if (c.npods + 1 > c.apods) reasons |= 1u << RB_PODS;
  if (pf & PF_HAS_REQUEST) {
    if (c.q[0] < h->req_cpu + c.q[4]) reasons |= 1u << RB_CPU;
    if (c.q[1] < h->req_mem + c.q[5]) reasons |= 1u << RB_MEM;
    if (c.q[2] < h->req_agpu + c.q[6]) reasons |= 1u << RB_AGPU;
    if (c.q[3] < h->req_eph + c.q[7]) reasons |= 1u << RB_EPH;
    const uint32_t sm = h->scalar_mask;
#pragma unroll
    for (int s = 0; s < KSB_MAX_SCALARS; ++s)
      if ((sm >> s) & 1u)
        if (c.q[10 + s] < h->scalar_req[s] + c.q[14 + s]) reasons |= 1u << (RB_SCALAR0 + s);
  }
  if (pf & PF_HAS_PORTS) {
    uint64_t hit = 0;
#pragma unroll
    for (int w = 0; w < KSB_PORT_WORDS; ++w) {
      hit |= (uint64_t)c.q[18 + w] & pv.ports[w];
      hit |= (uint64_t)c.q[22 + w] & pv.ports[KSB_PORT_WORDS + w];
      hit |= (uint64_t)c.q[26 + w] & pv.ports[2 * KSB_PORT_WORDS + w];
    }
    if (hit) reasons |= 1u << RB_PORTS;
  }
  if (reasons == 0 && h->n_ext) {
    uint32_t local[KSB_MAX_EXT];
    reasons |= select_devices_cached(T, &c, pv, picks ? picks : local, keep);
  }
  *reasons_out = reasons;
  *base_out = 0;
  if (!want_score) return;
  const int64_t acpu = c.q[0], amem = c.q[1];
  const int64_t cpu = h->nz_cpu + c.q[8];
  const int64_t mem = h->nz_mem + c.q[9];
  int32_t lr, bra;
  lr_bra(cpu, mem, acpu, amem, c.rcp[0], c.rcp[1], &lr, &bra);
  uint64_t avoid_hit = (uint64_t)c.q[30] & h->avoid_mask;
  if (POL && T.tw > 1)  // the words above 0 of the preferAvoid set: static columns, read in place
    for (int w = 0; w < T.tw - 1; ++w) avoid_hit |= T.avoid[(size_t)(w + 1) * T.n_pad + c.row] & pv.tw_ext[2 * (T.tw - 1) + w];
  const int32_t av = avoid_hit ? 0 : 10;
  *base_out = T.w_lr * lr + T.w_bra * bra + T.w_avoid * av;
  if constexpr (POL) *base_out += policy_score(T, h, cpu, mem, acpu, amem, c.sscore, c.row);
}

// One warp: load row r's mutable columns into a cache entry (one round of independent L2 loads). In two steps, so
// that a warp filling two entries has both rows' loads in flight before it consumes either.
__device__ __forceinline__ void cache_fill_issue(const KsbNodeTable& T, int r, int lane, int64_t* v_out, uint32_t* v2_out) {
  const size_t np = T.n_pad;
  // columns of one family are consecutive, n_pad apart, in the order of q[] (device_types.h): lanes 0..17 the
  // int64 family from alloc_cpu on, lanes 18..29 the twelve port words, lane 30 preferAvoid
  const int64_t* p = lane < 18   ? T.alloc_cpu + (size_t)lane * np + r
                     : lane == 30 ? reinterpret_cast<const int64_t*>(T.avoid) + r  // word 0 of the preferAvoid set
                                  : reinterpret_cast<const int64_t*>(T.port_any) + (size_t)(lane - 18) * np + r;
  // lanes 0..6: the u32 family from alloc_pods on (alloc_pods, num_pods, flags, zone, dev_avail[2], static_score)
  const uint32_t* p2 = reinterpret_cast<const uint32_t*>(T.alloc_pods) + (size_t)(lane < 5 + KSB_MAX_RES ? lane : 0) * np + r;
  int64_t v = 0;
  uint32_t v2 = 0;
  if (lane < 31) v = __ldcg(p);
  if (lane < 5 + KSB_MAX_RES) v2 = __ldcg(p2);
  *v_out = v;
  *v2_out = v2;
}
__device__ __forceinline__ void cache_fill_store(RowCache* c, int r, int lane, int64_t v, uint32_t v2) {
  if (lane < 31) c->q[lane] = v;
  // apods, npods, flags, zone, avail[0..1], sscore are consecutive 32-bit fields of the entry
  if (lane < 5 + KSB_MAX_RES) reinterpret_cast<uint32_t*>(&c->apods)[lane] = v2;
  if (lane == 31) c->row = r;
  if (lane >= 24 && lane < 24 + KSB_MAX_EXT) c->mid[lane - 24] = 0;
  if (lane < 2) c->rcp[lane] = fast_rcp((double)v);
}
__device__ __forceinline__ void cache_fill(const KsbNodeTable& T, int r, RowCache* c, int lane) {
  int64_t v;
  uint32_t v2;
  cache_fill_issue(T, r, lane, &v, &v2);
  cache_fill_store(c, r, lane, v, v2);
}

// What the decider warp hands to the writer warp for one pod: the chosen row's mutable columns after the
// assume step (a copy of the row-cache entry) and the result record.
struct CommitRec {
  int64_t q[31];
  int32_t npods;
  uint32_t avail[KSB_MAX_RES];
  int32_t row;   // local row on shard `rank`, -1: nothing to write
  int32_t rank;
  KsbResult res;
};

// Decider side of the assume step (NodeInfo.AddPod, schedulercache/node_info.go:329-356): the row-cache entry and,
// in the same pass, the record the writer warp copies to the table. Lane l owns q[l]; the header's int64 fields
// are read as an array (req 0..3, nz 4..5, add 6..9, scalar_req 10..13, scalar_add 14..17).
__device__ __forceinline__ void commit_and_record(RowCache* c, CommitRec* rec, const PodView& pv, const uint32_t* picks,
                                                  int row, int rank, int lane) {
  const KsbPodHeader* h = pv.h;
  const int64_t* hq = reinterpret_cast<const int64_t*>(h);
  if (lane < 31) {
    const int ai = (lane >= 4 && lane < 8) ? lane + 2 : ((lane == 8 || lane == 9) ? lane - 4 : ((lane >= 14 && lane < 18) ? lane : -1));
    int64_t v = c->q[lane];
    if (ai >= 0) v += hq[ai];
    if ((h->flags & PF_HAS_PORTS) && lane >= 18 && lane < 30) v = (int64_t)((uint64_t)v | pv.ports[3 * KSB_PORT_WORDS + (lane - 18)]);
    c->q[lane] = v;
    rec->q[lane] = v;
  } else {
    const int32_t npods = c->npods + 1;
    c->npods = npods;
    rec->npods = npods;
    uint32_t take[KSB_MAX_RES];
#pragma unroll
    for (int s = 0; s < KSB_MAX_RES; ++s) take[s] = 0;
    for (int i = 0; i < h->n_ext; ++i) {
      const int rs = h->ext[i].rslot;
#pragma unroll
      for (int s = 0; s < KSB_MAX_RES; ++s)  // no dynamic index: the array stays in registers
        if (rs == s) take[s] |= picks[i];
    }
#pragma unroll
    for (int s = 0; s < KSB_MAX_RES; ++s) {
      const uint32_t av = c->avail[s] & ~take[s];
      c->avail[s] = av;
      rec->avail[s] = av;
    }
    rec->row = row;
    rec->rank = rank;
  }
}

// Writer side: the row's mutable columns from the record to the table (NodeInfo.AddPod, node_info.go:329-356,
// and the bind echo of device ids, pkg/registry/core/pod/storage/storage.go:180-187).
__device__ __forceinline__ void write_through(const KsbNodeTable& T, const CommitRec& rec, int r, const KsbPodHeader* h, int lane) {
  const size_t np = T.n_pad;
  // requested / nonzero columns (lanes 4..9) and touched scalar columns (14..17): q[] order = column order
  const bool own = (lane >= 4 && lane < 10) || (lane >= 14 && lane < 18 && h->scalar_add[lane - 14] != 0);
  if (own) T.alloc_cpu[(size_t)lane * np + r] = rec.q[lane];
  if ((h->flags & PF_HAS_PORTS) && lane >= 18 && lane < 30) T.port_any[(size_t)(lane - 18) * np + r] = (uint64_t)rec.q[lane];
  if (lane == 31) {
    T.num_pods[r] = rec.npods;
    if (h->n_ext)
#pragma unroll
      for (int s = 0; s < KSB_MAX_RES; ++s) T.dev_avail[(size_t)s * np + r] = rec.avail[s];
  }
}

// Position of the (n+1)-th set bit of w counted from bit 31 downwards (w has more than n set bits).
__device__ __forceinline__ int nth_set_from_top(uint32_t w, int n) {
  uint32_t r = __brev(w);  // n-th set bit from the bottom of r
  int p = 0;
#pragma unroll
  for (int width = 16; width >= 1; width >>= 1) {
    const int c = __popc(r & ((1u << width) - 1u));
    if (n >= c) {
      n -= c;
      p += width;
      r >>= width;
    }
  }
  return 31 - p;
}

// Rank-select helpers of phase 3, written for one THREAD so that a warp can run one pod per lane.
//
// Pod g's candidates for selectHost are the rows that held its maximum score when the group started (the
// "snapshot holders": per 256-row segment a count and a 256-bit mask from phase 2, reduced here to suffix counts
// `cum`), minus the changed rows that stopped being holders (D), plus the changed rows that became holders (A).
// Changed rows live in the row cache; r0[i] is the number of snapshot holders strictly above the row of cache
// entry i (rows are name-ordered, selectHost walks them in descending order, api/types.go:218-223).

// The rem-th candidate (0-based) in descending row order. Returns the cache entry if it is an added row, else -1
// with *q = its rank among the snapshot holders (deleted ones included in the count).
__device__ __forceinline__ int select_merged(uint32_t A, uint32_t D, int rem, const int32_t* r0, const int32_t* erow, int* q) {
  uint32_t m = A | D;
  int adds = 0, dels = 0;
  while (m) {
    int best = -1, brow = -1;
    for (uint32_t t = m; t; t &= t - 1) {  // the remaining patch with the largest row
      const int b = __ffs(t) - 1;
      const int r = erow[b];
      if (r > brow) {
        brow = r;
        best = b;
      }
    }
    m &= ~(1u << best);
    const int before = r0[best] - dels + adds;  // candidates above this patch's position
    if (rem < before) break;                    // the answer is a snapshot holder above it
    if ((A >> best) & 1u) {
      if (rem == before) return best;
      adds += 1;
    } else {
      dels += 1;
    }
  }
  *q = rem - adds + dels;
  return -1;
}

// dynamic shared memory carve-up
struct Smem {
  KsbPodHeader* hdr;   // [kMaxG]
  uint64_t* pool;      // [kGroupPoolWords]
  int32_t* pool_off;   // [kMaxG] offset of pod g's pool inside `pool`
  int32_t* zone;       // [kMaxG][zone_stride] phase 1: block-local sums; phase 2: global sums; phase 3: t_r0
  int32_t* part;       // [kMaxG][kWarps][8] per-warp partials of the Reduce inputs
  KsbNorms* norm;      // [kMaxG]
  int32_t* max_zone;   // [kMaxG]
  int32_t* wstat;      // [kMaxG][kWarps][3] per-warp (max, count, holder mask) of one segment
  int32_t* cum;        // [kSegCache] block 0: [pod][segment] snapshot holders in segments >= this one
  RowCache* rows;      // [kCache] block 0: rows changed since the snapshot the group was evaluated on
  RowCache* post;      // [kMaxG] block 0: the window's chosen rows after their pod's assume step, not yet final
  int32_t* fe_hist;    // [KSB_REASON_BITS] block 0: FitError histogram correction for changed rows
  int32_t* r0;         // [kMaxG][kCache] block 0: snapshot holders of pod g above the row of cache entry i
  int32_t* t_r0;       // [kMaxG][kMaxG] block 0: the same for the window's rows (aliases `zone`)
  CommitRec* recs;     // [kMaxG] block 0: decider -> writer hand-off (aliases part + wstat)
};
constexpr int kSmemParts = 13;

__device__ __host__ __forceinline__ size_t smem_layout(size_t* offs, int zone_stride) {
  size_t off = 0;
  int i = 0;
  auto take = [&](size_t bytes) {
    offs[i++] = off;
    off += (bytes + 15) & ~(size_t)15;
  };
  take(sizeof(KsbPodHeader) * kMaxG);
  take((size_t)kGroupPoolWords * 8);
  take(kMaxG * 4);
  take((size_t)kMaxG * zone_stride * 4);  // zone; phase 3 reuses it as t_r0 (zone_stride >= 64 > kMaxG)
  take((size_t)kMaxG * kWarps * 8 * 4);   // part: phase 1/2 reduction scratch ...
  take((size_t)kMaxG * kWarps * 3 * 4);   // wstat: ... contiguous with part; phase 3 reuses both as S.recs
  take(sizeof(KsbNorms) * kMaxG);
  take(kMaxG * 4);
  take((size_t)kSegCache * 4);
  take(sizeof(RowCache) * kCache);
  take(sizeof(RowCache) * kMaxG);
  take((size_t)KSB_REASON_BITS * 4);
  take((size_t)kMaxG * kCache * 4);
  static_assert(sizeof(CommitRec) * kMaxG <= (size_t)kMaxG * kWarps * 11 * 4, "commit records alias part + wstat");
  return off;
}

__device__ __forceinline__ Smem carve(unsigned char* base, int zone_stride) {
  size_t o[kSmemParts];
  smem_layout(o, zone_stride);
  Smem s;
  s.hdr = (KsbPodHeader*)(base + o[0]);
  s.pool = (uint64_t*)(base + o[1]);
  s.pool_off = (int32_t*)(base + o[2]);
  s.zone = (int32_t*)(base + o[3]);
  s.part = (int32_t*)(base + o[4]);
  s.wstat = (int32_t*)(base + o[5]);
  s.norm = (KsbNorms*)(base + o[6]);
  s.max_zone = (int32_t*)(base + o[7]);
  s.cum = (int32_t*)(base + o[8]);
  s.rows = (RowCache*)(base + o[9]);
  s.post = (RowCache*)(base + o[10]);
  s.fe_hist = (int32_t*)(base + o[11]);
  s.r0 = (int32_t*)(base + o[12]);
  s.t_r0 = (int32_t*)(base + o[3]);
  s.recs = (CommitRec*)(base + o[4]);
  return s;
}

}  // namespace

// PIPE: the pipelined schedule (see `pipe` below) as a separate instantiation, so that the plain one carries none
// of its state in registers.
// POL: a Policy enabled plugins outside the DefaultProvider set (ksb_policy_set), or a taint / preferAvoid set has grown
// beyond 64 entries; the default instantiations carry none of that code.
template <bool PIPE, bool MULTI, bool POL>
__global__ void __launch_bounds__(kThreads) ksb_schedule_batch_kernel(KsbBatchArgs A) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  const Smem S = carve(smem_raw, A.zone_stride);
  __shared__ int32_t s_gc;            // pods in the current group
  __shared__ int32_t s_consumed;      // pods of the group that were committed
  __shared__ KsbPeer s_peers[KSB_MAX_RANKS];  // rank 0, block 0: the shard table, read once
  // one rank: the only shard is this GPU's own, described by the kernel parameters themselves
#define PEER(x) (MULTI ? s_peers[(x)] : A.self)
  __shared__ volatile int32_t s_ready[kMaxG]; // deciders -> writer: record g is complete
  __shared__ volatile int32_t s_stop;         // deciders -> writer: pods decided in this group (-1: still deciding)
  // block 0, phase 3 (see there)
  __shared__ int32_t s_gmax0[kMaxG], s_k0[kMaxG];  // per pod: maximum score at group start and the rows holding it
  __shared__ int32_t s_lrel[kMaxG];                // per pod: earlier pods of the group that bump lastNodeIndex
  __shared__ uint32_t s_snap[kMaxG], s_now[kMaxG]; // per pod: cache entries whose row held / holds its maximum
  __shared__ int32_t s_erow[kCache];               // global row of cache entry i
  __shared__ int32_t s_pick[kMaxG];                // window: the row each pod picked
  __shared__ uint32_t t_dirty[kMaxG], t_now[kMaxG], t_snap[kMaxG], t_spec[kMaxG];  // window: column-pass outcomes per pick
  __shared__ int32_t t_idx[kMaxG];                 // window: cache entry of the picked row, -1 none, -2 no pick
#if KSB_FINE_TIMING == 1
  __shared__ uint32_t t_cause[4][kMaxG];
  for (int i = threadIdx.x; i < 4 * kMaxG; i += kThreads) (&t_cause[0][0])[i] = 0;
#endif
  __shared__ volatile uint32_t s_special;          // pods the one-pod path must decide
  __shared__ uint32_t s_prio, s_bad0;              // pods ranked by score; window pods whose pick failed
  __shared__ volatile int32_t s_ncr, s_cut;        // rows in the cache; the group ends early
  __shared__ uint32_t s_b1_target;             // block 0, pipelined: barrier count at which the evaluation has read the table

  const KsbNodeTable& T = A.table;
  const int tid = threadIdx.x;
  const int lane = tid & 31;
  const int warp = tid >> 5;
  const int nblk = gridDim.x;
  const int b = blockIdx.x;
  const int nz = T.n_zones;
  const int zs = A.zone_stride;
  const int nseg = A.nseg;
  const int G = A.group_size;
  const size_t np = T.n_pad;
  uint32_t epoch = 0, xepoch = 0;
  uint32_t epoch1 = 0;  // arrivals at the barrier between phases 1 and 2 (its own counter, see reduce_barrier_x)
  constexpr bool multi = MULTI;
  uint64_t last_index = *A.last_node_index;  // used by warp 0 of block 0 only
  int gi = 0;
  // Pipelined launches (A.pipeline, one rank): in iteration k block 0 commits the group evaluated in iteration k-1
  // ("commit side") while the other blocks evaluate the next one ("evaluation side") against the table as it stood
  // before that commit; the commit's table writes are held back until the evaluation has read the table.
  constexpr bool pipe = PIPE;
  constexpr int xb = PIPE ? 1 : 0;  // the block that runs the cross-GPU leg of the barriers (block 0 of rank 0 commits)
  const bool committer = b == 0 && A.rank == 0;
  const bool lag = pipe && committer;  // this block works on the commit-side group
  int ev_p0 = 0, ev_slot = 0;          // evaluation side: first pod and scratch half of the group
  bool cm_valid = false, cm_carry = false;  // commit side: a group is pending; it was evaluated next to another commit
  int cm_p0 = 0, cm_G = 0, cm_slot = 0, cm_floor = 0;
  // The block that shares an SM with block 0 would compete with the decider warp for issue slots: it sits the
  // evaluation out. group_done[4] = block 0's SM, group_done[5] = index of that block (0: none).
  int shun_b = 0;
  uint32_t x2 = 0, x1 = 0;  // numbers of the iteration-ending barriers (sharded, pipelined) and of the Reduce exchanges
  if (MULTI) {  // the cross-GPU flags are never reset: their numbers are carried from launch to launch
    xepoch = A.xstate[0];
    x2 = A.xstate[1];
    x1 = A.xstate[2];
  }
  if (pipe) {
    // (sharded: only rank 0 has a committing block, so only its blocks take part; every rank runs the barriers)
    uint32_t smid;
    asm volatile("mov.u32 %0, %%smid;" : "=r"(smid));
    if (A.rank == 0 && b == 0 && tid == 0) {
      A.group_done[4] = (int32_t)smid;
      A.group_done[5] = 0;
    }
    grid_barrier(A, epoch, xepoch, true, xb);
    if (A.rank == 0 && b != 0 && tid == 0 && (int32_t)smid == __ldcv(A.group_done + 4)) A.group_done[5] = b;
    grid_barrier(A, epoch, xepoch, true, xb);
    shun_b = A.rank == 0 ? __ldcv(A.group_done + 5) : 0;
    if (nblk < 4) shun_b = 0;
  }
  int cm_gid = 0;    // block 0: sequence number of the group being committed (row-cache tags)
#ifndef KSB_WCAP_MAX
#define KSB_WCAP_MAX kDeciders
#endif
  // One column per deciding warp and round: wider windows (two columns in a row per warp) decided more pods per round but
  // fewer per microsecond (cfg4 at 50,000 nodes: 2.87 -> 2.70 us/pod; about one (pick, later pod) pair in thirty is dirty)
  const int wcap = KSB_WCAP_MAX;
  // block 0, phase 3: window size; grows while whole windows come out right, shrinks when they do not

  if (MULTI && blockIdx.x == 0 && A.rank == 0) {
    const uint32_t* src = reinterpret_cast<const uint32_t*>(A.peers);
    uint32_t* dst = reinterpret_cast<uint32_t*>(s_peers);
    for (int i = threadIdx.x; i < A.n_ranks * (int)(sizeof(KsbPeer) / 4); i += kThreads) dst[i] = src[i];
    __syncthreads();
  }
  if (multi) {
    // sharded: rank 0 clears the shared reduction targets itself and an opening barrier publishes that before any
    // rank adds to them
    if (A.rank == 0 && blockIdx.x == 0) {
      for (int i = threadIdx.x; i < 2 * kMaxG * 8; i += kThreads)
        reinterpret_cast<int32_t*>(A.norms)[i] = ((i & 7) == 7) ? -1 : 0;
      for (int i = threadIdx.x; i < 2 * kMaxG * A.zone_stride; i += kThreads) A.zone_sum[i] = 0;
      for (int i = threadIdx.x; i < 2 * kMaxG * KSB_REASON_BITS; i += kThreads) A.ghist[i] = 0;
      if (threadIdx.x < 2) A.group_done[threadIdx.x] = 0;
    }
    if (A.rank != 0 && (int)blockIdx.x == xb) {  // this GPU's accumulators
      for (int i = threadIdx.x; i < 2 * kMaxG * 8; i += kThreads)
        reinterpret_cast<int32_t*>(A.acc_norms)[i] = ((i & 7) == 7) ? -1 : 0;
      for (int i = threadIdx.x; i < 2 * kMaxG * A.zone_stride; i += kThreads) A.acc_zone_sum[i] = 0;
      for (int i = threadIdx.x; i < 2 * kMaxG * KSB_REASON_BITS; i += kThreads) A.acc_ghist[i] = 0;
      if (threadIdx.x < 8) A.l_group_done[threadIdx.x] = 0;
    }
    grid_barrier(A, epoch, xepoch, true, xb);
  }
  // measurement state lives in shared memory (thread 0 of each block updates it; block 0's copy is reported): kept
  // in registers it would cost every thread of the kernel some thirty of them
  __shared__ long long s_tk[40];  // [0..7] cycles per stage, [8] time mark, [9] groups, [10] groups cut short
  if (tid < 40) s_tk[tid] = 0;
  if (tid == 0) s_ncr = 0;
  __syncthreads();
  if (tid == 0) s_tk[8] = clock64();
#define KSB_TICK(i)                                               \
  do {                                                            \
    if (threadIdx.x == 0) {                                       \
      const long long now_ = clock64();                           \
      s_tk[i] += now_ - s_tk[8];                                  \
      s_tk[8] = now_;                                             \
      if (blockIdx.x == 0) A.counters[14] = ((uint64_t)gi << 16) | (i); \
    }                                                             \
  } while (0)
  while (ev_p0 < A.n_pods || cm_valid) {
    const bool ev_valid = ev_p0 < A.n_pods;
    const bool have_group = lag ? cm_valid : ev_valid;  // this block has a group to stage
    const int p0 = lag ? cm_p0 : ev_p0;
    const int slot = lag ? cm_slot : ev_slot;
    // pods may join the group only if every counter they read was last written before `dep_floor`: the start of
    // the group itself, or of the group being committed next to this evaluation
    const int dep_floor = lag ? cm_floor : ((pipe && cm_valid) ? cm_p0 : ev_p0);
    const size_t half = pipe ? (size_t)slot * (size_t)G : 0;  // first pod index of this group's scratch half
    // this group's half of the per-(pod, row) and per-(pod, segment) scratch arrays (recomputed where used: held in
    // registers across the loop body they cost more than the multiply)
#define raw_s (A.raw + half * np)
#define score_s (A.score + half * np)
#define reasons_s (A.reasons + half * np)
    // ---- stage the group's requirement blocks; every block derives the same group extent
    if (!have_group) {
      if (tid == 0) s_gc = 0;
      __syncthreads();
    } else {
      const int avail = min(G, A.n_pods - p0);
      const uint4* src = reinterpret_cast<const uint4*>(A.headers + p0);
      uint4* dst = reinterpret_cast<uint4*>(S.hdr);
      const int n16 = avail * (int)(sizeof(KsbPodHeader) / 16);
      for (int i = tid; i < n16; i += kThreads) dst[i] = src[i];
      __syncthreads();
      if (warp == 0) {
        bool ok = false;
        if (lane < avail) {
          const KsbPodHeader* h = S.hdr + lane;
          int rel_end = (int)(h->pool_off - S.hdr[0].pool_off) + pool_words(h, T.W, T.tw);
          ok = rel_end <= kGroupPoolWords && __ldg(A.dep + p0 + lane) < dep_floor;
          S.pool_off[lane] = (int32_t)(h->pool_off - S.hdr[0].pool_off);
        }
        uint32_t bal = __ballot_sync(0xffffffffu, ok);
        int gc = __ffs(~bal) - 1;  // leading run of admissible pods
        if (gc < 0) gc = 32;
        // pipelined: 0 when the first pod itself depends on the group being committed (the evaluation waits a turn)
        if (lane == 0) s_gc = pipe ? min(gc, avail) : max(1, min(gc, avail));
      }
      __syncthreads();
      const int gc = s_gc;
      const KsbPodHeader* hl = S.hdr + max(0, gc - 1);
      const int words = gc > 0 ? S.pool_off[gc - 1] + pool_words(hl, T.W, T.tw) : 0;
      const uint64_t* psrc = A.pool + S.hdr[0].pool_off;
      for (int i = tid; i < words; i += kThreads) S.pool[i] = psrc[i];
      for (int i = tid; i < gc * zs; i += kThreads) S.zone[i] = 0;
      for (int i = tid; i < gc * kWarps * 8; i += kThreads) {
        int f = i & 7;
        S.part[i] = (f == 7) ? -1 : 0;  // fit, max_na, max_tt, max_sc, ipa_max, ipa_min, have_zones, one_row
      }
      __syncthreads();
    }
    const int Gc = s_gc;
    // Block 0 learns the evaluated group's size from here. Both published numbers are indexed by the parity of the
    // iteration, not of the group: a block that runs ahead into the next iteration then never overwrites a value that a
    // slower block has yet to read (scratch halves do not alternate in an iteration that evaluates nothing).
    if (pipe && b == 1 && tid == 0) A.l_group_done[2 + (gi & 1)] = Gc;  // (every rank derives the same number)
    const bool do_eval = !lag && Gc > 0 && !(shun_b != 0 && b == shun_b) && !(pipe && b == 0);
    KSB_TICK(0);
    // Reduce inputs of the group: accumulated per GPU (acc_*: the targets themselves on rank 0), pushed to rank 0's
    // targets at the barrier after phase 1, read back as totals (tot_*: again the targets themselves on rank 0)
    // (recomputed where used: held in registers across the loop body they would be spilled)
#define cur_norms (A.norms + (size_t)slot * kMaxG)
#define cur_zsum (A.zone_sum + (size_t)slot * kMaxG * zs)
#define cur_acc_norms (A.acc_norms + (size_t)slot * kMaxG)
#define cur_acc_zsum (A.acc_zone_sum + (size_t)slot * kMaxG * zs)
#define cur_tot_norms (A.tot_norms + (size_t)slot * kMaxG)
#define cur_tot_zsum (A.tot_zone_sum + (size_t)slot * kMaxG * zs)
#define cur_acc_ghist (A.acc_ghist + (size_t)slot * kMaxG * KSB_REASON_BITS)

    // work split for phases 1 and 2: the (segment, pod) pairs, pod index fastest, are dealt to the blocks in
    // contiguous runs of equal length, so every block is busy whatever the ratio of segments to blocks
    const int items = nseg * Gc;
    // pipelined: block 0 does not evaluate, nor does the block on its SM
    const int eb = pipe ? b - 1 - ((shun_b != 0 && b > shun_b) ? 1 : 0) : b;
    const int enb = pipe ? nblk - 1 - (shun_b != 0 ? 1 : 0) : nblk;
    // (items <= 2048 segments x 32 pods and a few hundred blocks: the products fit 32 bits)
    const int it0 = do_eval ? (int)((uint32_t)items * (uint32_t)eb / (uint32_t)enb) : 0;
    const int it1 = do_eval ? (int)((uint32_t)items * (uint32_t)(eb + 1) / (uint32_t)enb) : 0;
    const bool active = it1 > it0;
    const int seg_lo = active ? it0 / Gc : 0, seg_hi = active ? (it1 - 1) / Gc : -1;

    // ---- phase 1: predicates + Map
    if (active) {
      for (int seg = seg_lo; seg <= seg_hi; ++seg) {
        const int r = seg * kSeg + tid;
        const bool live = r < A.eval_n;
        const int g_lo = seg == seg_lo ? it0 - seg_lo * Gc : 0;
        const int g_hi = seg == seg_hi ? it1 - 1 - seg_hi * Gc : Gc - 1;
        RowCommon rc;
        if (live) load_row_common<false>(T, r, rc);
        size_t at = (half + (size_t)g_lo) * np + (size_t)r;  // this thread's slot in the (pod, row) stashes
        for (int g = g_lo; g <= g_hi; ++g, at += np) {
          const KsbPodHeader* h = S.hdr + g;
          int32_t fit = 0, mna = 0, mtt = 0, msc = 0, ipa = 0, hz = 0;
          if (live && !(h->flags & PF_SCHED_ERROR)) {
            const PodView pv = make_view(h, S.pool + S.pool_off[g], T.W);
            EvalOut o;
            eval_node<false, POL>(T, r, rc, pv, o, nullptr);
            int4 raw;
            if (o.reasons == 0) {
              fit = 1;
              mna = o.na;
              mtt = o.tt;
              msc = (int32_t)o.sc;
              ipa = o.ipa;
              if (o.zone) {
                hz = 1;
                if (o.sc) atomicAdd(&S.zone[g * zs + o.zone], (int32_t)o.sc);
              }
              raw = make_int4(o.base, o.na, (o.tt << 24) | (int32_t)(o.sc & 0xffffffu), o.ipa);
            } else {
              raw = make_int4(INT32_MIN, 0, 0, 0);
            }
            A.raw[at] = raw;
            A.reasons[at] = o.reasons;
            if constexpr (MULTI) {  // what phase 3 reads of a changed row lives on rank 0, at the global row
              const size_t gat = (half + (size_t)g) * (size_t)A.gs_np + (size_t)(A.gs_row0 + r);
              A.gs_base0[gat] = raw.x;
              if (A.rank != 0) A.gs_reasons[gat] = o.reasons;
            }
          }
          // Reduce inputs of this warp's 32 rows: ballots for the counts and flags, a redux only for the values the pod
          // can have at all (uniform per pod); this warp's own slot, so no atomics
          const uint32_t fb = __ballot_sync(0xffffffffu, fit != 0);
          if (fb) {
            int32_t* pp = S.part + (g * kWarps + warp) * 8;
            const bool any_z = __any_sync(0xffffffffu, hz != 0);
            int32_t vna = 0, vtt = 0, vsc = 0, vmax = 0, vmin = 0;
            if (h->n_pref) vna = warp_max(mna);
            if (h->soft_untol || (POL && T.tw > 1)) vtt = warp_max(mtt);
            if (h->spread_col >= 0) vsc = warp_max(msc);
            if (h->n_score) {
              vmax = warp_max(max(0, ipa));
              vmin = warp_min(min(0, ipa));
            }
            if (lane == 0) {
              pp[0] += __popc(fb);
              pp[1] = max(pp[1], vna);
              pp[2] = max(pp[2], vtt);
              pp[3] = max(pp[3], vsc);
              pp[4] = max(pp[4], vmax);
              pp[5] = min(pp[5], vmin);
              pp[6] = max(pp[6], any_z ? 1 : 0);
              pp[7] = max(pp[7], T.row_base + seg * kSeg + warp * 32 + (31 - __clz(fb)));  // highest fitting row
            }
          }
        }
      }
    }
    __syncthreads();
    if (active) {
      for (int i = tid; i < Gc * 8; i += kThreads) {
        const int g = i >> 3, f = i & 7;
        int32_t v = (f == 7) ? -1 : 0;
        for (int w = 0; w < kWarps; ++w) {
          int32_t x = S.part[(g * kWarps + w) * 8 + f];
          if (f == 0) v += x;
          else if (f == 5) v = min(v, x);
          else v = max(v, x);
        }
        int32_t* dst = reinterpret_cast<int32_t*>(cur_acc_norms + g) + f;
        if (f == 0) { if (v) atomicAdd(dst, v); }
        else if (f == 5) { if (v) atomicMin(dst, v); }
        else if (f == 7) { if (v >= 0) atomicMax(dst, v); }
        else { if (v) atomicMax(dst, v); }
      }
      for (int g = warp; g < Gc; g += kWarps)
        for (int z = lane; z < nz; z += 32)
          if (S.zone[g * zs + z]) atomicAdd(&cur_acc_zsum[g * zs + z], S.zone[g * zs + z]);
    }
    KSB_TICK(1);
    {
      // Sharded: once this GPU's grid has arrived, its crossing block sends the GPU's partial Reduce inputs of the group
      // (at most 32 pods x 8 values, and the interned zones' sums) to every rank and combines the totals from what
      // every rank sent here. Each value travels as one 8-byte word {value, number of this exchange}: a word is
      // complete when its number matches, so there is no fence, no flag and no NVLink read on the path; the inbox
      // alternates between two halves, so a word is overwritten only after its reader has sent the exchange after.
      auto exch1 = [&](uint32_t tag) {
        if constexpr (MULTI) {
          const int R = A.n_ranks;
          const size_t hslot = (size_t)(tag & 1u) * KSB_MAX_RANKS;  // inbox half of this exchange
          auto word = [&](int32_t v) { return ((unsigned long long)(uint32_t)v << 32) | (unsigned long long)tag; };
          const int nzv = Gc * nz;
          if (tid < Gc * 8) {
            int32_t* src = reinterpret_cast<int32_t*>(cur_acc_norms) + tid;
            const int32_t vn = __ldcg(src);
            *src = (tid & 7) == 7 ? -1 : 0;
            for (int q = 0; q < R; ++q)
              reinterpret_cast<volatile unsigned long long*>(A.xpeer[q] + A.xo_norms)[(hslot + A.rank) * kMaxG * 8 + tid] = word(vn);
          }
          for (int i0 = 0; i0 < nzv; i0 += 4 * kThreads) {
            int32_t vz[4] = {0, 0, 0, 0};
#pragma unroll
            for (int j = 0; j < 4; ++j) {
              const int i = i0 + tid + j * kThreads;
              if (i < nzv) {
                const int at = (i / nz) * zs + (i % nz);
                vz[j] = __ldcg(cur_acc_zsum + at);
                if (vz[j]) cur_acc_zsum[at] = 0;
              }
            }
            for (int q = 0; q < R; ++q) {
              volatile unsigned long long* dz =
                  reinterpret_cast<volatile unsigned long long*>(A.xpeer[q] + A.xo_zone) + (hslot + A.rank) * (size_t)kMaxG * zs;
#pragma unroll
              for (int j = 0; j < 4; ++j) {
                const int i = i0 + tid + j * kThreads;
                if (i < nzv) dz[(i / nz) * zs + (i % nz)] = word(vz[j]);
              }
            }
          }
          // ... and what every rank sent here
          const char* inbox = A.xpeer[A.rank];
          long long t0 = 0;
          uint32_t spins = 0;
          bool dead = false;
          auto take = [&](const unsigned long long* w) -> int32_t {
            unsigned long long v;
            for (;;) {
              asm volatile("ld.relaxed.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(w) : "memory");
              if ((uint32_t)v == tag || dead) break;
              if ((++spins & 1023u) == 0) {
                const long long now = clock64();
                if (t0 == 0) t0 = now;
                if (now - t0 > 6000000000ll || ld_acquire((const uint32_t*)A.error_flag) != 0) {
                  atomicCAS(A.error_flag, 0, 3);
                  dead = true;
                }
              }
            }
            return (int32_t)(v >> 32);
          };
          if (tid < Gc * 8) {
            const int f = tid & 7;
            const unsigned long long* base = reinterpret_cast<const unsigned long long*>(inbox + A.xo_norms) + hslot * kMaxG * 8 + tid;
            int32_t v = take(base);
            for (int q = 1; q < R; ++q) {
              const int32_t x = take(base + (size_t)q * kMaxG * 8);
              if (f == 0) v += x;
              else if (f == 5) v = min(v, x);
              else v = max(v, x);
            }
            reinterpret_cast<int32_t*>(cur_tot_norms)[tid] = v;
          }
          for (int i = tid; i < nzv; i += kThreads) {
            const int at = (i / nz) * zs + (i % nz);
            const unsigned long long* base = reinterpret_cast<const unsigned long long*>(inbox + A.xo_zone) + hslot * (size_t)kMaxG * zs + at;
            int32_t v = 0;
            for (int q = 0; q < R; ++q) v += take(base + (size_t)q * kMaxG * zs);
            cur_tot_zsum[at] = v;
          }
        }
      };
      if constexpr (MULTI) reduce_barrier_x(A, epoch1, x1, !lag, xb, exch1);
      else grid_barrier_b1(A, epoch1, !lag);
    }
    if (lag && tid == 0) s_b1_target = epoch1;  // the writer warp holds its table writes until this barrier is complete
    KSB_TICK(2);

    // ---- phase 2: Reduce + weighted sum; per-segment (max, tie count) of every pod
    for (int i = tid; i < Gc * 8; i += kThreads)
      reinterpret_cast<int32_t*>(S.norm)[i] = __ldcv(reinterpret_cast<const int32_t*>(cur_tot_norms) + i);
    for (int g = warp; g < Gc; g += kWarps)
      for (int z = lane; z < zs; z += 32) S.zone[g * zs + z] = z < nz ? __ldcv(cur_tot_zsum + g * zs + z) : 0;
    __syncthreads();
    if (tid < Gc) {
      int32_t mz = 0;
      for (int z = 0; z < nz; ++z) mz = max(mz, S.zone[tid * zs + z]);
      S.max_zone[tid] = mz;
    }
    __syncthreads();
    ScoreLut* lut = reinterpret_cast<ScoreLut*>(S.post);  // phase 3's window scratch is idle here
    static_assert(sizeof(ScoreLut) * kMaxG <= sizeof(RowCache) * kMaxG, "score tables alias the window scratch");
    if (active) {
      // the pods this block scores: one run inside a segment, or all of them
      const int lg_lo = seg_lo == seg_hi ? it0 - seg_lo * Gc : 0;
      const int lg_hi = seg_lo == seg_hi ? it1 - 1 - seg_hi * Gc : Gc - 1;
      for (int i = tid; i < (lg_hi - lg_lo + 1) * 32; i += kThreads) {
        const int g = lg_lo + (i >> 5), e = i & 31;
        const KsbNorms nm = S.norm[g];
        ScoreLut& L = lut[g];
        if (e < kLutSc) {
          const double f = spread_node_f(nm.max_sc, e);
          L.fs[e] = f;
          L.fsz[e] = spread_node_part(f);
        } else if (e < kLutSc + kLutZone) {
          const int z = e - kLutSc;
          L.fz[z] = spread_zone_term(S.max_zone[g], z < zs ? S.zone[g * zs + z] : 0);
        } else {
          L.tts[e - kLutSc - kLutZone] = taint_score(nm.max_tt, e - kLutSc - kLutZone);
        }
      }
      static_assert(kLutSc + kLutZone + kLutTt == 32, "one table entry per lane");
    }
    __syncthreads();
    if (active) {
      for (int seg = seg_lo; seg <= seg_hi; ++seg) {
        const int r = seg * kSeg + tid;
        const bool live = r < A.eval_n;
        const uint32_t zone = live ? T.zone[r] : 0u;
        const int g_lo = seg == seg_lo ? it0 - seg_lo * Gc : 0;
        const int g_hi = seg == seg_hi ? it1 - 1 - seg_hi * Gc : Gc - 1;
        size_t at = (half + (size_t)g_lo) * np + (size_t)r;
        for (int g = g_lo; g <= g_hi; ++g, at += np) {
          const KsbPodHeader* h = S.hdr + g;
          const KsbNorms nm = S.norm[g];
          const bool prio = !(h->flags & (PF_SCHED_ERROR | PF_PRIO_ERROR)) && nm.fit_count >= 2;
          if (nm.fit_count == 0 && !(h->flags & PF_SCHED_ERROR)) {
            // no node fits: histogram of failure reasons for FitError.Error() (generic_scheduler.go:70-88),
            // one ballot per reason bit present in the warp; lane b owns bit b
            const uint32_t m = live ? A.reasons[at] : 0u;
            uint32_t un = __reduce_or_sync(0xffffffffu, m);
            int32_t acc = 0;
            while (un) {
              const int bit = __ffs(un) - 1;
              un &= un - 1;
              const int c = __popc(__ballot_sync(0xffffffffu, (m >> bit) & 1u));
              if (lane == bit) acc = c;
            }
            if (acc) {
              atomicAdd(cur_acc_ghist + g * KSB_REASON_BITS + lane, acc);
            }
          }
          if (!prio) continue;  // uniform per (block, g)
          int32_t s = INT32_MIN;
          if (live) {
            const int4 raw = A.raw[at];
            if (raw.x != INT32_MIN) {
              s = final_score(T, nm, lut[g], S.zone + g * zs, S.max_zone[g], raw.x, raw.y, (int32_t)((uint32_t)raw.z >> 24),
                              (uint32_t)raw.z & 0xffffffu, raw.w, zone);
              if constexpr (POL)
                if (T.ext_score) s += T.ext_score[r];  // combinedScores of the extenders (generic_scheduler.go:608-632)
            }
            if constexpr (MULTI) A.gs_score[(half + (size_t)g) * (size_t)A.gs_np + (size_t)(A.gs_row0 + r)] = s;
            else A.score[at] = s;
          }
          int32_t wm = warp_max(s);
          const uint32_t wmask = __ballot_sync(0xffffffffu, s == wm && s != INT32_MIN);
          if (lane == 0) {
            S.wstat[(g * kWarps + warp) * 3] = wm;
            S.wstat[(g * kWarps + warp) * 3 + 1] = __popc(wmask);
            S.wstat[(g * kWarps + warp) * 3 + 2] = (int32_t)wmask;
          }
        }
        __syncthreads();
        if (tid >= g_lo && tid <= g_hi) {
          const int g = tid;
          const KsbPodHeader* h = S.hdr + g;
          const bool prio = !(h->flags & (PF_SCHED_ERROR | PF_PRIO_ERROR)) && S.norm[g].fit_count >= 2;
          if (prio) {
            int32_t m = INT32_MIN, c = 0;
            for (int w = 0; w < kWarps; ++w) {
              int32_t wm = S.wstat[(g * kWarps + w) * 3], wc = S.wstat[(g * kWarps + w) * 3 + 1];
              if (wm > m) {
                m = wm;
                c = wc;
              } else if (wm == m) {
                c += wc;
              }
            }
            if (m == INT32_MIN) c = 0;
            // straight into the arrays phase 3 reads, numbered by global segment: this GPU's own with one rank and on
            // rank 0, rank 0's over NVLink on the other ranks (their blocks fence system-wide before the barrier that
            // ends the iteration anyway, for the per-row results)
            const size_t sat = (half + (size_t)g) * (size_t)A.gs_nseg + (size_t)(A.seg_gbase + seg);
            A.gs_seg_max[sat] = m;
            A.gs_seg_cnt[sat] = c;
            // which rows hold the segment's maximum: word w = rows 32w..32w+31 of the segment
            uint32_t mw[kWarps];
#pragma unroll
            for (int w = 0; w < kWarps; ++w)
              mw[w] = (S.wstat[(g * kWarps + w) * 3] == m && m != INT32_MIN) ? (uint32_t)S.wstat[(g * kWarps + w) * 3 + 2] : 0u;
            uint4* mp = reinterpret_cast<uint4*>(A.gs_seg_mask + sat * kWarps);
            mp[0] = make_uint4(mw[0], mw[1], mw[2], mw[3]);
            mp[1] = make_uint4(mw[4], mw[5], mw[6], mw[7]);
          }
        }
        __syncthreads();
      }
    }
    KSB_TICK(3);
    // FitError histograms of the group: this GPU's partial counts go to rank 0 before phase 3 reads them (plain
    // schedule: at the barrier before phase 3; pipelined: at the barrier that ends the iteration)
    // how far the commit got (block 0 of rank 0 writes it before it arrives): rank 0 stores the two words into every
    // rank's copy on its way through the barrier, so nobody reads them over NVLink
    auto push_done = [&] {
      if constexpr (MULTI) {
        if (A.rank == 0 && tid < 2) {
          const int32_t v = __ldcv(A.group_done + tid);
          for (int q = 1; q < A.n_ranks; ++q) reinterpret_cast<volatile int32_t*>(A.xpeer[q] + A.xo_gdone)[tid] = v;
        }
      }
    };
    auto push_hist = [&] {
      if constexpr (MULTI) {
      if (A.rank == 0) return;
      // FitError histograms of the group: this GPU's partial counts go to rank 0 before phase 3 reads them
      for (int i = tid; i < Gc * KSB_REASON_BITS; i += kThreads) {
        const int32_t v = __ldcg(cur_acc_ghist + i);
        if (v) {
          atomicAdd_system(A.ghist + (size_t)slot * kMaxG * KSB_REASON_BITS + i, v);
          cur_acc_ghist[i] = 0;
        }
      }
      }
    };
    // (pipelined: block 0 decides while the others are in phases 1-2; sharded: ranks other than 0 stored results into rank 0's memory)
    if (!pipe) grid_barrier_x(A, epoch, xepoch, true, xb, push_hist, NoXfer(), MULTI && A.rank != 0);
    KSB_TICK(4);

    // ---- phase 3: block 0 of rank 0 commits the group's pods in order. Rows of other shards are reached
    //      through the peer table (NVLink P2P loads/stores); with one rank every peer access is local.
    //
    // What selectHost + AssumePod need from pod to pod is small: the rows chosen earlier in the group (the row
    // cache), and for every later pod whether such a row still is, or has become, a holder of that pod's maximum.
    // That relation is kept as two bit masks per pod over the cache entries (s_snap: holder at group start,
    // s_now: holder now) plus r0 (snapshot holders above the entry's row), filled by COLUMN PASSES: whenever an
    // entry is created or changed, one warp evaluates every later pod on it, one pod per lane.
    // Pods are then decided in WINDOWS: one warp picks a host for up to 32 consecutive pods at once, one pod per
    // lane, each pick assuming that the window's earlier picks leave its pod's holder set alone; the deciding
    // warps apply each pick's assume step to a scratch copy of its row and run the column pass on it, which
    // tells which later pods of the window were wrong to assume so. Pods before the first such pod are final;
    // the window restarts at that pod, which now knows every row it depends on. A pod whose situation is not the
    // plain one (no fit, one fit, an error, a changed row that stopped fitting or that beats the snapshot's
    // maximum) is decided alone by the exact one-pod path below ("serial"), as in a group of one.
    if (committer && Gc > 0) {
      const int nsegT = A.nseg_total;
      const int sps = A.segs_per_shard;
      const int R = MULTI ? A.n_ranks : 1;
      cm_gid += 1;
      // rank-select over pod g's snapshot holders: the row of the holder with q holders above it
      auto select_snapshot = [&](int g, int q) -> int {
        const int32_t* cum = S.cum + g * nsegT;
        int lo = 0, hi = nsegT - 1;
        while (lo < hi) {  // the last segment whose suffix count exceeds q
          const int mid = (lo + hi + 1) >> 1;
          if (cum[mid] > q) lo = mid;
          else hi = mid - 1;
        }
        int within = q - (lo + 1 < nsegT ? cum[lo + 1] : 0);
        const int pr = 0;
        const KsbPeer& P = PEER(pr);
        const int ls = lo - P.seg_base;
        const uint4* mp = reinterpret_cast<const uint4*>(P.seg_mask + ((half + g) * P.nseg + ls) * kWarps);
        const uint4 m0 = __ldcv(mp), m1 = __ldcv(mp + 1);
        const uint32_t w[kWarps] = {m0.x, m0.y, m0.z, m0.w, m1.x, m1.y, m1.z, m1.w};
        // the word holding the answer: the highest word whose holders, counted from the top, exceed `within`
        int wsel = -1, skip = 0, acc = 0;
        uint32_t word = 0;
#pragma unroll
        for (int wi = kWarps - 1; wi >= 0; --wi) {
          const int c = __popc(w[wi]);
          if (wsel < 0 && within < acc + c) {
            wsel = wi;
            skip = within - acc;
            word = w[wi];
          }
          acc += c;
        }
        const int row = wsel < 0 ? -1 : P.table.row_base + ls * kSeg + wsel * 32 + nth_set_from_top(word, skip);
        return row;
      };
      // Column pass (one warp): every pod g >= g_from of the group, one per lane, on one changed row. Per pod: is the
      // row a holder of the pod's maximum now / was it at group start, does the pod need the serial path because of
      // it, and how many snapshot holders lie above the row. Split in two so that the loads of the pod's group-start
      // results on the row (col_issue) are in flight while the row itself is fetched and committed.
      struct ColIn {
        int32_t base0, score0, above;
        uint32_t reasons0;
        uint4 m0, m1;
        bool act, seg_holds;
      };
      // Groups of up to 16 pods (always so when pipelined) run two columns per pass: lanes 0..15 the pods on one
      // row, lanes 16..31 the same pods on another. `g` is the lane's pod, `valid` whether its column exists.
      // (kept off: with the pipelined schedule two-column passes gave placements that changed from run to run, about
      // one run in forty, and no measurable gain; one column per pass is what scripts/repeat_check.py runs clean)
      constexpr bool dual = false;
      auto col_issue = [&](bool valid, int prank, int crow, int g, int g_from) -> ColIn {
        ColIn in;
        in.act = valid && g >= g_from && g < Gc && ((s_prio >> g) & 1u);
        in.base0 = INT32_MIN;
        in.score0 = 0;
        in.reasons0 = 0;
        in.above = 0;
        in.seg_holds = false;
        in.m0 = make_uint4(0, 0, 0, 0);
        in.m1 = in.m0;
        if (in.act) {
          const KsbPeer& P = PEER(prank);
          // sharded: the group-start results stay on the GPU that evaluated the row (three loads over NVLink per changed
          // row and pod; pushing every rank's results to rank 0 instead made NVLink writes the bottleneck)
          if constexpr (MULTI) {  // every rank stored these into rank 0's arrays at the global row
            const size_t at = (half + g) * (size_t)A.gs_np + (size_t)crow;
            in.base0 = __ldcv(A.gs_base0 + at);
            in.reasons0 = __ldcv(A.gs_reasons + at);
            in.score0 = __ldcv(A.gs_score + at);
          } else {
            const size_t at = (half + g) * (size_t)P.raw_np + (size_t)crow;
            in.base0 = __ldcv(&P.raw[at].x);
            in.reasons0 = __ldcv(P.reasons + at);
            in.score0 = __ldcv(P.score + at);
          }
          const int ls = crow / kSeg;
          const int gseg = P.seg_base + ls;
          const int32_t* cum = S.cum + g * nsegT;
          in.above = gseg + 1 < nsegT ? cum[gseg + 1] : 0;
          in.seg_holds = cum[gseg] > in.above;
          if (in.seg_holds) {  // the segment holds snapshot holders: its mask tells how many lie above the row
            const uint4* mp = reinterpret_cast<const uint4*>(P.seg_mask + ((half + g) * P.nseg + ls) * kWarps);
            in.m0 = __ldcv(mp);
            in.m1 = __ldcv(mp + 1);
          }
        }
        return in;
      };
      // per lane: its pod on its column's row; the callers ballot the three outcomes
      auto col_eval = [&](RowCache* c, const ColIn& in, int g, int32_t* r0_out, bool* bit_now, bool* bit_snap, bool* spec) {
        *bit_now = false;
        *bit_snap = false;
        *spec = false;
        if (in.act) {
          const KsbPodHeader* h = S.hdr + g;
          const PodView pv = make_view(h, S.pool + S.pool_off[g], T.W);
          const KsbPeer& P = PEER(c->pad & 0xff);
          int32_t r0v = in.above;
          if (in.seg_holds) {
            const uint32_t w[kWarps] = {in.m0.x, in.m0.y, in.m0.z, in.m0.w, in.m1.x, in.m1.y, in.m1.z, in.m1.w};
            const int rel = c->row % kSeg;
            const int wi0 = rel >> 5;
#pragma unroll
            for (int wi = 0; wi < kWarps; ++wi) {
              if (wi > wi0) r0v += __popc(w[wi]);
              else if (wi == wi0) r0v += __popc(w[wi] & ~((2u << (rel & 31)) - 1u));
            }
          }
          r0_out[g] = r0v;
          uint32_t new_reasons;
          int32_t base;
          eval_cached<POL>(P.table, *c, pv, in.reasons0 & ~kDynMask, &new_reasons, &base, nullptr, true, false);
          const bool was_fit = in.base0 != INT32_MIN;
          const int32_t gm = s_gmax0[g];
          const int32_t cold = was_fit ? in.score0 : INT32_MIN;
          const int32_t cnew =
              (was_fit && new_reasons == 0) ? (int32_t)((uint32_t)cold - (uint32_t)in.base0 + (uint32_t)base) : INT32_MIN;
          *bit_snap = cold == gm;
          *bit_now = cnew == gm;
          *spec = (was_fit && new_reasons != 0) || cnew > gm;
        }
      };
      // one column, one pod per lane
      auto column_pass = [&](RowCache& c, int g_from, uint32_t* now_m, uint32_t* snap_m, uint32_t* spec_m, int32_t* r0_out) {
        const ColIn in = col_issue(true, c.pad & 0xff, c.row, lane, g_from);
        bool bn, bs, sp;
        col_eval(&c, in, lane, r0_out, &bn, &bs, &sp);
        *now_m = __ballot_sync(0xffffffffu, bn);
        *snap_m = __ballot_sync(0xffffffffu, bs);
        *spec_m = __ballot_sync(0xffffffffu, sp);
      };
      // make a column pass's outcome part of the later pods' state, as cache entry `t`
      auto apply_column = [&](int t, int g_from, bool is_new, uint32_t now_m, uint32_t snap_m, uint32_t spec_m, const int32_t* r0_src) {
        const int g = lane;
        if (g >= g_from && g < Gc && ((s_prio >> g) & 1u)) {
          const uint32_t bit = 1u << t;
          if ((now_m >> g) & 1u) atomicOr(&s_now[g], bit);
          else atomicAnd(&s_now[g], ~bit);
          if (is_new) {
            if ((snap_m >> g) & 1u) atomicOr(&s_snap[g], bit);
            else atomicAnd(&s_snap[g], ~bit);
            S.r0[g * kCache + t] = r0_src[g];
          }
        }
        if (lane == 0 && spec_m) atomicOr(const_cast<uint32_t*>(&s_special), spec_m);
      };

      {
        // Rows in the cache. Pipelined: this group was evaluated before the previous group's rows reached the
        // table, so those rows (tagged with that group's number) stay in the cache and count as changed from the
        // first pod on; older entries are dropped. All warps: read the kept entries that move, then write them.
        const int have = s_ncr;
        const int prev = cm_gid - 1;
        const bool keep = pipe && cm_carry && lane < have && (S.rows[lane].pad >> 8) == prev;
        const uint32_t km = __ballot_sync(0xffffffffu, keep);
        constexpr int kWordsPerRow = (int)(sizeof(RowCache) / 4);
        static_assert(kWordsPerRow <= 96 && kCache <= 4 * kWarps, "an entry is moved by one warp in three words per lane");
        uint32_t mv[4][3];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const int i = warp + j * kWarps;
          if ((km >> i) & 1u) {
            const uint32_t* src = reinterpret_cast<const uint32_t*>(S.rows + i);
#pragma unroll
            for (int x = 0; x < 3; ++x) mv[j][x] = lane + 32 * x < kWordsPerRow ? src[lane + 32 * x] : 0u;
          }
        }
        __syncthreads();
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const int i = warp + j * kWarps;
          if ((km >> i) & 1u) {
            const int d = __popc(km & ((1u << i) - 1u));
            uint32_t* dst = reinterpret_cast<uint32_t*>(S.rows + d);
#pragma unroll
            for (int x = 0; x < 3; ++x)
              if (lane + 32 * x < kWordsPerRow) dst[lane + 32 * x] = mv[j][x];
          }
        }
        if (tid == 0) s_ncr = __popc(km);
        __syncthreads();
        if (warp == 0 && lane < __popc(km)) s_erow[lane] = PEER(S.rows[lane].pad & 0xff).table.row_base + S.rows[lane].row;
      }
      // Per pod ranked by score, one pod per warp: the maximum over all segments, and per segment the number of
      // rows holding it in this and all higher segments (suffix counts; entry 0 is the pod's tie count k0).
      for (int g = warp; g < Gc; g += kWarps) {
        const KsbPodHeader* h = S.hdr + g;
        const bool prio = !(h->flags & (PF_SCHED_ERROR | PF_PRIO_ERROR)) && S.norm[g].fit_count >= 2;
        if (!prio) {
          if (lane == 0) {
            s_gmax0[g] = INT32_MIN;
            s_k0[g] = 0;
          }
          continue;
        }
        int32_t* cum = S.cum + g * nsegT;
        int32_t gmax0, carry = 0;
        if (nsegT <= 8 * 32 && A.self.nseg == nsegT && (A.self.nseg & 3) == 0) {
          // up to 256 segments (65,536 rows; sharded: rank 0's arrays hold every rank's segments): lane l takes segments 8l..8l+7 with four 16-byte loads, sums
          // its own eight counts serially and one warp scan joins the lanes (the row stride of the per-pod statistics
          // is a multiple of four segments, so the vectors are aligned)
          const int32_t* pmax = A.self.seg_max + (half + g) * A.self.nseg;
          const int32_t* pcnt = A.self.seg_cnt + (half + g) * A.self.nseg;
          int32_t vmax[8], vcnt[8];
#pragma unroll
          for (int hh = 0; hh < 2; ++hh) {
            const int i0 = lane * 8 + hh * 4;
            int4 a4 = make_int4(INT32_MIN, INT32_MIN, INT32_MIN, INT32_MIN), c4 = make_int4(0, 0, 0, 0);
            if (i0 < nsegT) {  // nsegT may end inside the vector: the statistics arrays are padded to the stride
              a4 = __ldcv(reinterpret_cast<const int4*>(pmax + i0));
              c4 = __ldcv(reinterpret_cast<const int4*>(pcnt + i0));
            }
            vmax[hh * 4 + 0] = a4.x; vmax[hh * 4 + 1] = a4.y; vmax[hh * 4 + 2] = a4.z; vmax[hh * 4 + 3] = a4.w;
            vcnt[hh * 4 + 0] = c4.x; vcnt[hh * 4 + 1] = c4.y; vcnt[hh * 4 + 2] = c4.z; vcnt[hh * 4 + 3] = c4.w;
          }
          int32_t m = INT32_MIN;
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            if (lane * 8 + j >= nsegT) vmax[j] = INT32_MIN;
            m = max(m, vmax[j]);
          }
          gmax0 = warp_max(m);
          int32_t suf[8];  // holders in this lane's segments j..7
          int32_t run = 0;
#pragma unroll
          for (int j = 7; j >= 0; --j) {
            run += vmax[j] == gmax0 ? vcnt[j] : 0;
            suf[j] = run;
          }
          int32_t incl = run;  // + the lanes above
#pragma unroll
          for (int d = 1; d < 32; d <<= 1) {
            const int32_t t = __shfl_down_sync(0xffffffffu, incl, d);
            if (lane + d < 32) incl += t;
          }
          const int32_t above = incl - run;
#pragma unroll
          for (int j = 0; j < 8; ++j)
            if (lane * 8 + j < nsegT) cum[lane * 8 + j] = above + suf[j];
          carry = __shfl_sync(0xffffffffu, incl, 0);
        } else {
          int32_t m = INT32_MIN;
          for (int i = lane; i < nsegT; i += 32) {
            const int q = 0;
            const KsbPeer& P = PEER(q);
            const int32_t v = __ldcv(P.seg_max + (half + g) * P.nseg + (i - P.seg_base));
            cum[i] = v;
            m = max(m, v);
          }
          gmax0 = warp_max(m);
          __syncwarp();
          for (int top = nsegT; top > 0; top -= 32) {
            const int i = top - 1 - lane;  // lane 0 = highest segment
            int32_t cnt = 0;
            if (i >= 0 && cum[i] == gmax0) {
              const int q = 0;
              const KsbPeer& P = PEER(q);
              cnt = __ldcv(P.seg_cnt + (half + g) * P.nseg + (i - P.seg_base));
            }
            int32_t incl = cnt;
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) {
              const int32_t t = __shfl_up_sync(0xffffffffu, incl, d);
              if (lane >= d) incl += t;
            }
            __syncwarp();
            if (i >= 0) cum[i] = carry + incl;
            carry += __shfl_sync(0xffffffffu, incl, 31);
          }
        }
        if (lane == 0) {
          s_gmax0[g] = gmax0;
          s_k0[g] = carry;
        }
      }
      if (tid < kMaxG) {
        s_ready[tid] = 0;
        s_snap[tid] = 0;
        s_now[tid] = 0;
      }
      if (warp == 0) {
        // which pods take the plain path, and how far lastNodeIndex has moved when each pod's turn comes: only pods
        // that reach selectHost bump it (generic_scheduler.go:144-146, 186-189)
        bool plain = false;
        if (lane < Gc) {
          const KsbPodHeader* h = S.hdr + lane;
          plain = !(h->flags & (PF_SCHED_ERROR | PF_PRIO_ERROR)) && S.norm[lane].fit_count >= 2;
        }
        const uint32_t pm = __ballot_sync(0xffffffffu, plain);
        s_lrel[lane] = __popc(pm & ((1u << lane) - 1u));
        if (lane == 0) {
          s_prio = pm;
          s_special = A.commit ? ~pm : 0xffffffffu;
          s_stop = -1;
          s_cut = 0;
        }
      }
      __syncthreads();
      {  // carried rows: their column pass against this group's pods
        const int ncr = s_ncr;
        if (dual) {
          for (int i = warp * 2; i < ncr; i += kWarps * 2) {
            const int sub = lane >> 4, g = lane & 15;
            const bool valid = i + sub < ncr;
            RowCache* c = S.rows + (valid ? i + sub : i);
            const ColIn in = col_issue(valid, c->pad & 0xff, c->row, g, 0);
            bool bn, bs, sp;
            col_eval(c, in, g, S.t_r0 + (i + sub) * kMaxG, &bn, &bs, &sp);
            const uint32_t nm_ = __ballot_sync(0xffffffffu, bn), sm_ = __ballot_sync(0xffffffffu, bs), xm_ = __ballot_sync(0xffffffffu, sp);
            __syncwarp();
            apply_column(i, 0, true, nm_ & 0xffffu, sm_ & 0xffffu, xm_ & 0xffffu, S.t_r0 + i * kMaxG);
            if (i + 1 < ncr) apply_column(i + 1, 0, true, nm_ >> 16, sm_ >> 16, xm_ >> 16, S.t_r0 + (i + 1) * kMaxG);
          }
        } else {
          for (int i = warp; i < ncr; i += kWarps) {
            uint32_t nm_, sm_, xm_;
            column_pass(S.rows[i], 0, &nm_, &sm_, &xm_, S.t_r0 + i * kMaxG);
            __syncwarp();
            apply_column(i, 0, true, nm_, sm_, xm_, S.t_r0 + i * kMaxG);
          }
        }
      }
      __syncthreads();
      long long t3_mark = 0;
      if (tid == 0) {  // phase-3 breakdown (ksb_stats.stage_cycles[7..11]): rounds, window pods, one-pod-path pods, set-up and deciding cycles
        t3_mark = clock64();
        s_tk[14] += t3_mark - s_tk[8];
      }
#if KSB_FINE_TIMING == 1
      long long f3_mark = clock64();
#define KSB_F3(i)                         \
  do {                                    \
    if (tid == 0) {                       \
      const long long now_ = clock64();   \
      s_tk[16 + (i)] += now_ - f3_mark;   \
      f3_mark = now_;                     \
    }                                     \
  } while (0)
#define KSB_G(i)                          \
  do {                                    \
    if (tid == 0) {                       \
      const long long now_ = clock64();   \
      s_tk[24 + (i)] += now_ - g3_mark;   \
      g3_mark = now_;                     \
    }                                     \
  } while (0)
#define KSB_G0()                          \
  do {                                    \
    if (tid == 0) g3_mark = clock64();    \
  } while (0)
      long long g3_mark = 0;
#else
#define KSB_F3(i) ((void)0)
#define KSB_G(i) ((void)0)
#define KSB_G0() ((void)0)
#endif
#define KSB_BAR3() asm volatile("bar.sync 1, %0;" ::"n"(kDeciders * 32) : "memory")
      if (warp < kDeciders) {
        int f = 0;          // pods decided so far
        const uint64_t L0 = last_index;
        while (f < Gc) {
          const uint32_t special = s_special;
          const int ncr = s_ncr;
          KSB_F3(0);
          KSB_BAR3();  // everyone holds this round's view before anyone changes it
          KSB_F3(1);
          if ((special >> f) & 1u) {
            // ================= one pod, exact (warp 0) =================
            if (warp == 0) {
              const int g = f;
              const KsbPodHeader* h = S.hdr + g;
              const PodView pv = make_view(h, S.pool + S.pool_off[g], T.W);
              const KsbNorms nm = S.norm[g];
              const bool sched_err = (h->flags & PF_SCHED_ERROR) != 0;
              const bool prio = !sched_err && !(h->flags & PF_PRIO_ERROR) && nm.fit_count >= 2;
              const uint64_t last_g = L0 + (uint64_t)s_lrel[g];
              KsbResult res;
              res.status = 0;
              res.row = -1;
              res.num_filtered = nm.fit_count;  // lowered below when rows left the filtered list inside the group
              res.fiterr_slot = -1;
              res.score = 0;
#pragma unroll
              for (int i = 0; i < KSB_MAX_EXT; ++i) res.device_mask[i] = 0;
              int chosen = -1, chosen_rank = 0;  // local row on shard chosen_rank
              bool cut = false;
              if (sched_err) {
                res.status = 3;
              } else {
                // (1) this pod on the rows changed since the snapshot, from the row cache
                uint32_t old_reasons = 0, new_reasons = 0;
                int32_t cnew = INT32_MIN, cold = INT32_MIN;
                int crow = -1, crank = 0;
                bool flip = false;
                if (lane < ncr) {
                  RowCache& c = S.rows[lane];
                  crow = c.row;
                  crank = c.pad & 0xff;
                  const KsbPeer& P = PEER(crank);
                  int32_t base0;  // INT32_MIN = the row did not fit at group start
                  if constexpr (MULTI) {  // stored into rank 0's arrays, at the global row, by the rank that evaluated it
                    const size_t at = (half + g) * (size_t)A.gs_np + (size_t)crow;
                    base0 = __ldcv(A.gs_base0 + at);
                    old_reasons = __ldcv(A.gs_reasons + at);
                    if (prio) cold = __ldcv(A.gs_score + at);
                  } else {
                    const size_t at = (half + g) * (size_t)P.raw_np + (size_t)crow;
                    base0 = __ldcv(&P.raw[at].x);
                    old_reasons = __ldcv(P.reasons + at);
                    if (prio) cold = __ldcv(P.score + at);
                  }
                  int32_t base;
                  eval_cached<POL>(P.table, c, pv, old_reasons & ~kDynMask, &new_reasons, &base, nullptr, prio);
                  const bool was_fit = base0 != INT32_MIN;
                  flip = was_fit && new_reasons != 0;  // a fit row turned unfit: the Reduce inputs may have moved
                  if (prio && was_fit && new_reasons == 0)
                    // final_score = base + terms that depend only on group-start values: the new total is the old one
                    // with the base score exchanged
                    cnew = (int32_t)((uint32_t)cold - (uint32_t)base0 + (uint32_t)base);
                  if (!prio || !was_fit) cold = INT32_MIN;
                }
                int fit_now = nm.fit_count;
                {
                  // A row that fit at group start and no longer does leaves the filtered list. The Reduce inputs of the
                  // snapshot stay valid if the row cannot have been a holder of any of them: then it simply stops being
                  // a candidate. Otherwise speculation is dropped and this pod opens the next group.
                  const uint32_t fb = __ballot_sync(0xffffffffu, flip);
                  if (fb) {
                    bool safe = true;
                    if (flip) {
                      const KsbPeer& P = PEER(crank);
                      // the Map values of the row (sharded: kept by the rank that evaluated it, a remote read on this rare path)
                      const int own = MULTI ? min(R - 1, (crow / kSeg) / sps) : 0;
                      const KsbPeer& O = PEER(own);
                      const int4 raw = __ldcv(O.raw + (half + g) * (MULTI ? (size_t)O.raw_np : (size_t)P.table.n_pad) +
                                              (crow - (MULTI ? O.raw_row0 : 0)));
                      const int32_t na = raw.y, tt = (int32_t)((uint32_t)raw.z >> 24), ipa = raw.w;
                      const uint32_t sc = (uint32_t)raw.z & 0xffffffu;
                      safe = (nm.max_na == 0 || na < nm.max_na) && (nm.max_tt == 0 || tt < nm.max_tt) && sc == 0 &&
                             (nm.ipa_max == nm.ipa_min || (ipa > nm.ipa_min && ipa < nm.ipa_max));
                    }
                    fit_now -= __popc(fb);
                    if (!__all_sync(0xffffffffu, safe) || fit_now < 2) cut = true;
                  }
                }
                res.num_filtered = fit_now;
                if (cut) {
                  // nothing
                } else if (nm.fit_count >= 2 && (h->flags & PF_PRIO_ERROR)) {
                  res.status = 3;
                } else if (nm.fit_count == 0) {
                  // FitError (generic_scheduler.go:133-139): the histogram over all rows was built in phase 2; here
                  // only the correction for rows whose reasons changed since the group started
                  res.status = 2;
                  int slotv = 0;
                  if (lane == 0) slotv = atomicAdd(A.fiterr_count, 1);
                  slotv = __shfl_sync(0xffffffffu, slotv, 0);
                  if (slotv < A.fiterr_cap) {
                    S.fe_hist[lane] = 0;
                    __syncwarp();
                    if (lane < ncr) {
                      uint32_t m = old_reasons;
                      while (m) {
                        int bit = __ffs(m) - 1;
                        m &= m - 1;
                        atomicAdd(&S.fe_hist[bit], -1);
                      }
                      m = new_reasons;
                      while (m) {
                        int bit = __ffs(m) - 1;
                        m &= m - 1;
                        atomicAdd(&S.fe_hist[bit], 1);
                      }
                    }
                    __syncwarp();
                    A.fiterr_hist[(size_t)slotv * KSB_REASON_BITS + lane] =
                        __ldcg(A.ghist + ((size_t)slot * kMaxG + g) * KSB_REASON_BITS + lane) + S.fe_hist[lane];
                    res.fiterr_slot = slotv;
                  }
                  __syncwarp();
                } else if (nm.fit_count == 1) {
                  // generic_scheduler.go:144-146: no selectHost, counter untouched
                  chosen_rank = 0;
                  chosen = nm.one_row - PEER(chosen_rank).table.row_base;
                } else {
                  // (2) selectHost (generic_scheduler.go:177-192): lastNodeIndex % k over the rows tied at the maximum,
                  //     in descending row order: the snapshot's holders patched with the changed rows
                  const int32_t gmax0 = s_gmax0[g];
                  const int32_t gmax = max(gmax0, warp_max(cnew));
                  int pick = -1, pick_rank = 0;
                  if (gmax > gmax0) {
                    // a changed row now beats every row of the snapshot: the ties are changed rows only
                    const bool cand = lane < ncr && cnew == gmax;
                    const uint32_t cb = __ballot_sync(0xffffffffu, cand);
                    const int k = __popc(cb);
                    const int rem = (int)(last_g % (uint64_t)k);
                    const int grow = cand ? PEER(crank).table.row_base + crow : -1;
                    int above = 0;
                    for (uint32_t mm = cb; mm; mm &= mm - 1) {
                      const int gr = __shfl_sync(0xffffffffu, grow, __ffs(mm) - 1);
                      above += (gr > grow) ? 1 : 0;
                    }
                    const uint32_t hit = __ballot_sync(0xffffffffu, cand && above == rem);
                    const int src = __ffs(hit) - 1;
                    pick = __shfl_sync(0xffffffffu, crow, src);
                    pick_rank = __shfl_sync(0xffffffffu, crank, src);
                  } else {
                    const uint32_t Am = __ballot_sync(0xffffffffu, lane < ncr && cnew == gmax && cold != gmax);
                    const uint32_t Dm = __ballot_sync(0xffffffffu, lane < ncr && cold == gmax && cnew != gmax);
                    const int32_t k = s_k0[g] + __popc(Am) - __popc(Dm);
                    if (k <= 0) {
                      cut = true;  // every holder of the group-start maximum changed and fell below it
                    } else {
                      const int rem = (last_g >> 32) == 0 ? (int)((uint32_t)last_g % (uint32_t)k) : (int)(last_g % (uint64_t)k);
                      int q = 0;
                      const int ent = select_merged(Am, Dm, rem, S.r0 + g * kCache, s_erow, &q);
                      int grow = ent >= 0 ? s_erow[ent] : select_snapshot(g, q);
                      if (grow >= 0) {
                        pick_rank = 0;
                        pick = grow - PEER(pick_rank).table.row_base;
                      }
                    }
                  }
                  if (!cut) {
                    if (pick >= 0) {
                      chosen = pick;
                      chosen_rank = pick_rank;
                      res.score = gmax;
                    } else {
                      res.status = 3;  // only reachable after a barrier timeout
                    }
                  }
                }
              }
              if (!cut) {
                // (3) the assume step on the chosen row: the row cache is updated and the write-through (table columns,
                //     class counters, result record) handed to the writer warp
                CommitRec& rec = S.recs[g];
                int idx = -1;
                bool fresh = false;
                if (chosen >= 0 && A.commit) {
                  const bool mine = lane < ncr && S.rows[lane].row == chosen && (S.rows[lane].pad & 0xff) == chosen_rank;
                  const uint32_t bal = __ballot_sync(0xffffffffu, mine);
                  if (bal) idx = __ffs(bal) - 1;
                }
                if (chosen >= 0) {
                  const KsbPeer& P = PEER(chosen_rank);
                  uint32_t picks[KSB_MAX_EXT] = {0, 0, 0, 0};
                  if (A.commit) {
                    if (idx < 0) {
                      if (ncr >= kCache) {
                        cut = true;  // the cache is full: this pod opens the next group (not reachable: groups are sized for it)
                      } else {
                        idx = ncr;
                        fresh = true;
                        cache_fill(P.table, chosen, S.rows + idx, lane);
                      }
                    }
                    if (!cut) {
                      if (lane == 31) S.rows[idx].pad = chosen_rank | (cm_gid << 8);  // shard, and the group that last changed it
                      __syncwarp();
                      if (h->n_ext) {
                        // the chosen row fits, so the device selection succeeds: only its picks are needed. Class tests the
                        // entry does not hold yet are run by the whole warp, one device per lane.
                        RowCache* ce = S.rows + idx;
                        for (int i = 0; i < h->n_ext; ++i) {
                          const int rs = h->ext[i].rslot;
                          const uint16_t id = h->ext[i].cls_id;
                          if (rs < KSB_MAX_RES && h->ext[i].num > 0 && id != 0 && ce->mid[i] != id) {
                            const uint32_t cls = P.table.dev_class[((size_t)rs * P.table.n_pad + chosen) * KSB_MAX_DEV + lane];
                            const uint64_t* cm = pv.extcls + i * KSB_CLASS_WORDS;
                            const uint32_t match = __ballot_sync(0xffffffffu, ((cm[cls >> 6] >> (cls & 63)) & 1ull) != 0);
                            if (lane == 0) {
                              ce->mm[i] = match;
                              ce->mid[i] = id;
                            }
                          }
                        }
                        __syncwarp();
                        if (lane == 0) select_devices_cached(P.table, ce, pv, picks);
#pragma unroll
                        for (int i = 0; i < KSB_MAX_EXT; ++i) picks[i] = __shfl_sync(0xffffffffu, picks[i], 0);
                      }
                      commit_and_record(S.rows + idx, &rec, pv, picks, chosen, chosen_rank, lane);
                    }
                  } else {
                    if (h->n_ext) {
                      if (lane == 0) {
                        EvalOut o;
                        RowCommon rc;
                        load_row_common<true>(P.table, chosen, rc);
                        eval_node<true, POL>(P.table, chosen, rc, pv, o, picks);
                      }
#pragma unroll
                      for (int i = 0; i < KSB_MAX_EXT; ++i) picks[i] = __shfl_sync(0xffffffffu, picks[i], 0);
                    }
                    if (lane == 31) rec.row = -1;
                  }
                  res.row = P.table.row_base + chosen;
#pragma unroll
                  for (int i = 0; i < KSB_MAX_EXT; ++i) res.device_mask[i] = picks[i];
                } else if (lane == 31) {
                  rec.row = -1;
                }
                if (!cut) {
                  if (lane == 0) rec.res = res;
                  __syncwarp();
                  __threadfence_block();
                  if (lane == 0) s_ready[g] = 1;
                  if (idx >= 0) {
                    // the changed row against every later pod of the group
                    if (lane == 0) {
                      s_erow[idx] = res.row;
                      if (fresh) s_ncr = ncr + 1;
                    }
                    uint32_t nm_, sm_, xm_;
                    column_pass(S.rows[idx], g + 1, &nm_, &sm_, &xm_, S.t_r0);
                    __syncwarp();
                    apply_column(idx, g + 1, fresh, nm_, sm_, xm_, S.t_r0);
                  }
                }
              }
              if (cut && lane == 0) s_cut = 1;
            }
            KSB_F3(7);
            KSB_BAR3();
            if (s_cut) break;
            if (tid == 0) s_tk[13] += 1;
            f += 1;
            continue;
          }
          // ================= a window of plain pods =================
          int W = __ffs((int)(special >> f)) - 1;  // plain pods in a row from f on
          if (W < 0) W = 32;
          W = min(min(W, wcap), min(Gc - f, kCache - ncr));
          if (W <= 0) {  // no room left in the row cache: the rest of the group opens the next one
            if (tid == 0) s_cut = 1;
            KSB_BAR3();
            break;
          }
          // (1) picks, one pod per lane
          if (warp == 0) {
            int pick = -1;
            bool bad = false;
            KSB_G0();
            if (lane < W) {
              const int g = f + lane;
              const uint32_t sn = s_snap[g], nw = s_now[g];
              const uint32_t Am = nw & ~sn, Dm = sn & ~nw;
              const int32_t k = s_k0[g] + __popc(Am) - __popc(Dm);
              if (k <= 0) {
                bad = true;
              } else {
                const uint64_t last_g = L0 + (uint64_t)s_lrel[g];
                const int rem = (last_g >> 32) == 0 ? (int)((uint32_t)last_g % (uint32_t)k) : (int)(last_g % (uint64_t)k);
                int q = 0;
                __syncwarp(__activemask());
                KSB_G(0);
                const int ent = select_merged(Am, Dm, rem, S.r0 + g * kCache, s_erow, &q);
                __syncwarp(__activemask());
                KSB_G(1);
                pick = ent >= 0 ? s_erow[ent] : select_snapshot(g, q);
                if (pick < 0) bad = true;
              }
              s_pick[lane] = pick;
            }
            __syncwarp();
            KSB_G(2);
            const uint32_t bm = __ballot_sync(0xffffffffu, bad);
            if (lane == 0) s_bad0 = bm;  // these pods go to the serial path (which ends the group if it must)
          }
          KSB_F3(2);
          KSB_BAR3();
          KSB_F3(1);
          // (2) per pick: its row after the assume step, and that row against the later pods (two picks per pass
          //     when the group has at most 16 pods)
          const int cpw = dual ? 2 : 1;
          for (int w0 = warp * cpw; w0 < W; w0 += kDeciders * cpw) {
            // which column this lane evaluates, and its pod
            const int sub = dual ? (lane >> 4) : 0;
            const int gl = dual ? (lane & 15) : lane;
            int my_e = -1, my_idx = -2, my_rank = 0, my_lrow = 0;  // this lane's column
            int e_[2] = {-1, -1}, idx_[2] = {-2, -2}, rank_[2] = {0, 0}, lrow_[2] = {0, 0};
#pragma unroll
            for (int s2 = 0; s2 < 2; ++s2) {
              const int w = w0 + s2;
              if (s2 < cpw && w < W) {
                const int e = s_pick[w];
                if (e >= 0) {
                  const int erank = 0;
                  const uint32_t hit = __ballot_sync(0xffffffffu, lane < ncr && s_erow[lane] == e);
                  e_[s2] = e;
                  idx_[s2] = hit ? __ffs(hit) - 1 : -1;
                  rank_[s2] = erank;
                  lrow_[s2] = e - PEER(erank).table.row_base;
                }
              }
              if (sub == s2) {
                my_e = e_[s2];
                my_idx = idx_[s2];
                my_rank = rank_[s2];
                my_lrow = lrow_[s2];
              }
            }
            const int my_w = w0 + sub;
            KSB_G0();
            const ColIn cin = col_issue(my_e >= 0, my_rank, my_lrow, gl, f + my_w + 1);
            KSB_G(3);
            int64_t fv[2] = {0, 0};
            uint32_t fv2[2] = {0, 0};
#pragma unroll
            for (int s2 = 0; s2 < 2; ++s2)  // rows not in the cache yet: both fetches in flight together
              if (s2 < cpw && w0 + s2 < W && e_[s2] >= 0 && idx_[s2] < 0)
                cache_fill_issue(PEER(rank_[s2]).table, lrow_[s2], lane, &fv[s2], &fv2[s2]);
#pragma unroll
            for (int s2 = 0; s2 < 2; ++s2) {
              const int w = w0 + s2;
              if (!(s2 < cpw && w < W)) continue;
              const int e = e_[s2];
              if (e < 0) continue;
              const int g = f + w;
              const int idx = idx_[s2], erank = rank_[s2], lrow = lrow_[s2];
              const KsbPodHeader* h = S.hdr + g;
              const PodView pv = make_view(h, S.pool + S.pool_off[g], T.W);
              const KsbPeer& P = PEER(erank);
              RowCache* pc = S.post + w;
              if (idx >= 0) {
                const uint32_t* src = reinterpret_cast<const uint32_t*>(S.rows + idx);
                uint32_t* dst = reinterpret_cast<uint32_t*>(pc);
                for (int x = lane; x < (int)(sizeof(RowCache) / 4); x += 32) dst[x] = src[x];
              } else {
                cache_fill_store(pc, lrow, lane, fv[s2], fv2[s2]);
              }
              __syncwarp();
              if (lane == 31) pc->pad = erank | (cm_gid << 8);
              __syncwarp();
              uint32_t picks[KSB_MAX_EXT] = {0, 0, 0, 0};
              if (h->n_ext) {
                for (int i = 0; i < h->n_ext; ++i) {
                  const int rs = h->ext[i].rslot;
                  const uint16_t id = h->ext[i].cls_id;
                  if (rs < KSB_MAX_RES && h->ext[i].num > 0 && id != 0 && pc->mid[i] != id) {
                    const uint32_t cls = P.table.dev_class[((size_t)rs * P.table.n_pad + lrow) * KSB_MAX_DEV + lane];
                    const uint64_t* cm = pv.extcls + i * KSB_CLASS_WORDS;
                    const uint32_t match = __ballot_sync(0xffffffffu, ((cm[cls >> 6] >> (cls & 63)) & 1ull) != 0);
                    if (lane == 0) {
                      pc->mm[i] = match;
                      pc->mid[i] = id;
                    }
                  }
                }
                __syncwarp();
                if (lane == 0) select_devices_cached(P.table, pc, pv, picks);
#pragma unroll
                for (int i = 0; i < KSB_MAX_EXT; ++i) picks[i] = __shfl_sync(0xffffffffu, picks[i], 0);
              }
              CommitRec& rec = S.recs[g];
              commit_and_record(pc, &rec, pv, picks, lrow, erank, lane);
              if (lane == 0) {
                KsbResult res;
                res.status = 0;
                res.row = e;
                res.num_filtered = S.norm[g].fit_count;
                res.fiterr_slot = -1;
                res.score = s_gmax0[g];
#pragma unroll
                for (int i = 0; i < KSB_MAX_EXT; ++i) res.device_mask[i] = picks[i];
                rec.res = res;
              }
            }
            __syncwarp();
            KSB_G(4);
            bool bn, bs, sp;
            col_eval(S.post + my_w, cin, gl, S.t_r0 + my_w * kMaxG, &bn, &bs, &sp);
            __syncwarp();
            KSB_G(5);
            // later pods of the window whose pick assumed this row's standing unchanged, and were wrong (or chose the
            // same row, whose copy here lacks this pod)
            bool dirty = false;
            if (my_e >= 0 && gl > f + my_w && gl < f + W) {
              const bool used = my_idx >= 0 ? ((s_now[gl] >> my_idx) & 1u) != 0 : bs;
              dirty = sp || bn != used || s_pick[gl - f] == my_e;
#if KSB_FINE_TIMING == 1
              // why windows end early: per (pick, later pod) cause masks, read for the first dirty pod in step (3)
              if (sp) atomicOr(&t_cause[0][my_w], 1u << gl);
              if (bn != used) atomicOr(&t_cause[bn ? 1 : 2][my_w], 1u << gl);
              if (s_pick[gl - f] == my_e) atomicOr(&t_cause[3][my_w], 1u << gl);
#endif
            }
            const uint32_t nm_ = __ballot_sync(0xffffffffu, bn), sm_ = __ballot_sync(0xffffffffu, bs);
            const uint32_t xm_ = __ballot_sync(0xffffffffu, sp), dm = __ballot_sync(0xffffffffu, dirty);
#if KSB_FINE_TIMING == 1 && defined(KSB_CAUSES)
            {  // (pick, later pod of the window) pairs: all, the row held the pod's maximum at group start, and it no longer does
              const uint32_t inwin = __ballot_sync(0xffffffffu, my_e >= 0 && gl > f + my_w && gl < f + W);
              if (lane == 0) {
                atomicAdd(reinterpret_cast<unsigned long long*>(&s_tk[38]), (unsigned long long)__popc(inwin));
                atomicAdd(reinterpret_cast<unsigned long long*>(&s_tk[39]), (unsigned long long)__popc(inwin & sm_));
                atomicAdd(reinterpret_cast<unsigned long long*>(&s_tk[36]), (unsigned long long)__popc(inwin & sm_ & ~nm_));
              }
            }
#endif
            if (lane < cpw && w0 + lane < W) {
              const int w = w0 + lane;
              const int sh = dual ? 16 * lane : 0;
              const uint32_t keep = dual ? 0xffffu : 0xffffffffu;
              const int idx = lane == 0 ? idx_[0] : idx_[1];
              t_dirty[w] = idx == -2 ? 0u : (dm >> sh) & keep;
              t_now[w] = (nm_ >> sh) & keep;
              t_snap[w] = (sm_ >> sh) & keep;
              t_spec[w] = (xm_ >> sh) & keep;
              t_idx[w] = idx;
            }
            // (slot 6 of the granular marks holds the wait for the writer warp)
          }
          KSB_F3(3);
          KSB_BAR3();
          KSB_F3(4);
          // (3) how many of the window's pods are final
          int n_ok;
          uint32_t newmask;
          {
            const uint32_t d = lane < W ? t_dirty[lane] : 0u;
            uint32_t dall = __reduce_or_sync(0xffffffffu, d) | (s_bad0 << f);
            dall >>= f;
            n_ok = dall ? min(W, __ffs((int)dall) - 1) : W;
            newmask = __ballot_sync(0xffffffffu, lane < n_ok && t_idx[lane] == -1);
#if KSB_FINE_TIMING == 1
            if (warp == 0) {
              const int dpod = f + n_ok;  // the pod that ended the window (if any)
              uint32_t c[4];
#pragma unroll
              for (int k = 0; k < 4; ++k) {
                const uint32_t m = lane < W ? t_cause[k][lane] : 0u;
                c[k] = __reduce_or_sync(0xffffffffu, m);
                if (lane < W) t_cause[k][lane] = 0;
              }
              if (lane == 0) {
                if (n_ok < W) {
                  for (int k = 0; k < 4; ++k)
                    if ((c[k] >> dpod) & 1u) s_tk[32 + k] += 1;
                }
                s_tk[37] += W;
              }
            }
#endif
          }
          // (4) their rows enter the cache, their records go to the writer, the later pods learn about them
          for (int w = warp; w < n_ok; w += kDeciders) {
            const int g = f + w;
            const int idx = t_idx[w];
            const bool fresh = idx < 0;
            const int t = fresh ? ncr + __popc(newmask & ((1u << w) - 1u)) : idx;
            {
              const uint32_t* src = reinterpret_cast<const uint32_t*>(S.post + w);
              uint32_t* dst = reinterpret_cast<uint32_t*>(S.rows + t);
              for (int x = lane; x < (int)(sizeof(RowCache) / 4); x += 32) dst[x] = src[x];
            }
            if (lane == 0) s_erow[t] = s_pick[w];
            apply_column(t, g + 1, fresh, t_now[w], t_snap[w], t_spec[w], S.t_r0 + w * kMaxG);
            __syncwarp();
            __threadfence_block();
            if (lane == 0) s_ready[g] = 1;
          }
          if (tid == 0) {
            s_ncr = ncr + __popc(newmask);
            if (s_bad0) atomicOr(const_cast<uint32_t*>(&s_special), s_bad0 << f);
          }

          if (tid == 0) {
            s_tk[11] += 1;
            s_tk[12] += n_ok;
          }
          f += n_ok;
          KSB_F3(5);
          KSB_BAR3();
          KSB_F3(1);
        }
        if (tid == 0) {
          s_tk[15] += clock64() - t3_mark;
          s_consumed = f;
          last_index = L0 + (uint64_t)(f < Gc ? s_lrel[f] : s_lrel[Gc - 1] + ((s_prio >> (Gc - 1)) & 1u));
          __threadfence_block();
          s_stop = f;
        }
        if (warp == 0) {
          last_index = __shfl_sync(0xffffffffu, last_index, 0);
          if (lane < R) *PEER(lane).last_node_index = last_index;  // every rank's host reads its own copy
        }
      } else {
        // writer warp: table write-through, class counters and result records of every decided pod, in order
        bool table_open = false;
        for (int g = 0; g < Gc; ++g) {
          while (s_ready[g] == 0) {
            const int st = s_stop;
            if (st >= 0 && g >= st) break;
            __nanosleep(20);
          }
          if (s_ready[g] == 0) break;  // the group ended before this pod
          __threadfence_block();
          const CommitRec& rec = S.recs[g];
          const KsbPodHeader* h = S.hdr + g;
          // (sharded: this warp writes rank 0's copies only; the other ranks' are written by all warps once the group is
          //  decided, below, off this warp's serial path)
          if (lane == 0) PEER(0).results[p0 + g] = rec.res;
          if (rec.row >= 0) {
            // Pipelined: the other blocks may be reading the table for the next group while these rows are written.
            // That is harmless: every row written here stays in the row cache for that group (the carried rows), so
            // phase 3 re-evaluates it from the cache and only uses the group-start results on it to take its
            // contribution to the Reduce inputs and tie counts back out, which works for whatever mix of old and new
            // columns the evaluation happened to read: requests only grow, so a mix fits whenever the new row does.
            // (kHoldTableWrites keeps the conservative schedule: no write before every block has read the table.)
            if (kHoldTableWrites && pipe && !table_open) {
              if (lane == 0) spin_until<false>(A, A.barrier + 1, s_b1_target, 7);
              __syncwarp();
              table_open = true;
            }
            const PodView pv = make_view(h, S.pool + S.pool_off[g], T.W);
            const KsbPeer& P = PEER(rec.rank);
            const int chosen = rec.row;
            write_through(P.table, rec, chosen, h, lane);
            // sharded: rank 0 holds every row (phase 3 works on that copy); the rank that evaluates the row gets the
            // same columns over NVLink so that its next evaluation sees them
            const int own = MULTI ? min(R - 1, (chosen / kSeg) / sps) : 0;
            // the class counters a bind bumps (SelectorSpread column entries, inter-pod-affinity table slots): no pod
            // of this group reads them. SelectorSpread columns are sharded with the rows; affinity tables are
            // replicated on every rank.
            const size_t pn = P.table.n_pad;
            for (int i = lane; i < h->n_spread_inc; i += 32) {
              atomicAdd(P.table.spread + (size_t)pv.spread_inc[i] * pn + chosen, 1u);
            }
            for (int i = lane; i < h->n_commit; i += 32) {
              const uint64_t cm = pv.commits[2 * i];
              const int32_t delta = (int32_t)(int64_t)pv.commits[2 * i + 1];
              const uint32_t off = (uint32_t)cm;
              const uint32_t d = P.table.topo[((uint32_t)(cm >> 32) & 0xffu) * pn + chosen];
              int32_t* tb = PEER(0).table.tables;
              if (d) atomicAdd(tb + off + d, delta);
              if ((cm >> 40) & 1ull) atomicAdd(tb + off, delta);  // slot 0: total matching pods anywhere
            }
          }
          __syncwarp();
        }
      }
#undef KSB_BAR3
      if constexpr (MULTI) {
        // Sharded: what the other ranks keep of this group's commits, written by all warps at once, a pod per warp:
        // the result record on every rank, the committed row's columns and SelectorSpread counters on the rank that
        // evaluates the row, the affinity-table counters on every rank. Of several pods of the group on one row the warp
        // of the last one writes the row (all of them, in order).
        __syncthreads();
        const int decided = s_consumed;
        for (int g = warp; g < decided; g += kWarps) {
          const CommitRec& rec = S.recs[g];
          const KsbPodHeader* h = S.hdr + g;
          if (lane >= 1 && lane < R) PEER(lane).results[p0 + g] = rec.res;
          if (rec.row < 0) continue;
          const PodView pv = make_view(h, S.pool + S.pool_off[g], T.W);
          const int chosen = rec.row;
          const int own = min(R - 1, (chosen / kSeg) / sps);
          if (own != 0) {
            bool later = false;
            for (int g2 = g + 1; g2 < decided; ++g2) later = later || S.recs[g2].row == chosen;
            const KsbPeer& O = PEER(own);
            const int orow = chosen - own * sps * kSeg;
            if (!later) {
              // write_through stores only the columns its own pod touched (ports, scalars, devices), so the earlier
              // pods of the group on this row are replayed first, in order, by this warp
              for (int g1 = 0; g1 < g; ++g1)
                if (S.recs[g1].row == chosen) write_through(O.table, S.recs[g1], orow, S.hdr + g1, lane);
              write_through(O.table, rec, orow, h, lane);
            }
            for (int i = lane; i < h->n_spread_inc; i += 32)
              atomicAdd_system(O.table.spread + (size_t)pv.spread_inc[i] * O.table.n_pad + orow, 1u);
          }
          const size_t pn = PEER(0).table.n_pad;
          for (int i = lane; i < h->n_commit; i += 32) {
            const uint64_t cm = pv.commits[2 * i];
            const int32_t delta = (int32_t)(int64_t)pv.commits[2 * i + 1];
            const uint32_t off = (uint32_t)cm;
            const uint32_t d = PEER(0).table.topo[((uint32_t)(cm >> 32) & 0xffu) * pn + chosen];
            for (int q = 1; q < R; ++q) {
              int32_t* tb = PEER(q).table.tables;
              if (d) atomicAdd_system(tb + off + d, delta);
              if ((cm >> 40) & 1ull) atomicAdd_system(tb + off, delta);
            }
          }
        }
      }
#if KSB_FINE_TIMING == 1
      const long long drain0 = clock64();
#endif
      __syncthreads();
#if KSB_FINE_TIMING == 1
      if (tid == 0) s_tk[30] += clock64() - drain0;  // deciders done, writer warp still writing
#endif
      // publish how far the group got
      if (tid == 0) {
        A.group_done[gi & 1] = s_consumed;
        s_tk[9] += 1;
        if (s_consumed < Gc) s_tk[10] += 1;
      }
      if (!pipe) {  // reset the other scratch half for the next group
        KsbNorms* nn = A.norms + (size_t)(slot ^ 1) * kMaxG;
        for (int i = tid; i < kMaxG * 8; i += kThreads) reinterpret_cast<int32_t*>(nn)[i] = ((i & 7) == 7) ? -1 : 0;
        int32_t* zn = A.zone_sum + (size_t)(slot ^ 1) * kMaxG * zs;
        for (int i = tid; i < kMaxG * zs; i += kThreads) zn[i] = 0;
        int32_t* gh = A.ghist + (size_t)(slot ^ 1) * kMaxG * KSB_REASON_BITS;
        for (int i = tid; i < kMaxG * KSB_REASON_BITS; i += kThreads) gh[i] = 0;
      }
    }
    KSB_TICK(5);
    if (pipe && committer && Gc > 0) {
      // this group's scratch half is free again (its contents were copied to shared memory before phase 3)
      KsbNorms* nn = A.norms + (size_t)slot * kMaxG;
      for (int i = tid; i < kMaxG * 8; i += kThreads) reinterpret_cast<int32_t*>(nn)[i] = ((i & 7) == 7) ? -1 : 0;
      int32_t* zn = A.zone_sum + (size_t)slot * kMaxG * zs;
      for (int i = tid; i < kMaxG * zs; i += kThreads) zn[i] = 0;
      int32_t* gh = A.ghist + (size_t)slot * kMaxG * KSB_REASON_BITS;
      for (int i = tid; i < kMaxG * KSB_REASON_BITS; i += kThreads) gh[i] = 0;
    }
    // pipelined: the evaluation side has read the table and finished phase 2; the commit side has decided and written
    // stores into other GPUs' memory since the last barrier: the writer warp's copies (block 0 of rank 0) and, pipelined,
    // the evaluation results of ranks other than 0
    if (pipe) {
      if constexpr (MULTI) end_barrier_x(A, epoch, x2, xb, push_hist, committer, gi & 1);
      else grid_barrier(A, epoch, xepoch, true, xb);
    }
    else grid_barrier_x(A, epoch, xepoch, true, xb, push_done, NoXfer(), MULTI && committer, MULTI && A.rank == 0);
    KSB_TICK(6);
    if (__ldcg(A.error_flag) != 0) break;  // only after a barrier timeout: stop instead of spinning
    if (!pipe) {
      const int adv = __ldcv(A.l_group_done + (gi & 1));
      if (adv <= 0) {
        if (tid == 0) atomicCAS(A.error_flag, 0, 5);
        break;
      }
      ev_p0 += adv;
      ev_slot ^= 1;
    } else {
      // every block draws the same conclusions from the two published numbers
      const int ge = ev_valid ? __ldcv(A.l_group_done + 2 + (gi & 1)) : 0;  // pods evaluated in this iteration
      const int adv = cm_valid ? __ldcv(A.l_group_done + (gi & 1)) : 0;    // pods committed in this iteration
      if (cm_valid && adv < cm_G) {
        // speculation was dropped inside the committed group: everything evaluated after it is void. The next
        // evaluation starts at the first uncommitted pod, against a table that holds every commit so far.
        if (committer) {  // the void evaluation's half of the reduction scratch
          KsbNorms* nn = A.norms + (size_t)(cm_slot ^ 1) * kMaxG;
          for (int i = tid; i < kMaxG * 8; i += kThreads) reinterpret_cast<int32_t*>(nn)[i] = ((i & 7) == 7) ? -1 : 0;
          int32_t* zn = A.zone_sum + (size_t)(cm_slot ^ 1) * kMaxG * zs;
          for (int i = tid; i < kMaxG * zs; i += kThreads) zn[i] = 0;
          int32_t* gh = A.ghist + (size_t)(cm_slot ^ 1) * kMaxG * KSB_REASON_BITS;
          for (int i = tid; i < kMaxG * KSB_REASON_BITS; i += kThreads) gh[i] = 0;
        }
        grid_barrier(A, epoch, xepoch, true, xb);
        ev_p0 = cm_p0 + adv;
        cm_valid = false;
      } else if (ge > 0) {
        cm_carry = cm_valid;
        cm_floor = cm_valid ? cm_p0 : ev_p0;
        cm_p0 = ev_p0;
        cm_G = ge;
        cm_slot = ev_slot;
        cm_valid = true;
        ev_p0 += ge;
        ev_slot ^= 1;
      } else {
        cm_valid = false;  // nothing was evaluated: end of the stream, or the first pod waits for the commit above
      }
    }
    gi += 1;
  }
  if (b == 0 && tid == 0) {
#if KSB_FINE_TIMING < 2
    for (int i = 0; i < 7; ++i) A.counters[2 + i] = (uint64_t)s_tk[i];  // block 0's cycles per stage
#endif
#if KSB_FINE_TIMING != 3
    for (int i = 0; i < 5; ++i) A.counters[9 + i] = (uint64_t)s_tk[11 + i];  // phase-3 breakdown
#endif
#if KSB_FINE_TIMING == 1
    // instead: cycles of thread 0 in phase 3 by step: picks, own columns, wait for the other columns, finalize, one-pod path
    A.counters[9] = (uint64_t)s_tk[18];
    A.counters[10] = (uint64_t)s_tk[19];
    A.counters[11] = (uint64_t)(s_tk[20] + s_tk[17]);
    A.counters[12] = (uint64_t)s_tk[21];
    A.counters[13] = (uint64_t)s_tk[23];
    for (int i = 0; i < 7; ++i) A.counters[2 + i] = (uint64_t)s_tk[24 + i];  // granular: picks 0-2, columns 3-5, [6] wait for the writer warp
#ifdef KSB_CAUSES  // instead: why windows ended early (special, became holder, stopped being holder, same pick, bad pick) and pods tried
    for (int i = 0; i < 4; ++i) A.counters[2 + i] = (uint64_t)s_tk[32 + i] * 1965ull;
    A.counters[6] = (uint64_t)s_tk[38] * 1965ull;  // pairs
    A.counters[7] = (uint64_t)s_tk[39] * 1965ull;  // pairs whose row was a holder for the later pod
    A.counters[8] = (uint64_t)s_tk[36] * 1965ull;  // ... and stopped being one
#endif
#endif
    if (A.rank == 0) {
      A.counters[0] = (uint64_t)s_tk[9];
      A.counters[1] = (uint64_t)s_tk[10];
    }
  }
  if (multi && b == xb && tid == 0) {
    A.xstate[0] = xepoch;
    A.xstate[1] = x2;
    A.xstate[2] = x1;
  }
#if KSB_FINE_TIMING == 2
  if (b == (pipe ? 1 : 0) && tid == 0)  // FINE_TIMING=2: an evaluating block's stage cycles instead of block 0's
    for (int i = 0; i < 7; ++i) A.counters[2 + i] = (uint64_t)s_tk[i];
#endif
#if KSB_FINE_TIMING == 3
  if (multi && b == xb && tid == 0)
    for (int i = 0; i < 7; ++i) {
      A.counters[2 + i] = (uint64_t)g_xdbg[i];
      g_xdbg[i] = 0;
    }
  if (multi && b == xb && tid == 0) {  // the same block's stage cycles next to them
    A.counters[9] = (uint64_t)s_tk[0];
    A.counters[10] = (uint64_t)s_tk[1];
    A.counters[11] = (uint64_t)s_tk[2];
    A.counters[12] = (uint64_t)s_tk[3];
    A.counters[13] = (uint64_t)s_tk[6];
  }
#endif
#undef KSB_TICK
#undef PEER
#undef cur_norms
#undef cur_zsum
#undef cur_acc_norms
#undef cur_acc_zsum
#undef cur_tot_norms
#undef cur_tot_zsum
#undef cur_acc_ghist
#undef raw_s
#undef score_s
#undef reasons_s
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

extern "C" size_t ksb_batch_smem_bytes(int zone_stride) {
  size_t o[kSmemParts];
  return smem_layout(o, zone_stride);
}

static const void* batch_kernel(bool pipe, bool multi, bool pol) {
  static const void* const fns[8] = {
      (const void*)ksb_schedule_batch_kernel<false, false, false>, (const void*)ksb_schedule_batch_kernel<true, false, false>,
      (const void*)ksb_schedule_batch_kernel<false, true, false>,  (const void*)ksb_schedule_batch_kernel<true, true, false>,
      (const void*)ksb_schedule_batch_kernel<false, false, true>,  (const void*)ksb_schedule_batch_kernel<true, false, true>,
      (const void*)ksb_schedule_batch_kernel<false, true, true>,   (const void*)ksb_schedule_batch_kernel<true, true, true>};
  return fns[(pipe ? 1 : 0) + (multi ? 2 : 0) + (pol ? 4 : 0)];
}

extern "C" int ksb_launch_batch(const KsbBatchArgs* args, int grid, size_t smem_bytes, cudaStream_t stream) {
  void* params[] = {(void*)args};
  // (the instantiations with the Policy plugins also carry the words above 0 of the taint / preferAvoid sets)
  const bool pol = args->table.w_mr != 0 || args->table.w_rl != 0 || args->table.w_il != 0 || args->table.has_static != 0 || args->table.has_presence != 0 ||
                   args->table.tw > 1 || args->table.ext_keep != nullptr || args->table.ext_score != nullptr;
  const void* fn = batch_kernel(args->pipeline != 0, args->n_ranks > 1, pol);
  cudaError_t e = cudaLaunchCooperativeKernel(fn, dim3(grid), dim3(kThreads), params,
                                              smem_bytes, stream);
  return (int)e;
}

extern "C" int ksb_batch_kernel_limits(int device, size_t smem_bytes, int* max_grid, int* threads) {
  int per_sm = 0, sms = 0;
  // every instantiation (plain / pipelined, one GPU / sharded, with / without Policy plugins) shares one grid size: the
  // smallest occupancy of them all
  const void* fns[8];
  for (int k = 0; k < 8; ++k) fns[k] = batch_kernel((k & 1) != 0, (k & 2) != 0, (k & 4) != 0);
  cudaError_t e = cudaSuccess;
  per_sm = 1 << 30;
  for (const void* fn : fns) {
    e = cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_bytes);
    if (e != cudaSuccess) return (int)e;
    int occ = 0;
    e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, fn, kThreads, smem_bytes);
    if (e != cudaSuccess) return (int)e;
    per_sm = occ < per_sm ? occ : per_sm;
  }
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

extern "C" int ksb_max_pool_words() { return kMaxPoolWords; }

// A label bit interned after the table went up (first pod whose selector mentions it) is set on the device instead of
// re-sending the 64-bit column it lives in: from the per-key column of parsed numeric values for Gt / Lt bits
// (selector.go:214-233), from a list of rows for everything that matches few rows.
__global__ void ksb_label_threshold_kernel(uint64_t* col, const int64_t* num, const uint8_t* valid, int n, uint64_t mask, int greater,
                                           int64_t x) {
  const int r = blockIdx.x * blockDim.x + threadIdx.x;
  if (r < n && valid[r] && (greater ? num[r] > x : num[r] < x)) col[r] |= mask;
}
__global__ void ksb_label_scatter_kernel(uint64_t* col, const int32_t* rows, int k, uint64_t mask) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < k) col[rows[i]] |= mask;  // a row appears once in the list
}
extern "C" int ksb_launch_label_threshold(uint64_t* col, const int64_t* num, const uint8_t* valid, int n, uint64_t mask, int greater,
                                          int64_t x, cudaStream_t stream) {
  if (n <= 0) return 0;
  ksb_label_threshold_kernel<<<(n + 255) / 256, 256, 0, stream>>>(col, num, valid, n, mask, greater, x);
  return (int)cudaGetLastError();
}
extern "C" int ksb_launch_label_scatter(uint64_t* col, const int32_t* rows, int k, uint64_t mask, cudaStream_t stream) {
  if (k <= 0) return 0;
  ksb_label_scatter_kernel<<<(k + 255) / 256, 256, 0, stream>>>(col, rows, k, mask);
  return (int)cudaGetLastError();
}

// ---------------------------------------------------------------------------------------------
// Scratch for ksb_replay_last_batch_noalloc (include/ksched_b200.h).
//
// Some users of this library drive it from places where asking the CUDA runtime for device memory
// is off the table: a control loop whose memory budget was fixed at start-up (a MIG slice or an MPS
// client with a hard limit, where a late cudaMalloc may simply fail), or a callback running on
// somebody else's stream where cudaMalloc's implicit device-wide synchronisation would stall
// unrelated work. For them the batch kernel's working memory lives here, in one array that is part
// of the loaded module: every GPU that loads the library carries these bytes until the process exits.
// The limits below bound what fits; larger jobs must use ksb_replay_last_batch. To keep the array
// small this path runs the kernel with groups of one pod (same results, by construction).
//
// All calls share the one array and it keeps whatever the previous call left behind, so the host
// side re-zeroes the counters the kernel accumulates into before each launch, and two calls whose
// kernels overlap on the same GPU would corrupt each other.
#define KSB_NOALLOC_MAX_ROWS 65536    // padded node rows
#define KSB_NOALLOC_MAX_ZONES 1024    // zone-sum slots
#define KSB_NOALLOC_MAX_GRID 1024     // thread blocks
namespace {
constexpr size_t na_align(size_t x) { return (x + 255) & ~(size_t)255; }
constexpr size_t kNaSegs = KSB_NOALLOC_MAX_ROWS / kSeg;
constexpr size_t kNaNorms = 0;
constexpr size_t kNaZone = kNaNorms + na_align(2 * kMaxG * sizeof(KsbNorms));
constexpr size_t kNaZoneBig = 2 * (size_t)kMaxG * KSB_NOALLOC_MAX_ZONES * 4;
constexpr size_t kNaRawBase = kNaZone + na_align(kNaZoneBig);
constexpr size_t kNaScore = kNaRawBase + na_align((size_t)KSB_NOALLOC_MAX_ROWS * 16);
constexpr size_t kNaReasons = kNaScore + na_align((size_t)KSB_NOALLOC_MAX_ROWS * 4);
constexpr size_t kNaSegMax = kNaReasons + na_align((size_t)KSB_NOALLOC_MAX_ROWS * 4);
constexpr size_t kNaSegCnt = kNaSegMax + na_align(kNaSegs * 4);
constexpr size_t kNaSegMask = kNaSegCnt + na_align(kNaSegs * 4);
constexpr size_t kNaBarrier = kNaSegMask + na_align(kNaSegs * 32);
constexpr size_t kNaError = kNaBarrier + 256;
constexpr size_t kNaFitCount = kNaError + 256;
constexpr size_t kNaLastIndex = kNaFitCount + 256;
constexpr size_t kNaGroupDone = kNaLastIndex + 256;
constexpr size_t kNaCounters = kNaGroupDone + 256;
constexpr size_t kNaRelease = kNaCounters + 256;  // counters: 16 x 8 bytes
constexpr size_t kNaXbar = kNaRelease + 256;
constexpr size_t kNaXstate = kNaXbar + 256;
constexpr size_t kNaPeer = kNaXstate + 256;
constexpr size_t kNaGhist = kNaPeer + na_align(sizeof(KsbPeer));
constexpr size_t kNaFitHist = kNaGhist + na_align(2 * (size_t)kMaxG * KSB_REASON_BITS * 4);
constexpr size_t kNaTotal = kNaFitHist + na_align((size_t)KSB_MAX_FITERR * KSB_REASON_BITS * 4);
}  // namespace

static __device__ __align__(256) unsigned char ksb_noalloc_scratch[kNaTotal];  // about 2.4 MiB

// In host code `ksb_noalloc_scratch` is only a stand-in for the symbol, not a device address, so the
// address each GPU gave the array is looked up once per device (with that device current) and cached.
extern "C" int ksb_noalloc_bind(int device, int n_pad, int zone_stride, int grid, KsbBatchArgs* a, char* err,
                                size_t err_len) {
  static unsigned char* base_of[64] = {nullptr};
  auto fail = [&](const char* msg) {
    if (err && err_len) snprintf(err, err_len, "%s", msg);
    return -1;
  };
  if (device < 0 || device >= 64) return fail("device ordinal out of range");
  if (n_pad > KSB_NOALLOC_MAX_ROWS || zone_stride > KSB_NOALLOC_MAX_ZONES || grid > KSB_NOALLOC_MAX_GRID) {
    if (err && err_len)
      snprintf(err, err_len,
               "batch exceeds the no-allocation limits: %d padded node rows (max %d), %d zone slots (max %d), "
               "%d blocks (max %d); use ksb_replay_last_batch",
               n_pad, KSB_NOALLOC_MAX_ROWS, zone_stride, KSB_NOALLOC_MAX_ZONES, grid, KSB_NOALLOC_MAX_GRID);
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
  a->group_size = 1;  // one pod per group: the stashes need one row set, not 32
  a->norms = (KsbNorms*)(b + kNaNorms);
  a->zone_sum = (int32_t*)(b + kNaZone);
  a->raw = (int4*)(b + kNaRawBase);
  a->score = (int32_t*)(b + kNaScore);
  a->reasons = (uint32_t*)(b + kNaReasons);
  a->seg_max = (int32_t*)(b + kNaSegMax);
  a->seg_cnt = (int32_t*)(b + kNaSegCnt);
  a->seg_mask = (uint32_t*)(b + kNaSegMask);
  a->barrier = (uint32_t*)(b + kNaBarrier);
  a->error_flag = (int32_t*)(b + kNaError);
  a->fiterr_count = (int32_t*)(b + kNaFitCount);
  a->last_node_index = (uint64_t*)(b + kNaLastIndex);
  a->group_done = (int32_t*)(b + kNaGroupDone);
  a->counters = (uint64_t*)(b + kNaCounters);
  a->ghist = (int32_t*)(b + kNaGhist);
  a->release = (uint32_t*)(b + kNaRelease);
  a->xbar = (uint32_t*)(b + kNaXbar);
  a->xstate = (uint32_t*)(b + kNaXstate);
  a->peers = (const KsbPeer*)(b + kNaPeer);
  a->fiterr_hist = (int32_t*)(b + kNaFitHist);
  a->fiterr_cap = KSB_MAX_FITERR;
  a->acc_norms = a->norms;
  a->tot_norms = a->norms;
  a->acc_zone_sum = a->zone_sum;
  a->tot_zone_sum = a->zone_sum;
  a->acc_ghist = a->ghist;
  a->l_group_done = a->group_done;
  a->gs_seg_max = a->seg_max;
  a->gs_seg_cnt = a->seg_cnt;
  a->gs_seg_mask = a->seg_mask;
  a->gs_np = n_pad;
  a->gs_score = a->score;  // (read by sharded kernels only: one rank reads score / raw / reasons)
  a->gs_reasons = a->reasons;
  a->gs_base0 = nullptr;
  a->gs_row0 = 0;
  a->gs_nseg = a->nseg;
  a->seg_gbase = 0;
  a->eval_n = a->table.n;
  return 0;
}

// ------------------------------------------------------------------ preemption
// selectVictimsOnNode (core/generic_scheduler.go:883-953) for every node where preemption might help (:957-991), one
// thread per row: take every lower-priority pod off the node's aggregates, check that the preemptor fits, order the
// potential victims by decreasing priority exactly as the reference's sort.Sort does (go1.9 src/sort/sort.go,
// restated below: the order among equal priorities decides who is reprieved), then put them back one by one as long
// as the preemptor still fits. pickOneNodeForPreemption (:654-758) runs on the host over the per-row statistics.
// Scope (enforced by Engine::preempt): no inter-pod affinity in play, no host ports on the preemptor, no PDBs.
namespace {
constexpr uint32_t kUnresolvable = (1u << RB_SELECTOR) | (1u << RB_HOSTNAME) | (1u << RB_TAINTS) | (1u << RB_LABEL_PRESENCE) |
                                   (1u << RB_NOT_READY) | (1u << RB_NET_UNAVAILABLE) | (1u << RB_UNSCHEDULABLE);
// bits that removing pods can clear, and bits podFitsOnNode never sets (the fork's device selection runs after it)
constexpr uint32_t kByRemoval = 0x1ffu;  // RB_PODS .. RB_SCALAR3
constexpr uint32_t kNotPodFits = (0xfu << RB_EXT0) | (1u << RB_EXT_LIMITS);

struct NodeAgg {
  int64_t cpu, mem, agpu, eph, scalar[KSB_MAX_SCALARS];
  int32_t pods;
};

struct GoSortDev {  // sort.Sort over idx[0..n): Less(i, j) = priority(idx[i]) > priority(idx[j])
  int32_t* idx;
  const KsbVictimPod* pods;
  __device__ bool less(int i, int j) const { return pods[idx[i]].priority > pods[idx[j]].priority; }
  __device__ void swap(int i, int j) {
    const int32_t t = idx[i];
    idx[i] = idx[j];
    idx[j] = t;
  }
  __device__ void insertionSort(int a, int b) {
    for (int i = a + 1; i < b; i++)
      for (int j = i; j > a && less(j, j - 1); j--) swap(j, j - 1);
  }
  __device__ void siftDown(int lo, int hi, int first) {
    int root = lo;
    for (;;) {
      int child = 2 * root + 1;
      if (child >= hi) return;
      if (child + 1 < hi && less(first + child, first + child + 1)) child++;
      if (!less(first + root, first + child)) return;
      swap(first + root, first + child);
      root = child;
    }
  }
  __device__ void heapSort(int a, int b) {
    const int first = a, lo = 0, hi = b - a;
    for (int i = (hi - 1) / 2; i >= 0; i--) siftDown(i, hi, first);
    for (int i = hi - 1; i >= 0; i--) {
      swap(first, first + i);
      siftDown(lo, i, first);
    }
  }
  __device__ void medianOfThree(int m1, int m0, int m2) {
    if (less(m1, m0)) swap(m1, m0);
    if (less(m2, m1)) {
      swap(m2, m1);
      if (less(m1, m0)) swap(m1, m0);
    }
  }
  __device__ void doPivot(int lo, int hi, int* midlo, int* midhi) {
    const int m = (int)((unsigned)(lo + hi) >> 1);
    if (hi - lo > 40) {
      const int s = (hi - lo) / 8;
      medianOfThree(lo, lo + s, lo + 2 * s);
      medianOfThree(m, m - s, m + s);
      medianOfThree(hi - 1, hi - 1 - s, hi - 1 - 2 * s);
    }
    medianOfThree(lo, m, hi - 1);
    const int pivot = lo;
    int a = lo + 1, c = hi - 1;
    for (; a < c && less(a, pivot); a++) {
    }
    int b = a;
    for (;;) {
      for (; b < c && !less(pivot, b); b++) {
      }
      for (; b < c && less(pivot, c - 1); c--) {
      }
      if (b >= c) break;
      swap(b, c - 1);
      b++;
      c--;
    }
    bool protect = hi - c < 5;
    if (!protect && hi - c < (hi - lo) / 4) {
      int dups = 0;
      if (!less(pivot, hi - 1)) {
        swap(c, hi - 1);
        c++;
        dups++;
      }
      if (!less(b - 1, pivot)) {
        b--;
        dups++;
      }
      if (!less(m, pivot)) {
        swap(m, b - 1);
        b--;
        dups++;
      }
      protect = dups > 1;
    }
    if (protect) {
      for (;;) {
        for (; a < b && !less(b - 1, pivot); b--) {
        }
        for (; a < b && less(a, pivot); a++) {
        }
        if (a >= b) break;
        swap(a, b - 1);
        a++;
        b--;
      }
    }
    swap(pivot, b - 1);
    *midlo = b - 1;
    *midhi = c;
  }
  // quickSort(data, 0, n, maxDepth(n)) with the recursion on the smaller side kept on an explicit stack (the ranges
  // are disjoint, so the order in which they are finished does not change the result)
  __device__ void sort(int n) {
    int depth = 0;
    for (int i = n; i > 0; i >>= 1) depth++;
    int stack[3 * 24];
    int sp = 0;
    stack[sp++] = 0;
    stack[sp++] = n;
    stack[sp++] = depth * 2;
    while (sp > 0) {
      int maxDepth = stack[--sp], b = stack[--sp], a = stack[--sp];
      bool done = false;
      while (b - a > 12) {
        if (maxDepth == 0) {
          heapSort(a, b);
          done = true;
          break;
        }
        maxDepth--;
        int mlo, mhi;
        doPivot(a, b, &mlo, &mhi);
        if (mlo - a < b - mhi) {
          stack[sp++] = a;  // the smaller side: at most lg(n) ranges wait at a time
          stack[sp++] = mlo;
          stack[sp++] = maxDepth;
          a = mhi;
        } else {
          stack[sp++] = mhi;
          stack[sp++] = b;
          stack[sp++] = maxDepth;
          b = mlo;
        }
      }
      if (done) continue;
      if (b - a > 1) {
        for (int i = a + 6; i < b; i++)
          if (less(i, i - 6)) swap(i, i - 6);
        insertionSort(a, b);
      }
    }
  }
};
}  // namespace

__global__ void ksb_preempt_kernel(KsbPreemptArgs A) {
  const KsbNodeTable& T = A.table;
  int r = blockIdx.x * blockDim.x + threadIdx.x;
  if (A.only_row >= 0) {
    if (r != 0) return;
    r = A.only_row;
  }
  if (r >= T.n) return;
  const size_t np = T.n_pad;
  const uint32_t mask = A.reasons[r];
  int32_t state = 0, nv = 0, highest = 0;
  int64_t sum = 0;
  if (!(mask & kUnresolvable)) {  // nodesWherePreemptionMightHelp
    state = 1;
    if (!(mask & ~(kByRemoval | kNotPodFits))) {  // nothing fails that removing pods cannot change
      NodeAgg g;
      g.cpu = T.req_cpu[r];
      g.mem = T.req_mem[r];
      g.agpu = T.req_agpu[r];
      g.eph = T.req_eph[r];
      for (int s = 0; s < KSB_MAX_SCALARS; ++s) g.scalar[s] = T.req_scalar[s * np + r];
      g.pods = T.num_pods[r];
      const int64_t acpu = T.alloc_cpu[r], amem = T.alloc_mem[r], aagpu = T.alloc_agpu[r], aeph = T.alloc_eph[r];
      const int32_t apods = T.alloc_pods[r];
      auto fits = [&]() {  // PodFitsResources (predicates.go:630-687)
        if (g.pods + 1 > apods) return false;
        if (!A.has_request) return true;
        if (acpu < A.req_cpu + g.cpu || amem < A.req_mem + g.mem || aagpu < A.req_agpu + g.agpu || aeph < A.req_eph + g.eph) return false;
        for (int s = 0; s < KSB_MAX_SCALARS; ++s)
          if (((A.scalar_mask >> s) & 1u) && T.alloc_scalar[s * np + r] < A.scalar_req[s] + g.scalar[s]) return false;
        return true;
      };
      auto remove = [&](const KsbVictimPod& p) {  // NodeInfo.RemovePod: ephemeral storage is never given back (node_info.go:396-398)
        g.cpu -= p.cpu;
        g.mem -= p.mem;
        g.agpu -= p.agpu;
        for (int s = 0; s < KSB_MAX_SCALARS; ++s) g.scalar[s] -= p.scalar[s];
        g.pods -= 1;
      };
      auto add = [&](const KsbVictimPod& p) {  // NodeInfo.AddPod
        g.cpu += p.cpu;
        g.mem += p.mem;
        g.agpu += p.agpu;
        g.eph += p.eph;
        for (int s = 0; s < KSB_MAX_SCALARS; ++s) g.scalar[s] += p.scalar[s];
        g.pods += 1;
      };
      int32_t idx[KSB_MAX_VICTIMS];
      int n = 0;
      bool overflow = false;
      for (int i = A.row_start[r]; i < A.row_start[r + 1]; ++i)
        if (A.pods[i].priority < A.priority) {  // NodeInfo.pods order (see oracle/preempt.hpp on the range-while-removing loop)
          if (n == KSB_MAX_VICTIMS) {
            overflow = true;
            break;
          }
          idx[n++] = i;
          remove(A.pods[i]);
        }
      if (overflow) {
        atomicExch(A.error_flag, 1);
      } else {
        GoSortDev gs{idx, A.pods};
        gs.sort(n);
        if (fits()) {
          state = 2;
          for (int k = 0; k < n; ++k) {  // reprieve from the highest priority down
            const KsbVictimPod& p = A.pods[idx[k]];
            add(p);
            if (!fits()) {
              remove(p);
              if (nv == 0) highest = p.priority;
              sum += (int64_t)p.priority + 2147483648ll;
              if (A.only_row >= 0) A.victim_list[nv] = idx[k];
              nv += 1;
            }
          }
        }
      }
    }
  }
  A.row_state[r] = state;
  A.row_victims[r] = nv;
  A.row_highest[r] = highest;
  A.row_sum[r] = sum;
}

extern "C" int ksb_launch_preempt(const KsbPreemptArgs* args, cudaStream_t stream) {
  const int n = args->only_row >= 0 ? 1 : args->table.n;
  if (n <= 0) return 0;
  ksb_preempt_kernel<<<(n + 127) / 128, 128, 0, stream>>>(*args);
  return (int)cudaGetLastError();
}

