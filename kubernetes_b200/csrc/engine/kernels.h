// Host/device interface of kernels.cu.
#pragma once
#include <cuda_runtime.h>
#include <stddef.h>
#include <stdint.h>

#include "device_types.h"

// One shard of the node table as seen from the GPU that runs phase 3 (rank 0). For rank 0 itself these are
// its own buffers; for the other ranks they are CUDA-IPC mappings of the peer's buffers (NVLink P2P).
struct KsbPeer {
  KsbNodeTable table;
  int4* raw;
  int32_t* score;
  uint32_t* reasons;
  int32_t* seg_max;
  int32_t* seg_cnt;
  uint32_t* seg_mask;
  KsbResult* results;
  uint64_t* last_node_index;
  int32_t nseg;      // 256-row segments of this shard
  int32_t seg_base;  // its first segment in the global numbering
  int32_t raw_np;    // row stride of `raw` (the rank's own evaluation range, padded)
  int32_t raw_row0;  // global row of raw's row 0
};

struct KsbBatchArgs {
  KsbNodeTable table;
  const KsbPodHeader* headers;  // [n_pods]
  const uint64_t* pool;
  const int32_t* dep;           // [n_pods] latest earlier pod writing a counter this pod reads, -1 if none
  int32_t n_pods;
  int32_t zone_stride;          // ints per pod in the zone-sum arrays (>= table.n_zones)
  int32_t commit;               // 1: apply the assume step after each pod (ksb_schedule_batch)
  int32_t group_size;           // pods per group, 1..KSB_MAX_GROUP
  int32_t nseg;                 // 256-row segments in this shard
  int32_t pipeline;             // 1: block 0 commits group k-1 while the other blocks evaluate group k (one rank only);
                                // raw/score/reasons/seg_* then hold two halves, one per parity of the group
  KsbResult* results;           // [n_pods]
  // scratch (global memory beyond inputs and outputs)
  KsbNorms* norms;              // [2][KSB_MAX_GROUP] ping-pong per group
  int32_t* zone_sum;            // [2][KSB_MAX_GROUP][zone_stride]
  int4* raw;                    // [group_size][n_pad] phase-1 Map results
  int32_t* score;               // [group_size][n_pad] phase-2 totals
  uint32_t* reasons;            // [group_size][n_pad] failure-reason masks
  int32_t* seg_max;             // [group_size][nseg]
  int32_t* seg_cnt;             // [group_size][nseg]
  uint32_t* seg_mask;           // [group_size][nseg][8] rows of the segment that hold its maximum, one bit per row
  uint32_t* barrier;            // [2] grid barrier counters, zero at launch: [1] counts the barrier between phases 1 and 2, [0] all others
  int32_t* error_flag;
  int32_t* fiterr_count;
  int32_t* fiterr_hist;         // [fiterr_cap][KSB_REASON_BITS] reason histograms of the launch's no-fit pods
  int32_t fiterr_cap;           // histograms the buffer holds (>= KSB_MAX_FITERR; the engine sizes it to the launch)
  uint64_t* last_node_index;    // genericScheduler.lastNodeIndex
  int32_t* group_done;          // [8] pods consumed by the group, ping-pong; pipelined: [2..3] size of the group evaluated, [4] block 0's SM, [5] the block sharing it
  uint64_t* counters;           // [16] groups run, groups cut short by a dropped speculation, stage cycles
  int32_t* ghist;               // [2][KSB_MAX_GROUP][KSB_REASON_BITS] reason histograms of no-fit pods, ping-pong
  // Where the per-segment statistics of phase 2 go: this GPU's own arrays with one rank; with several ranks rank 0's
  // arrays (every rank writes its segments' statistics there over NVLink, indexed by GLOBAL segment), because phase 3
  // runs on rank 0 against its own full copy of the node table. The per-(pod, row) results stay where they were
  // computed; phase 3 reads the few it needs (changed rows) from there.
  int32_t* gs_seg_max;          // [halves * group_size][gs_nseg]
  int32_t* gs_seg_cnt;
  uint32_t* gs_seg_mask;
  int32_t gs_np, gs_nseg;
  // Sharded: the per-(pod, row) results phase 3 reads of a changed row (group-start total, its bind-dependent part and
  // the reason mask) are stored by every rank straight into rank 0's arrays at the GLOBAL row, so that the commit
  // never reads over NVLink. One rank: score / reasons themselves (base0 unused: raw[].x is read instead).
  int32_t* gs_score;            // [halves * group_size][gs_np]
  int32_t* gs_base0;            // [halves * group_size][gs_np] raw[].x: INT32_MIN = the row did not fit
  uint32_t* gs_reasons;         // [halves * group_size][gs_np]
  int32_t gs_row0;              // global row of this rank's evaluation row 0
  int32_t seg_gbase;            // global index of this rank's first evaluated segment
  int32_t eval_n;               // rows this rank evaluates: rows [0, eval_n) of its table (rank 0's table holds all rows)
  // Reduce inputs: accumulated per GPU, pushed to rank 0's targets (norms / zone_sum / ghist above) at a barrier,
  // read back as totals. On rank 0 and with one rank all of these are the targets themselves.
  KsbNorms* acc_norms;
  KsbNorms* tot_norms;
  int32_t* acc_zone_sum;
  int32_t* tot_zone_sum;
  int32_t* acc_ghist;
  int32_t* l_group_done;        // this GPU's copy of group_done
  // sharding (n_ranks == 1: everything local). With n_ranks > 1 the reduction targets above (norms, zone_sum,
  // ghist, group_done) and xbar point into rank 0's memory on every rank.
  const KsbPeer* peers;         // [n_ranks], dereferenced by rank 0 only
  uint32_t* xbar;               // (unused since the flag barrier; kept so that the control block keeps its layout)
  // Cross-GPU barrier and Reduce exchange, all-to-all: every rank owns an inbox in its scratch buffer with one slot
  // per source rank; a rank stores its partial Reduce inputs into its slot of EVERY inbox, then its arrival number
  // into its word of every rank's flag array, and waits on its OWN flag array (local polling, no NVLink read on
  // the path); the totals are then combined from the local inbox. xpeer[q] = rank q's scratch buffer as mapped
  // here (own buffer for q == rank), xo_* the byte offsets of the inbox parts inside every scratch buffer.
  char* xpeer[KSB_MAX_RANKS];
  uint32_t xo_flag;             // uint32_t [KSB_MAX_RANKS]: arrival number last written by each rank
                                // (+64: [KSB_MAX_RANKS] "evaluation done" numbers, read by rank 0's committing block;
                                //  +128: two 8-byte words {pods consumed, number} "commit done", written by rank 0)
  uint32_t xo_norms;            // uint64_t [2][KSB_MAX_RANKS][KSB_MAX_GROUP * 8]: {value, exchange number} words
  uint32_t xo_zone;             // uint64_t [2][KSB_MAX_RANKS][KSB_MAX_GROUP][zone_stride]
  uint32_t xo_gdone;            // int32_t [8]: the rank's copy of group_done (rank 0 pushes words 0..1)
  uint32_t* release;            // local [2]: per counter, the arrival count up to which this GPU's blocks may proceed
  uint32_t* xstate;             // local [3]: the cross-GPU arrival numbers carried from launch to launch (the flags are never reset)
  int32_t n_ranks, rank;
  int32_t segs_per_shard;       // segments per shard (the last shard may hold fewer)
  int32_t nseg_total;           // segments over all shards
  // this GPU's own shard in the form of a peer entry (filled at launch). With one rank phase 3 reads it from the
  // kernel's parameter space instead of going through the peer table in shared memory.
  KsbPeer self;
};

extern "C" {
int ksb_launch_batch(const KsbBatchArgs* args, int grid, size_t smem_bytes, cudaStream_t stream);
int ksb_batch_kernel_limits(int device, size_t smem_bytes, int* max_grid, int* threads);
int ksb_launch_apply(const KsbNodeTable* T, const KsbPodHeader* header, const uint64_t* pool, int row,
                     const uint32_t* masks, cudaStream_t stream);
size_t ksb_batch_smem_bytes(int zone_stride);
int ksb_max_pool_words();
// ---- preemption (core/generic_scheduler.go:199-256, :654-991), one thread per node row
// A cached pod as selectVictimsOnNode needs it: what NodeInfo.RemovePod / AddPod move (node_info.go:329-419) and its
// priority. Pods of one node are consecutive, in NodeInfo.pods order.
struct KsbVictimPod {
  int64_t cpu, mem, agpu, eph, nz_cpu, nz_mem;
  int64_t scalar[KSB_MAX_SCALARS];
  int32_t priority;
  int32_t pad;
};
struct KsbPreemptArgs {
  KsbNodeTable table;
  const uint32_t* reasons;       // [n] FailedPredicateMap of the Schedule call that failed: one reason mask per row
  const KsbVictimPod* pods;      // CSR values
  const int32_t* row_start;      // [n + 1] CSR offsets
  // the preemptor (GetResourceRequest, predicates.go:630-687)
  int64_t req_cpu, req_mem, req_agpu, req_eph;
  int64_t scalar_req[KSB_MAX_SCALARS];
  uint32_t scalar_mask;
  int32_t has_request;
  int32_t priority;
  int32_t only_row;              // -1: every row, statistics only; else: this row, with its victim list
  // per row: state (0 preemption cannot help, 1 potential node, 2 candidate: the pod fits once victims are gone),
  // number of victims, highest victim priority; sum of (priority + 2^31) over the victims
  int32_t* row_state;
  int32_t* row_victims;
  int32_t* row_highest;
  int64_t* row_sum;
  int32_t* victim_list;          // only_row >= 0: the row's victims as indices into pods[], decreasing priority
  int32_t* error_flag;           // 1: a node holds more lower-priority pods than KSB_MAX_VICTIMS
};
#define KSB_MAX_VICTIMS 512
int ksb_launch_preempt(const KsbPreemptArgs* args, cudaStream_t stream);

// Set one label bit on the device (a bit interned after the table was uploaded): rows whose parsed numeric label value
// is greater / less than x, or the listed rows.
int ksb_launch_label_threshold(uint64_t* col, const int64_t* num, const uint8_t* valid, int n, uint64_t mask, int greater, int64_t x,
                               cudaStream_t stream);
int ksb_launch_label_scatter(uint64_t* col, const int32_t* rows, int k, uint64_t mask, cudaStream_t stream);
// Point the scratch fields of *a at the module's static device array (no allocation) and set the group
// size that array supports; -1 + text on error, including batches beyond the array's stated limits.
int ksb_noalloc_bind(int device, int n_pad, int zone_stride, int grid, KsbBatchArgs* a, char* err, size_t err_len);
}
