// Layout of the GPU-resident structure-of-arrays mirror of schedulercache.NodeInfo
// (reference: plugin/pkg/scheduler/schedulercache/node_info.go:35-78) and of the per-pod
// requirement block (reference: algorithm/predicates/metadata.go:45-101 predicateMetadata +
// algorithm/priorities/metadata.go:45-66 priorityMetadata). Shared by the host encoder and
// the kernels. Rows are node names in bytewise ascending order, so selectHost's tie order
// (descending host name, api/types.go:218-223) is descending row order.
#pragma once
#include <stdint.h>

#define KSB_MAX_SCALARS 4     // interned scalar resource names (nvidia.com/gpu, hugepages-*, ...)
#define KSB_MAX_RES 2         // interned extended (device) resource names
#define KSB_MAX_TOPO 8        // interned topology keys
#define KSB_PORT_WORDS 4      // 256 interned proto/port slots and 256 proto/ip/port slots
#define KSB_MAX_TERMS 8       // required node-affinity terms / preferred terms per pod
#define KSB_MAX_EXPRS 32      // label expressions per pod
#define KSB_CLASS_WORDS 4     // 256 interned device attribute classes
#define KSB_MAX_EXT 4         // pod.Spec.ExtendedResources entries per pod
#define KSB_MAX_DEV 32        // devices per (node, resource)
#define KSB_MAX_FITERR 4096   // no-fit pods per batch that keep a reason histogram
#define KSB_REASON_BITS 32
#define KSB_MAX_GROUP 32      // pods evaluated against one snapshot of the table
#define KSB_MAX_SEGMENTS 2048 // 256-row segments per shard (524,288 node rows)
#define KSB_MAX_TAINT_WORDS 3  // 192 distinct NoSchedule/NoExecute taints, PreferNoSchedule taints and preferAvoidPods references
#define KSB_MAX_RANKS 8        // shards of one node table (one NVSwitch domain)
#define KSB_SEG_CACHE 4096    // (group size) x (segments) bound: segment statistics held in shared memory

// node flags (predicates.go:1382-1441)
#define NF_NOT_READY 1u
#define NF_OUT_OF_DISK 2u
#define NF_NET_UNAVAILABLE 4u
#define NF_UNSCHEDULABLE 8u
#define NF_MEM_PRESSURE 16u
#define NF_DISK_PRESSURE 32u
#define NF_LABEL_PRESENCE 64u  // CheckNodeLabelPresence (Policy-enabled predicate) fails on this node

// failure-reason bits (error.go:33-54, :84-86)
enum {
  RB_PODS = 0, RB_CPU, RB_MEM, RB_AGPU, RB_EPH, RB_SCALAR0, RB_SCALAR1, RB_SCALAR2, RB_SCALAR3,
  RB_HOSTNAME, RB_PORTS, RB_SELECTOR, RB_TAINTS, RB_MEM_PRESSURE, RB_DISK_PRESSURE, RB_NOT_READY,
  RB_OUT_OF_DISK, RB_NET_UNAVAILABLE, RB_UNSCHEDULABLE, RB_IPA, RB_IPA_EXISTING_ANTI, RB_IPA_AFFINITY,
  RB_IPA_ANTI, RB_EXT0, RB_EXT1, RB_EXT2, RB_EXT3, RB_EXT_LIMITS, RB_LABEL_PRESENCE, RB_EXTENDER
};

// pod flags
#define PF_HAS_REQUEST 1u        // some request is non-zero (predicates.go:648-654 shortcut not taken)
#define PF_BEST_EFFORT 2u
#define PF_NODE_SELECTOR 4u      // expr 0 is the nodeSelector ALL-expression
#define PF_HAS_REQUIRED 8u       // required node affinity present
#define PF_HAS_PORTS 16u
#define PF_EXT_LIMITS_ERROR 32u  // PodExtendedResourceName failed for entry `ext_err_entry`
#define PF_PRIO_ERROR 64u        // a priority function would return an error (node_affinity.go:61-64)
#define PF_SCHED_ERROR 128u      // a predicate would return an error for every node
#define PF_NO_VOLUMES_OK 256u

// expression kinds
#define EK_ANY 0u   // (bits & mask) != 0
#define EK_NONE 1u  // (bits & mask) == 0
#define EK_ALL 2u   // (bits & mask) == mask

// inter-pod affinity check modes (predicates.go:1238-1346)
#define CM_FORBID_EXISTING 0u  // existing pod's required anti-affinity matches the incoming pod
#define CM_REQUIRE 1u          // own required affinity term
#define CM_FORBID_OWN 2u       // own required anti-affinity term
#define CM_FAIL_ALWAYS 3u      // own anti-affinity term with empty topologyKey

struct KsbNodeTable {
  int32_t n;         // rows held by this shard
  int32_t n_pad;     // column stride
  int32_t row_base;  // global row of local row 0
  int32_t n_total;   // rows over all shards
  int32_t w_mr, w_rl; // MostRequestedPriority / ResourceLimitsPriority weights (0: not enabled by the Policy)
  int32_t w_il;       // ImageLocalityPriority weight (0: not enabled); its per-node scores are columns of `spread`
  int32_t has_static; // a Policy configured NodeLabelPriority instances: static_score holds their per-node sum
  int32_t has_presence; // ... CheckNodeLabelPresence labels: NF_LABEL_PRESENCE marks the nodes that fail it
  // One-pod calls with scheduler-extender verdicts (core/generic_scheduler.go:333-351, :608-632), else null: the nodes
  // the extenders' Filter calls kept (one bit per row) and the sum over extenders of weight * Prioritize score per row
  const uint32_t* ext_keep;
  const int32_t* ext_score;
  int32_t W;         // label words
  int32_t tw;        // 64-bit words of each taint / preferAvoid set (columns taint_hard, taint_soft, avoid hold tw words each,
                     // word w at + w * n_pad; a pod's masks for the words above 0 trail its pool block)
  int32_t n_zones;   // interned zone keys (+1: id 0 = no zone)
  // weights (defaults.go:217-257)
  int32_t w_lr, w_bra, w_ss, w_na, w_tt, w_ipa, w_avoid;
  // int64 columns
  int64_t *alloc_cpu, *alloc_mem, *alloc_agpu, *alloc_eph;
  int64_t *req_cpu, *req_mem, *req_agpu, *req_eph;
  int64_t *nz_cpu, *nz_mem;
  int64_t *alloc_scalar;  // [KSB_MAX_SCALARS][n_pad]
  int64_t *req_scalar;    // [KSB_MAX_SCALARS][n_pad]
  int32_t *alloc_pods, *num_pods;
  uint32_t *flags;
  uint32_t *zone;         // zone id, 0 = none (pkg/util/node/node.go:116-133)
  uint64_t *labels;       // [W][n_pad]
  uint64_t *taint_hard;   // NoSchedule / NoExecute taints
  uint64_t *taint_soft;   // PreferNoSchedule taints
  uint64_t *avoid;        // preferAvoidPods controller refs
  uint64_t *port_any, *port_wild, *port_exact;  // [KSB_PORT_WORDS][n_pad]
  uint32_t *dev_avail;    // [KSB_MAX_RES][n_pad]
  int32_t *static_score;  // [n_pad] weighted sum of the node's NodeLabelPriority scores (the u32 column after dev_avail)
  uint16_t *dev_class;    // [KSB_MAX_RES][n_pad][KSB_MAX_DEV]
  uint32_t *topo;         // [KSB_MAX_TOPO][n_pad] domain ids, 0 = label absent
  uint32_t *spread;       // [spread classes][n_pad] matching-pod counts (selector_spreading.go:85-108)
  int32_t *tables;        // arena of per-domain count tables for inter-pod affinity
};

struct KsbExtEntry {
  int32_t num;
  uint8_t rslot;  // 255: resource name unknown to every node
  uint8_t pad;
  uint16_t cls_id;  // interned (resource slot, device-class mask); 0 = not interned. Equal ids = equal masks.
};

struct __attribute__((aligned(16))) KsbPodHeader {
  int64_t req_cpu, req_mem, req_agpu, req_eph;  // GetResourceRequest (predicates.go:583-621)
  int64_t nz_cpu, nz_mem;                       // getNonZeroRequests; also NodeInfo.AddPod's nonzero delta
  int64_t add_cpu, add_mem, add_agpu, add_eph;  // calculateResource (node_info.go:421-433)
  int64_t scalar_req[KSB_MAX_SCALARS];
  int64_t scalar_add[KSB_MAX_SCALARS];
  uint64_t hard_untol, soft_untol, avoid_mask;
  int64_t lim_cpu, lim_mem;  // the pod's cpu / memory limits (resource_limits.go:96-126), for ResourceLimitsPriority
  uint32_t flags;
  int32_t pinned_row;  // PodFitsHost: -1 no pin, -2 pinned to an unknown node, else global row
  int32_t spread_col;  // -1: pod has no selectors
  uint32_t pool_off;   // in u64 words
  int32_t pref_weight[KSB_MAX_TERMS];
  uint32_t expr_kind[2];  // 2 bits per expression
  uint8_t term_end[KSB_MAX_TERMS];
  uint8_t pref_end[KSB_MAX_TERMS];
  KsbExtEntry ext[KSB_MAX_EXT];
  uint8_t scalar_mask;  // bit s: scalar slot s is in podRequest.ScalarResources
  uint8_t n_exprs, n_req_terms, n_pref, n_ext, n_checks, n_score, n_commit, n_spread_inc;
  uint8_t ext_err_entry;
  int16_t img_col;  // ImageLocalityPriority: column of `spread` holding this image set's score per node, -1 none
  uint8_t pad[4];
  uint8_t expr_wm[KSB_MAX_EXPRS];  // per expression: which of the first four label words its mask touches (bit w = word w)
};

// pool words, in order:
//   exprs        n_exprs * W
//   ports        6 * KSB_PORT_WORDS            (if PF_HAS_PORTS): chk_any, chk_wild, chk_exact, add_any, add_wild, add_exact
//   ext classes  n_ext * KSB_CLASS_WORDS
//   checks       n_checks * 1                   table_off | topo << 32 | mode << 40 | self_match << 48
//   score terms  n_score * 2                    table_off | topo << 32 ; coef (int64)
//   commits      n_commit * 2                   table_off | topo << 32 | bump_total << 40 ; delta (int64)
//   spread inc   n_spread_inc * 1               column index

struct KsbResult {
  int32_t status;
  int32_t row;  // global row
  int32_t num_filtered;
  int32_t fiterr_slot;  // index into the reason-histogram buffer, -1 if none
  int64_t score;
  uint32_t device_mask[KSB_MAX_EXT];
};

struct KsbNorms {
  int32_t fit_count;
  int32_t max_na;
  int32_t max_tt;
  int32_t max_sc;
  int32_t ipa_max;
  int32_t ipa_min;
  int32_t have_zones;
  int32_t one_row;  // max fitting global row (the fitting row when fit_count == 1)
};
