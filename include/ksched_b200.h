/*
 * ksched_b200 — C-ABI of the B200 scheduling engine.
 *
 * This is the drop-in boundary for ONE path of rotorliu/kubernetes: the body of
 * genericScheduler.Schedule (findNodesThatFit -> PrioritizeNodes -> selectHost) with the
 * DefaultProvider predicates/priorities and the fork's device-attribute matching.
 * A Go shim implementing algorithm.ScheduleAlgorithm binds these symbols through cgo
 * (INTEGRATION.md shows the stub). Paths cited below are relative to
 * /root/reference/plugin/pkg/scheduler.
 *
 * Conventions
 *  - Objects cross the boundary as flat little-endian byte records (include/ksb_wire.hpp
 *    documents the layout): one buffer per Node / Pod / selector owner. cgo may not pass Go
 *    memory that itself holds Go pointers, so byte buffers are the idiomatic form; the
 *    engine copies what it needs before returning and never retains caller memory.
 *  - Every call returns 0 on success, <0 on error (text via ksb_last_error).
 *  - ksb_schedule* are called from one thread at a time (the reference calls Schedule from a
 *    single goroutine, scheduler.go:179). Mutators may be called from other threads and never
 *    wait for a scheduling call: one that arrives while ksb_schedule* / ksb_flush works on the
 *    table is queued (the lock it takes is held for the time of an append, like the cache's own
 *    mutex, schedulercache/cache.go:83-97) and returns 0; queued mutators are applied in arrival
 *    order when that call ends and before the next one starts, which is the reference's snapshot
 *    point (core/generic_scheduler.go:122). A mutator that finds the engine free and nothing
 *    queued is applied on the spot and reports its own failure; the failure of a queued one is
 *    logged and kept for ksb_deferred_errors (the reference's informer handlers only log such
 *    errors, factory/factory.go:554-778).
 *  - There is no CPU fallback: if no CUDA device is usable ksb_create fails.
 */
#ifndef KSCHED_B200_H_
#define KSCHED_B200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct ksb_engine ksb_engine;

/* Priority weights and capacities. Replaces the priorityConfigs / hardPodAffinityWeight
 * arguments of core.NewGenericScheduler (core/generic_scheduler.go:1013-1035) as resolved by
 * factory.CreateFromKeys (factory/factory.go:897-958) from defaults.go:217-257. */
typedef struct ksb_config {
  int32_t device;                   /* CUDA device ordinal */
  int32_t w_least_requested;        /* LeastRequestedPriority        (default 1) */
  int32_t w_balanced;               /* BalancedResourceAllocation    (default 1) */
  int32_t w_selector_spread;        /* SelectorSpreadPriority        (default 1) */
  int32_t w_node_affinity;          /* NodeAffinityPriority          (default 1) */
  int32_t w_taint_toleration;       /* TaintTolerationPriority       (default 1) */
  int32_t w_inter_pod_affinity;     /* InterPodAffinityPriority      (default 1) */
  int32_t w_prefer_avoid;           /* NodePreferAvoidPodsPriority   (default 10000) */
  int32_t hard_pod_affinity_weight; /* v1.DefaultHardPodAffinitySymmetricWeight = 1 */
  int32_t label_words;              /* 64-bit words of the per-node label bitset (default 4) */
  int32_t shard_rank;               /* node-table shard held by this engine (0 when unsharded) */
  int32_t shard_count;              /* number of shards (1 when unsharded) */
  int32_t group_size;               /* pods evaluated per speculative group (0 = default) */
  int32_t pipeline;                 /* overlap a group's commit with the next group's evaluation: 0 = automatic
                                       (on from 16,384 node rows, any shard count), 1 = on, -1 = off */
} ksb_config;

/* Fill *cfg with the DefaultProvider values (algorithmprovider/defaults/defaults.go:217-257). */
void ksb_config_default(ksb_config* cfg);

/* core.NewGenericScheduler (core/generic_scheduler.go:1013-1035). */
ksb_engine* ksb_create(const ksb_config* cfg, char* err, size_t err_len);
void ksb_destroy(ksb_engine* e);
const char* ksb_last_error(const ksb_engine* e);

/* ---- state: the decorator of schedulercache.Cache (schedulercache/interface.go:61-121) ---- */

/* Cache.AddNode / UpdateNode (schedulercache/cache.go:347-371) -> NodeInfo.SetNode
 * (schedulercache/node_info.go:469-491). buf = Node record. */
int32_t ksb_node_upsert(ksb_engine* e, const uint8_t* buf, size_t len);
/* Cache.RemoveNode (schedulercache/cache.go:373-387). */
int32_t ksb_node_remove(ksb_engine* e, const char* name, size_t name_len);
/* Cache.AssumePod / AddPod (schedulercache/cache.go:118-137, :223-260) -> NodeInfo.AddPod
 * (node_info.go:329-356). buf = Pod record with NodeName (and Assigned device ids) set. */
int32_t ksb_pod_add(ksb_engine* e, const uint8_t* buf, size_t len);
/* Cache.ForgetPod / RemovePod (schedulercache/cache.go:170-197, :283-312) -> NodeInfo.RemovePod
 * (node_info.go:359-419). The pod is identified by namespace/name. */
int32_t ksb_pod_remove(ksb_engine* e, const char* ns, size_t ns_len, const char* name, size_t name_len);

/* ---- schedulercache.Cache as entry points of its own (schedulercache/interface.go:61-121) ----
 * The pod state machine of schedulercache/cache.go (Initial -> Assumed -> Added -> Deleted, Assumed -> Expired; :44-66,
 * :118-462) inside the library, for a caller that wants the informer handlers (factory/factory.go:554-778:
 * addPodToCache, updatePodInCache, deletePodFromCache) to be one call each instead of keeping the states itself above
 * ksb_pod_add / ksb_pod_remove. Each call decides from the states at the time of the call and returns the reference's
 * own error text on a bad transition ("pod ns/name state wasn't initial but get assumed", ...; a text starting with
 * "fatal:" is where the reference calls glog.Fatalf); what the transition does to the table goes through the mutator
 * queue. buf = Pod record with NodeName set. Time is the caller's clock in seconds (the reference injects it the same
 * way, cache.go:426-431). From the first ksb_cache_* call on, pods placed by ksb_schedule_batch enter as assumed.
 * Nodes use ksb_node_upsert / ksb_node_remove (AddNode / UpdateNode / RemoveNode have no state machine). */
void ksb_cache_set_ttl(ksb_engine* e, double seconds);                                          /* schedulercache.New(ttl, ..): 30 s */
int32_t ksb_cache_assume_pod(ksb_engine* e, const uint8_t* buf, size_t len);                    /* cache.go:118-137 */
int32_t ksb_cache_finish_binding(ksb_engine* e, const uint8_t* buf, size_t len, double now_seconds);  /* :139-160 */
int32_t ksb_cache_forget_pod(ksb_engine* e, const uint8_t* buf, size_t len);                    /* :162-190 */
int32_t ksb_cache_add_pod(ksb_engine* e, const uint8_t* buf, size_t len);                       /* :223-257 */
int32_t ksb_cache_update_pod(ksb_engine* e, const uint8_t* old_buf, size_t old_len, const uint8_t* new_buf, size_t new_len); /* :259-285 */
int32_t ksb_cache_remove_pod(ksb_engine* e, const uint8_t* buf, size_t len);                    /* :287-312 */
int32_t ksb_cache_is_assumed_pod(ksb_engine* e, const uint8_t* buf, size_t len, int32_t* assumed);  /* :314-328 */
int32_t ksb_cache_cleanup_assumed(ksb_engine* e, double now_seconds);                           /* cleanupAssumedPods, :431-462 */
/* configFactory.skipPodUpdate (factory/factory.go:306-353): 1 when the pod is assumed and the update differs from the
 * assumed pod in nothing but NodeName (ResourceVersion and Annotations do not cross this boundary at all). */
int32_t ksb_cache_skip_pod_update(ksb_engine* e, const uint8_t* buf, size_t len, int32_t* skip);

/* The listers SelectorSpread reads (algorithm/priorities/metadata.go:69-96; algorithm/types.go
 * ServiceLister/ControllerLister/ReplicaSetLister/StatefulSetLister). kind: 0 service, 1 RC,
 * 2 ReplicaSet, 3 StatefulSet. buf = SelectorOwner record. */
int32_t ksb_selector_owner_upsert(ksb_engine* e, int32_t kind, const uint8_t* buf, size_t len);
int32_t ksb_selector_owner_remove(ksb_engine* e, int32_t kind, const char* ns, size_t ns_len, const char* name,
                                  size_t name_len);

/* ---- plugins a Policy enables next to the DefaultProvider set (api/types.go:36-125; registered by name in
 * algorithmprovider/defaults/defaults.go:93-116) ----
 * Call before the first ksb_schedule* (any later call re-lays the table). kind:
 *   KSB_POLICY_NODE_LABEL_PRIORITY  adds one NodeLabelPriority instance (priorities/node_label.go:29-60): label = key,
 *                                   `presence`, `weight`
 *   KSB_POLICY_LABEL_PRESENCE       adds `key` to the labels of the CheckNodeLabelPresence predicate
 *                                   (predicates/predicates.go:778-826); `presence` is the predicate's one flag
 *   KSB_POLICY_MOST_REQUESTED       MostRequestedPriority (priorities/most_requested.go:32-88) with `weight`
 *   KSB_POLICY_RESOURCE_LIMITS      ResourceLimitsPriority (priorities/resource_limits.go:34-126) with `weight`
 *   KSB_POLICY_CLEAR                back to the DefaultProvider set
 *   KSB_POLICY_IMAGE_LOCALITY       ImageLocalityPriority (priorities/image_locality.go:32-82) with `weight`: reads the
 *                                   image lists at the end of the node and pod records (ksb_wire.hpp)
 *   KSB_POLICY_SERVICE_SPREADING    ServiceSpreadingPriority (defaults.go:90-102) with `weight`. Its Map reads the selectors of
 *                                   the priorityMetadata whenever one exists (priorities/selector_spreading.go:71-76), and
 *                                   PrioritizeNodes always has one, so it scores exactly as SelectorSpreadPriority does
 *   KSB_POLICY_EQUAL                EqualPriority (core/generic_scheduler.go:643-652) with `weight`: 1 on every node. A factory
 *                                   whose Policy lists no priority at all passes weight 1 (generic_scheduler.go:517-529)
 * The other Policy-only plugins (ServiceAntiAffinity, the ServiceAffinity predicate) are not
 * implemented: a factory that is asked for them must refuse, as gpusched.New does for unknown keys. */
enum {
  KSB_POLICY_NODE_LABEL_PRIORITY = 0,
  KSB_POLICY_LABEL_PRESENCE = 1,
  KSB_POLICY_MOST_REQUESTED = 2,
  KSB_POLICY_RESOURCE_LIMITS = 3,
  KSB_POLICY_CLEAR = 4,
  KSB_POLICY_IMAGE_LOCALITY = 5,
  KSB_POLICY_SERVICE_SPREADING = 6,
  KSB_POLICY_EQUAL = 7
};
int32_t ksb_policy_set(ksb_engine* e, int32_t kind, const char* key, size_t key_len, int32_t presence, int32_t weight);

/* ---- preemption: ScheduleAlgorithm.Preempt (algorithm/scheduler_interface.go:48-61; genericScheduler.Preempt,
 * core/generic_scheduler.go:199-256) ----
 * For a pod whose ksb_schedule ended in KSB_FIT_ERROR. The call evaluates the pod once more (nothing is assumed) to have
 * its FailedPredicateMap, keeps the nodes where preemption might help (nodesWherePreemptionMightHelp, :957-991), finds on
 * every one of them the victims whose removal lets the pod fit (selectVictimsOnNode, :883-953: all lower-priority pods
 * off, then reprieved from the highest priority down, equal priorities in the order go1.9's sort.Sort leaves them) and
 * picks the node as pickOneNodeForPreemption does (:654-758: a node without victims, else the lowest highest victim
 * priority, the lowest priority sum, the fewest victims, the first). Where the reference walks a Go map the rows are
 * walked in node-name order. Nothing is removed: the caller deletes the victims and reports them with ksb_pod_remove.
 * Not implemented (the call fails with a text that says which): a preemptor with pod (anti-)affinity terms or host
 * ports, cached pods with required anti-affinity terms, PodDisruptionBudgets, nominated pods, extenders, sharded engines.
 * status: 0 node picked (node_index, num_victims; victims by ksb_preempt_victim in decreasing priority),
 *         1 no node qualifies, 2 preemption cannot help on any node: clear the pod's nominated node name (:227-231). */
typedef struct {
  int32_t status;
  int32_t node_index;
  int32_t num_victims;
  int32_t num_potential;   /* nodes where preemption might help */
  int32_t num_candidates;  /* of those, nodes on which the pod fits once victims are gone */
} ksb_preempt_result;
int32_t ksb_preempt(ksb_engine* e, const uint8_t* pod, size_t len, ksb_preempt_result* out);
/* "namespace/name" of victim i of the most recent ksb_preempt; valid until the next call on this engine. */
int32_t ksb_preempt_victim(ksb_engine* e, int32_t i, const char** key, size_t* len);

/* ---- the hot path ---- */

enum {
  KSB_SCHEDULED = 0,        /* host chosen */
  KSB_NO_NODES = 1,         /* ErrNoNodesAvailable (core/generic_scheduler.go:58,117) */
  KSB_FIT_ERROR = 2,        /* *core.FitError (core/generic_scheduler.go:133-139) */
  KSB_SCHEDULE_ERROR = 3    /* any other error the reference returns from Schedule */
};

#define KSB_MAX_EXT_ENTRIES 4
#define KSB_MAX_DEVICES 32

/* Return value of ScheduleAlgorithm.Schedule (algorithm/scheduler_interface.go:48-61):
 * (selectedMachine, v1.ExtendedResourceBinding, error). */
typedef struct ksb_result {
  int32_t status;          /* KSB_* */
  int32_t node_index;      /* row of the chosen node in bytewise name order, -1 if none */
  int32_t num_filtered;    /* len(filteredNodes) */
  int32_t num_nodes;       /* len(nodes) */
  int64_t score;           /* total score of the chosen host; 0 on the one-fit shortcut */
  /* binding: for pod.Spec.ExtendedResources[i], a bitmask over the node's devices of that
   * resource in ascending device-id order (resolve to ids with ksb_device_id). */
  uint32_t device_mask[KSB_MAX_EXT_ENTRIES];
} ksb_result;

/* genericScheduler.Schedule (core/generic_scheduler.go:109-161). Does NOT assume the pod:
 * the caller follows with ksb_pod_add exactly as scheduler.go:478 calls cache.AssumePod. */
int32_t ksb_schedule(ksb_engine* e, const uint8_t* pod, size_t len, ksb_result* out);

/* n sequential iterations of scheduleOne's algorithm + assume steps (scheduler.go:430-497,
 * bind always succeeding, device ids echoed as storage.go:180-187 would): semantically
 * identical to n times {ksb_schedule; ksb_pod_add with NodeName and Assigned filled in}.
 * pods = n Pod records back to back, offsets[i]..offsets[i+1] delimiting record i. */
int32_t ksb_schedule_batch(ksb_engine* e, const uint8_t* pods, const uint64_t* offsets, int32_t n, ksb_result* out);

/* Name of node row `index` (valid until the next node mutation). */
int32_t ksb_node_name(const ksb_engine* e, int32_t index, const char** name, size_t* len);
/* Id of device `bit` of extended resource `resource` on node row `index`. */
int32_t ksb_device_id(const ksb_engine* e, int32_t index, const char* resource, size_t resource_len, int32_t bit,
                      const char** id, size_t* len);
/* FitError.Error() text (core/generic_scheduler.go:70-88) for the most recent KSB_FIT_ERROR of
 * ksb_schedule, or for pod `i` of the most recent ksb_schedule_batch. */
int32_t ksb_fit_error_message(ksb_engine* e, int32_t i, char* buf, size_t buf_len);

/* Bring the device table up to date with every node / pod / selector-owner call made so far, now, instead of
 * inside the next ksb_schedule. Counterpart of the work cache.UpdateNodeNameToInfoMap
 * (schedulercache/cache.go:79-96, called at core/generic_scheduler.go:124) does at the start of each
 * Schedule; calling it after the informers' initial sync keeps that cost out of the first pod's latency. */
int32_t ksb_flush(ksb_engine* e);
/* Failures of mutators that were queued behind a scheduling call (see "Conventions"): returns their number and
 * writes their texts, one per line, into buf; the list is cleared. */
int32_t ksb_deferred_errors(ksb_engine* e, char* buf, size_t buf_len);

/* Recovery (the reference rebuilds its cache from LIST+WATCH after a restart,
 * plugin/cmd/kube-scheduler/app/server.go:593-598, factory/factory.go:948 WaitForCacheSync; this is the same for the
 * device mirror after a CUDA error or whenever the caller's cache and the engine may have drifted apart).
 * ksb_resync_begin forgets every node, pod and selector owner and clears the device's sticky error state; the caller
 * then replays its cache (ksb_node_upsert / ksb_selector_owner_upsert / ksb_pod_add for every object it holds) and
 * calls ksb_resync_end, which rebuilds the device table from those records (same work as ksb_flush). The dictionaries
 * (label bits, classes) and lastNodeIndex survive: they do not depend on cluster state. Not to be called while a
 * ksb_schedule* call is running. */
int32_t ksb_resync_begin(ksb_engine* e);
int32_t ksb_resync_end(ksb_engine* e);

/* FailedPredicateMap of the most recent single-pod ksb_schedule (core/generic_scheduler.go:283-362, the map that
 * FitError carries and nodesWherePreemptionMightHelp :957-991 reads): masks[i] has bit b set when node row i failed
 * for reason b; 0 = the node fitted. `n` must be at least the number of node rows (ksb_result.num_nodes). */
int32_t ksb_failed_predicates(ksb_engine* e, uint32_t* masks, int32_t n);
/* HostPriorityList of the most recent single-pod ksb_schedule (what PrioritizeNodes returns,
 * core/generic_scheduler.go:509-640, api/types.go:204-227): scores[i] = weighted total of node row i, INT64_MIN for a
 * row that was filtered out. The reference only prioritizes when at least two nodes fit (:144-149); with fewer the
 * call fails. `n` must be at least the number of node rows. Lets a test compare the whole list, as the reference's
 * priority tests do (priorities/test_util.go:45-61), not only the winner. */
int32_t ksb_scores(ksb_engine* e, int64_t* scores, int32_t n);
/* Text of failure reason `bit` as predicates/error.go:33-54, :84-86 spells it ("Insufficient cpu",
 * "MatchNodeSelector", ...); for the extended-resource bits, the text for the pod of the last call. Empty if unused. */
int32_t ksb_reason_name(ksb_engine* e, int32_t bit, char* buf, size_t buf_len);

/* The NodeInfo aggregates the cache keeps for node `name` (schedulercache/node_info.go:35-78), as the JSON object
 * {"requested": {"cpu", "memory", "gpu", "ephemeral", "scalars": {name: n}}, "nonzero": {"cpu", "memory"}, "pods": n,
 * "usedPorts": ["proto/ip/port"], "allocatable": {"cpu", "memory", "pods", "scalars": {name: n}},
 * "available": {resource: [device ids]}}, or "null" when no NodeInfo exists under that
 * name (cache.go:283-293 drops it with its last pod when the Node object is gone). The device columns of the row are
 * read back and compared with the host records first; a difference is an error. *len receives the text length;
 * returns 1 if buf was too small. */
int32_t ksb_node_info(ksb_engine* e, const char* name, size_t name_len, char* buf, size_t buf_len, size_t* len);

/* genericScheduler.lastNodeIndex (core/generic_scheduler.go:100): selectHost's round-robin counter. */
uint64_t ksb_get_last_node_index(const ksb_engine* e);
void ksb_set_last_node_index(ksb_engine* e, uint64_t v);

/* Counters for measurement (utiltrace steps of core/generic_scheduler.go:110-154 have no
 * numeric counterpart; these are the engine's own). */
typedef struct ksb_stats {
  uint64_t pods_scheduled;     /* pods that went through the device path */
  uint64_t kernel_launches;    /* CUDA kernels launched by the engine */
  uint64_t groups;             /* speculative groups executed */
  uint64_t group_restarts;     /* pods re-evaluated because a speculation was invalidated */
  uint64_t h2d_bytes;          /* host->device bytes copied */
  uint64_t d2h_bytes;          /* device->host bytes copied */
  double last_batch_kernel_ms; /* CUDA-event time of the kernels of the last ksb_schedule_batch */
  double last_batch_encode_ms; /* host time spent encoding the last batch */
  double last_batch_sync_ms;   /* host time spent bringing the device table up to date before the launch */
  double last_batch_decode_ms; /* host time spent decoding the wire records of the last batch */
  double last_batch_post_ms;   /* host time spent mirroring the last batch's placements into the host records */
  int32_t num_nodes;
  int32_t node_row_bytes;      /* bytes of node-table columns one (pod,node) evaluation may read */
  /* SM cycles block 0 spent in the last launch (after ksb_replay_last_batch: the replayed launch): staging, phase 1,
   * barrier 1, phase 2, barrier 2, phase 3, barrier 3; then, only in a build with FINE_TIMING=1, inside phase 3:
   * segment-statistics prologue, changed rows, select, commit, wait for the writer warp; the last entry is unused. */
  uint64_t stage_cycles[13];
} ksb_stats;
int32_t ksb_get_stats(const ksb_engine* e, ksb_stats* out);

/* ---- scheduler extenders (algorithm.SchedulerExtender; core/extender.go, wire form api/types.go:127-202) ----
 * The extenders are HTTP services the Go side calls (HTTPExtender.Filter / Prioritize with ExtenderArgs ->
 * ExtenderFilterResult / HostPriorityList); the engine takes their verdict for one pod. Sequence in the shim, as in
 * findNodesThatFit / PrioritizeNodes (core/generic_scheduler.go:333-351, :608-632): ksb_schedule for the nodes that pass
 * the predicates (ksb_failed_predicates: rows whose mask has no bit outside the device-selection reasons), the
 * extenders' Filter calls on that list and Prioritize calls on the outcome, then ksb_schedule_extended with
 *   keep_rows  one bit per node row (bit r of word r/32): the rows the Filter calls kept; NULL = no Filter verdict.
 *              A row that is not kept fails with the reason "extender" (the shim puts the extender's own message into
 *              FailedPredicateMap) before the fork's device selection runs, as in the reference;
 *   row_scores sum over the extenders of weight * score per node row (combinedScores); NULL = none.
 * Same result record as ksb_schedule; nothing is assumed. */
int32_t ksb_schedule_extended(ksb_engine* e, const uint8_t* pod, size_t len, const uint32_t* keep_rows, const int32_t* row_scores,
                              ksb_result* out);

/* Re-run the device part of the most recent ksb_schedule_batch `iters` times against a
 * restored copy of the node table it started from, with inputs already resident in HBM, and
 * return the average CUDA-event milliseconds per run (measurement only; state is restored). */
int32_t ksb_replay_last_batch(ksb_engine* e, int32_t iters, double* avg_ms, ksb_result* out);

/* ---- sharding the node table over the GPUs of one NVSwitch domain ----
 * No counterpart in the reference (its only parallelism is workqueue.Parallelize(16, ...) over nodes,
 * core/generic_scheduler.go:326,574). One engine per GPU is created with shard_rank / shard_count set; every
 * engine is fed every node, pod and selector owner. After the cluster is loaded each engine calls
 * ksb_shard_export (which fixes the capacities of the buffers peers will map), the callers exchange the
 * blobs (torch.distributed in bench.py), and each engine calls ksb_shard_connect with all of them. From then
 * on ksb_schedule_batch / ksb_replay_last_batch must be called on every rank with the same arguments; the
 * kernels synchronise over NVLink and every rank receives the same results. */
int32_t ksb_shard_export(ksb_engine* e, int32_t max_batch_pods, uint8_t* out, size_t cap, size_t* len);
int32_t ksb_shard_connect(ksb_engine* e, const uint8_t* blobs, const uint64_t* offsets, int32_t n_ranks);

/* No counterpart in the reference. One run of what ksb_replay_last_batch runs (the resident batch,
 * against the restored node table), for callers that must not allocate GPU memory while they work:
 * a process confined to a fixed memory budget (a MIG slice or an MPS client with a hard limit, where
 * a late cudaMalloc can fail). This entry point calls no cudaMalloc, cudaFree or other allocating
 * routine: the kernel's working memory is a fixed array inside the loaded library (about 2.4 MiB, held
 * by every GPU that loads the library until the process exits) and state is restored into buffers that
 * already exist.
 *  - It runs on the engine's own stream and is synchronous: the kernel writes the engine's result
 *    buffer, a device-to-host copy into `out` (one ksb_result per pod of the batch, caller-owned, pageable
 *    or pinned) follows, and the call returns after cudaStreamSynchronize. It cannot be captured into a
 *    CUDA graph and does not run on a caller's stream.
 *  - Valid only directly after the ksb_schedule_batch it repeats: any mutator, ksb_flush or layout
 *    rebuild in between makes it (and ksb_replay_last_batch) fail with "no batch to replay".
 *  - Limits: at most 65,536 padded node rows (node count rounded up to 256), 1,024 zone slots and
 *    1,024 thread blocks; a batch beyond them is rejected with an error naming the limit. Engines
 *    created with shard_count > 1 are refused.
 *  - The first call on a GPU looks up that GPU's address of the array (cudaGetSymbolAddress, with
 *    the engine's device current) and caches it.
 *  - Every call, from every engine in the process, shares the one array, and the array keeps what
 *    the last call left in it; the counters that must start at zero are cleared on each call. Two
 *    calls whose kernels overlap on one GPU would race on it: serialise them. */
int32_t ksb_replay_last_batch_noalloc(ksb_engine* e, ksb_result* out);

#ifdef __cplusplus
}
#endif
#endif /* KSCHED_B200_H_ */
