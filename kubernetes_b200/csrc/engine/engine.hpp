// Host side of the engine: dictionaries that turn strings into bit positions / ids, the host
// mirror of the structure-of-arrays node table, the per-pod requirement-block encoder, and the
// driver of the batch kernel. No scheduling decision is made here: fit, scores and host
// selection are computed only by the CUDA kernels in kernels.cu.
#pragma once
#include <cuda_runtime.h>

#include <array>
#include <map>
#include <memory>
#include <mutex>
#include <set>
#include <string>
#include <unordered_map>
#include <vector>

#include "../../../include/ksched_b200.h"
#include "device_types.h"
#include "kernels.h"
#include "model.hpp"
#include "selectors.hpp"

namespace ksb {

struct DevBuf {  // growable device buffer
  void* d = nullptr;
  size_t cap = 0;
  void ensure(size_t bytes);
  void release();
};

struct PinnedBuf {  // growable page-locked host buffer (staging for asynchronous copies)
  void* p = nullptr;
  size_t cap = 0;
  void ensure(size_t bytes);
  void release();
};

struct DeviceState {  // fork: ExtendedResourceCacheManagerImpl (schedulercache/extended_resources.go:39-207)
  std::map<std::string, ExtendedDevice> allocatable;  // ascending id = bit order
  std::set<std::string> available, used;
  bool domain_present = false;  // `available[rName]` exists
};

struct HostNode {
  Node node;
  bool has_node = false;
  int row = -1;
  std::set<std::string> pod_keys;
  std::vector<std::string> pod_order;  // NodeInfo.pods: appended by AddPod, RemovePod moves the last into the hole (node_info.go:329-419)
  std::set<std::string> used_ports;  // "proto/ip/port" (node_info.go:435-466)
  std::map<std::string, DeviceState> ext;
  int64_t req_cpu = 0, req_mem = 0, req_agpu = 0, req_eph = 0, nz_cpu = 0, nz_mem = 0;
  std::map<std::string, int64_t> req_scalar;
  std::string mem_pressure, disk_pressure;  // sticky ConditionStatus (node_info.go:478-488)
};

struct HostPod {
  Pod pod;  // node_name and extended[i].assigned filled in
};

void reslist_to_aggregates_public(const ResourceList& rl, int64_t* cpu, int64_t* mem, int64_t* agpu, int64_t* eph, int64_t* pods,
                                  std::map<std::string, int64_t>* scalars);  // node_info.go:90-111

struct PreemptOutcome {  // genericScheduler.Preempt's return values (core/generic_scheduler.go:199-256)
  int status = 0;  // 0 a node was picked, 1 no node qualifies, 2 preemption cannot help anywhere: clear the pod's nomination
  int row = -1;
  int num_potential = 0, num_candidates = 0;
  std::vector<std::string> victims;  // namespace/name, decreasing priority
};

struct SelClass {  // (namespaces, label selector) of a pod-affinity term
  std::set<std::string> nss;
  sel::Selector selector;
  std::vector<int> match_tables;  // MATCH tables over this class
};

enum TableKind { TK_MATCH = 0, TK_HELD_ANTI = 1, TK_HELD_W = 2 };

struct Table {
  int kind, selclass, topo;
  std::vector<int32_t> vals;  // [domains(topo) + 1]; slot 0 of MATCH tables = matching pods anywhere
  uint32_t off = 0;
};

struct SpreadClass {
  std::string ns;
  std::vector<sel::Selector> selectors;
  // An ImageLocalityPriority class shares the column store: its column holds the 0..10 score of this list of container
  // images on every node (priorities/image_locality.go:32-82), no pod ever matches it and no commit touches it.
  bool image_class = false;
  std::vector<std::string> images;
};

// Thrown by the dictionaries when a lookup would have to intern something new while the calling thread runs in
// read-only mode (the parallel part of the batch encoder): the pod is then encoded again on the main thread.
// A label bit beyond the current words was asked for: the table has to be re-laid with wider label columns before the
// pod can be encoded (caught by ksb_schedule* / ksb_pod_add, which grow the columns and start over).
struct GrowLabels : std::exception {
  const char* what() const noexcept override { return "label universe outgrew its words"; }
};

struct NeedsIntern {};

struct SigInfo {  // memo per distinct (namespace, labels) signature
  std::string ns;
  StrMap labels;
  std::vector<int8_t> spread_match;  // per spread class
  std::vector<int8_t> sel_match;     // per SelClass
  int spread_class = -2;             // -2 unknown, -1 none
  std::vector<int> spread_inc;       // spread classes the signature matches, valid for `spread_inc_classes` classes
  size_t spread_inc_classes = (size_t)-1;
  // where the cached pods with this signature are: node name -> {all, not being deleted}. Lets a new SelectorSpread
  // class or affinity table be counted from the signatures (hundreds) instead of from every cached pod.
  std::unordered_map<std::string, std::pair<int32_t, int32_t>> placed;
};

struct CommitEntry {  // one per-domain counter the assume step bumps
  int table;
  int32_t delta;
};

struct EncodedPod {
  KsbPodHeader h;
  std::vector<uint64_t> pool;
  std::vector<CommitEntry> commits;  // host-side form of the pool's commit list
  std::vector<int> spread_inc;
  std::vector<std::string> ext_reasons;  // failure text for RB_EXT0.. / RB_EXT_LIMITS
  SigInfo* sig = nullptr;  // the pod's (namespace, labels) signature record, so that the host mirror need not look it up again
};

class Engine {
 public:
  explicit Engine(const ksb_config& cfg);
  ~Engine();

  // mutators (schedulercache.Cache decorator)
  void node_upsert(const Node& n);
  void node_remove(const std::string& name);
  void pod_add(const Pod& p);
  void pod_remove(const std::string& ns, const std::string& name);
  void owner_upsert(int kind, const SelectorOwner& o);
  void owner_remove(int kind, const std::string& ns, const std::string& name);

  // hot path
  void schedule_batch(std::vector<Pod>& pods, ksb_result* out, bool assume);
  void flush();
  void resync_begin();
  void schedule_batch_once(std::vector<Pod>& batch, ksb_result* out, bool assume);
  // Plugins a Policy enables next to the DefaultProvider set (api/types.go:36-125): NodeLabelPriority instances and the
  // CheckNodeLabelPresence labels are per-node facts computed on the host (a score column, a node flag); the
  // MostRequestedPriority and ResourceLimitsPriority weights go to the kernel.
  struct NodeLabelPrio {
    std::string label;
    bool presence;
    int32_t weight;
  };
  std::vector<NodeLabelPrio> pol_node_label;
  std::vector<std::string> pol_presence_labels;
  bool pol_presence = true;
  int32_t w_most_requested = 0, w_resource_limits = 0, w_image_locality = 0;
  // ServiceSpreadingPriority (defaults.go:94-102) is SelectorSpread built over the service lister alone, but its Map takes
  // the selectors from the priorityMetadata whenever there is one (selector_spreading.go:71-76), and PrioritizeNodes always
  // has one (defaults.go:47-50): its weight simply adds to SelectorSpread's. EqualPriority (generic_scheduler.go:643-652)
  // is 1 on every node: its weight joins the per-node static score.
  int32_t w_service_spreading = 0, w_equal = 0;
  int image_class_of(const Pod& p);
  uint32_t image_score(const HostNode& hn, const std::vector<std::string>& images) const;
  void fill_image_scores(int row);
  void policy_set(int kind, const std::string& key, bool presence, int32_t weight);
  PreemptOutcome preempt(const Pod& p);
  // ksb_schedule with scheduler-extender verdicts: `keep` one bit per node row (null: no Filter verdict), `scores` the
  // extenders' weighted Prioritize sums per row (null: none)
  void schedule_extended(const Pod& p, const uint32_t* keep, const int32_t* scores, ksb_result* out);
  DevBuf b_ext_keep, b_ext_score;
  const uint32_t* call_ext_keep = nullptr;  // device pointers for the call in progress
  const int32_t* call_ext_score = nullptr;
  PreemptOutcome last_preempt;
  DevBuf b_pre_pods, b_pre_rows, b_pre_out;
  void warm_dictionaries();
  std::string node_info_json(const std::string& name);
  void grow_keep(DevBuf& b, size_t bytes);
  void failed_predicates(uint32_t* out, int32_t n_out);
  void scores(int64_t* out, int32_t n_out);
  std::string reason_name(int bit);
  void replay_last_batch(int iters, double* avg_ms, ksb_result* out);
  void replay_last_batch_noalloc(ksb_result* out);
  // sharding over the GPUs of one NVSwitch domain
  std::string shard_export(int max_batch_pods);
  void shard_connect(const std::vector<std::string>& blobs);
  std::string fit_error_message(int i);

  std::string last_error;
  ksb_config cfg;
  std::mutex mu;
  ksb_stats stats{};
  uint64_t last_node_index = 0;

  const std::string* node_name_at(int row) const;
  const std::string* device_id_at(int row, const std::string& resource, int bit) const;

 private:
  // ---- dictionaries
  int W;  // label words (64 bits each); grows by doubling, up to 16, when a selector mentions more distinct pairs
  int pending_W = 0;
  bool label_reset_pending = false;  // the universe is full at 16 words: the next rebuild starts the label dictionary afresh
  size_t batch_posted = 0;  // pods of the running batch already committed and mirrored (chunked batches)
  std::unordered_map<std::string, int> label_bits;  // "P\0k\0v" | "K\0k" | "G\0k\0n" | "L\0k\0n" -> bit
  struct LabelBitDef {
    char kind;
    std::string key, value;
    int64_t num = 0;
  };
  std::vector<LabelBitDef> label_defs;
  // inverted index over the retained node labels (rows by "key\0value" and by key): a label bit interned
  // later fills its column from the matching rows only instead of scanning every node
  std::map<std::array<uint64_t, KSB_CLASS_WORDS + 1>, uint16_t> ext_mask_ids;  // (device-class mask, slot) -> id
  // label -> rows index (built with the layout): rows holding a key=value pair; rows holding a key, with the value
  std::unordered_map<std::string, std::vector<int>> rows_by_pair;
  struct KeyRow {  // a row holding the key, with its value parsed once for the Gt / Lt bits (selector.go:214-233)
    int row;
    bool is_num;
    int64_t num;
  };
  std::unordered_map<std::string, std::vector<KeyRow>> rows_by_key;
  bool label_index_valid = false;
  void build_label_index();
  // Label bits interned after the table went up are set by a small kernel instead of a column upload: Gt / Lt bits from
  // a device column of the key's parsed values (created at ksb_flush for every key with a numeric value, or on first
  // use), bits matching few rows from a row list. For kernel-set Gt / Lt bits the host mirror is brought up to date
  // only when it is next needed as a whole (a column or full upload): `stale_host_bits`.
  struct NumCol {
    DevBuf buf;  // int64_t[n_pad] values, then uint8_t[n_pad] "parsed as an integer"
    bool valid = false;
  };
  std::map<std::string, NumCol> numcols;
  struct PendingFill {
    int bit;
    std::vector<int32_t> rows;  // scatter list; empty: threshold fill from numcols[label_defs[bit].key]
  };
  std::vector<PendingFill> pending_fills;
  std::vector<int> stale_host_bits;
  DevBuf b_fill_rows;
  void host_fill_bit(int bit);
  void settle_host_bits();
  bool ensure_numcol(const std::string& key);
  void run_pending_fills(const std::set<int>& uploaded_words);
  void invalidate_numcols();
  std::vector<Taint> hard_taints, soft_taints;
  std::unordered_map<std::string, int> hard_taint_idx, soft_taint_idx;
  std::unordered_map<std::string, int> avoid_idx;
  std::unordered_map<std::string, int> port_slot, port_exact_slot;
  std::unordered_map<std::string, int> scalar_idx;
  std::vector<std::string> scalar_names;
  std::unordered_map<std::string, int> res_idx;
  std::vector<std::string> res_names;
  std::unordered_map<std::string, int> class_idx;
  std::vector<StrMap> class_attrs;
  std::unordered_map<std::string, int> zone_idx;
  std::unordered_map<std::string, int> topo_idx;
  std::vector<std::string> topo_keys;
  std::vector<std::unordered_map<std::string, int>> topo_values;  // value -> domain id (1-based)

  static thread_local bool tl_readonly;  // this thread may only look the dictionaries up
  static void will_mutate() {
    if (tl_readonly) throw NeedsIntern();
  }
  int label_bit(char kind, const std::string& key, const std::string& value, int64_t num);
  bool node_has_label_bit(const Node& n, const LabelBitDef& d) const;
  int topo_key(const std::string& key);
  int topo_domain(int k, const Node& n);
  int scalar_slot(const std::string& name);
  int port_bit(const std::string& proto_port);
  int port_exact_bit(const std::string& triple);

  // ---- cluster state
  std::map<std::string, HostNode> nodes;  // by name: iteration order = row order
  std::vector<HostNode*> rows;
  std::unordered_map<std::string, HostPod> pods;
  std::vector<SelectorOwner> owners[4];
  bool layout_dirty = true;

  // ---- class registries
  std::vector<SelClass> selclasses;
  std::unordered_map<std::string, int> selclass_idx;
  std::vector<Table> tables;
  std::unordered_map<std::string, int> table_idx;
  std::vector<int> held_anti_tables, held_w_tables;
  bool arena_dirty = true;
  std::vector<int> appended_tables;  // tables created since the arena last went to the device, while it was otherwise up to date
  uint32_t arena_words = 0;
  std::vector<SpreadClass> spread_classes;
  std::unordered_map<std::string, int> spread_idx;
  std::unordered_map<std::string, std::unique_ptr<SigInfo>> sigs;
  SigInfo* sig_of(const Pod& p);
  bool sig_matches_sel(SigInfo* s, int selclass);
  bool sig_matches_spread(SigInfo* s, int cls);
  int selclass_of(const Pod& owner, const PodAffinityTerm& t, bool* ok);
  int table_of(int kind, int selclass, int topo);
  int spread_class_of(const Pod& p, SigInfo* s);
  void layout_arena();

  // ---- host mirror of the node table (same family layout as the device buffers)
  int n = 0, n_pad = 0;
  enum { I64_ALLOC_CPU = 0, I64_ALLOC_MEM, I64_ALLOC_AGPU, I64_ALLOC_EPH, I64_REQ_CPU, I64_REQ_MEM, I64_REQ_AGPU,
         I64_REQ_EPH, I64_NZ_CPU, I64_NZ_MEM, I64_ALLOC_SCALAR, I64_REQ_SCALAR = I64_ALLOC_SCALAR + KSB_MAX_SCALARS,
         I64_COLS = I64_REQ_SCALAR + KSB_MAX_SCALARS };
  enum { U32_ALLOC_PODS = 0, U32_NUM_PODS, U32_FLAGS, U32_ZONE, U32_DEV_AVAIL, U32_STATIC = U32_DEV_AVAIL + KSB_MAX_RES,
         U32_COLS = U32_STATIC + 1 };
  // taint / preferAvoid sets: TW 64-bit words each (1 until a 65th distinct entry is interned; growing re-lays the table)
  int TW = 1, pending_TW = 1;
  int u64_cols() const { return W + 3 * TW + 3 * KSB_PORT_WORDS; }
  int U64_HARD(int w = 0) const { return W + w; }
  int U64_SOFT(int w = 0) const { return W + TW + w; }
  int U64_AVOID(int w = 0) const { return W + 2 * TW + w; }
  int U64_PORT_ANY() const { return W + 3 * TW; }
  int U64_PORT_WILD() const { return W + 3 * TW + KSB_PORT_WORDS; }
  int U64_PORT_EXACT() const { return W + 3 * TW + 2 * KSB_PORT_WORDS; }
  int set_bit_slot(int have, const char* what);  // bit index for a new entry of such a set; grows TW when needed
  std::vector<int64_t> h_i64;
  std::vector<uint32_t> h_u32;
  std::vector<uint64_t> h_u64;
  std::vector<uint16_t> h_dev_class;  // [KSB_MAX_RES][n_pad][KSB_MAX_DEV]
  std::vector<uint32_t> h_topo;       // [KSB_MAX_TOPO][n_pad]
  std::vector<std::vector<uint32_t>> h_spread;
  int64_t& i64(int col, int row) { return h_i64[(size_t)col * n_pad + row]; }
  uint32_t& u32(int col, int row) { return h_u32[(size_t)col * n_pad + row]; }
  uint64_t& u64(int col, int row) { return h_u64[(size_t)col * n_pad + row]; }
  bool full_upload = true;
  std::set<int> dirty_rows;
  std::set<int> dirty_label_words, dirty_topo_cols, dirty_spread_cols;
  void rebuild_layout();
  void fill_row(int row);
  void fill_ports(int row);
  void fill_devices(int row);

  // ---- device
  cudaStream_t stream = nullptr;
  KsbNodeTable dT{};
  DevBuf b_i64, b_u32, b_u64, b_dev_class, b_topo, b_spread, b_tables, b_headers, b_pool, b_results, b_scratch,
      b_snapshot, b_one_header, b_one_pool, b_flush, b_dep, b_fiterr;
  size_t spread_cap_cols = 0;
  int grid = 0, threads = 0, max_grid = 0, group = 1, nseg = 0, zone_stride = 64;
  size_t smem_bytes = 0;
  bool pipe = false;  // the next launch overlaps commit and evaluation (prepare_geometry)
  std::vector<char> one_pod_buf;
  const char* one_pod_src = nullptr;  // one-pod call in flight: its staged header | dependency | pool block ...
  size_t one_pod_bytes = 0;           // ... and size (0: the pod block was uploaded to its own buffers)
  bool batch_pipe_ok = true;  // the last encoded batch has few pods that depend on a very recent pod
  size_t limits_for_smem = (size_t)-1;  // shared-memory size the cached max_grid / threads were queried for
  PinnedBuf pin_col;  // one column on its way up (first uses inside a Schedule call): a pageable source cost the first such copy 2 ms
  PinnedBuf pin_chunk[2], pin_ctl, pin_out;  // pod blocks in (two chunks in flight); control block in; control block + results out
  std::vector<int32_t> last_dep;              // dependency indices of the last batch, relative to the batch
  cudaEvent_t ev_start = nullptr, ev_stop = nullptr;
  // shard geometry: device rows [lbase, lbase + ln), device column stride ln_pad
  int lbase = 0, ln = 0, ln_pad = 0, rows_per_shard = 0;
  int eval_base = 0, eval_n = 0;  // rows this GPU evaluates: global rows [eval_base, eval_base + eval_n)
  size_t min_spread_cols = 0, min_table_bytes = (size_t)64 << 20;  // affinity tables: room up front, growing costs a Schedule call a millisecond
  bool shard_connected = false;
  struct PeerMap {
    void* p[9] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
    int ln = 0, ln_pad = 0, lbase = 0;
  };
  std::vector<PeerMap> peer_maps;    // rank 0: every other shard's buffers as mapped into this process
  size_t shard_results_cap = 0;
  int segs_per_shard = 1, nseg_total = 0, eval_nseg = 0;
  void prepare_geometry();
  void* peer0_scratch = nullptr;     // other ranks: rank 0's scratch buffer, mapped
  std::vector<void*> peer_scratch;   // every other rank's scratch buffer, mapped (barrier flags, Reduce inboxes)
  std::vector<void*> ipc_mappings;
  void plan_shard();
  void fill_table_pointers(KsbNodeTable* T, void* i64b, void* u32b, void* u64b, void* devb, void* topob, void* spreadb,
                           void* tablesb, size_t np) const;
  void upload_family(DevBuf& dst, const void* src, size_t elem, int cols);
  void upload_column(void* dev_base, const void* host_col, size_t elem, int dev_col);
  void upload_all();
  void upload_dirty();
  void sync_device();
  void setup_table_pointers();
  KsbBatchArgs make_args(int n_pods, bool commit);
  void launch(const KsbBatchArgs& a);
  std::vector<KsbNorms> launch_norms;  // host staging of launch(): per engine, the copies are issued under `mu`
  KsbPeer launch_self{};

  // ---- encoder
  void register_pod_classes(const Pod& p);
  void encode_pod(const Pod& p, EncodedPod* out);
  void compile_expr(const SelectorRequirement& r, bool* ok, std::vector<uint64_t>* mask, uint32_t* kind);
  void commit_lists(const Pod& p, SigInfo* s, std::vector<CommitEntry>* commits, std::vector<int>* spread_inc);
  // host mirror of commit_pod
  void apply_node_part_host(int row, const EncodedPod& e, const uint32_t* dev_masks);
  void apply_class_part_host(int row, const std::vector<CommitEntry>& commits, const std::vector<int>& spread_inc,
                             int sign);
  void host_add_pod(const Pod& p, HostNode& hn, const std::string* key = nullptr, SigInfo* sig = nullptr);
  void host_remove_pod(const Pod& p, HostNode& hn);

  // ---- last batch (for ksb_replay_last_batch and fit-error messages)
  std::vector<EncodedPod> last_encoded;
  std::vector<KsbResult> last_results;
  std::vector<int32_t> last_fiterr_hist;
  int last_n_nodes = 0;
  uint64_t last_index_before = 0;
  bool have_snapshot = false;
  void snapshot_dynamic();
  void restore_dynamic();
};

}  // namespace ksb
