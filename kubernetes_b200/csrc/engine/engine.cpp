// Host side of the engine (see engine.hpp). Reference paths are under
// /root/reference/plugin/pkg/scheduler unless noted.
#include "engine.hpp"

#include <algorithm>
#include <chrono>
#include <cstring>
#include <stdexcept>

namespace ksb {

thread_local bool Engine::tl_readonly = false;

#define CUDA_OK(call)                                                                          \
  do {                                                                                         \
    cudaError_t e_ = (call);                                                                   \
    if (e_ != cudaSuccess)                                                                     \
      throw std::runtime_error(std::string("CUDA error: ") + cudaGetErrorString(e_) + " at " + \
                               __FILE__ + ":" + std::to_string(__LINE__));                     \
  } while (0)

void DevBuf::ensure(size_t bytes) {
  if (bytes <= cap) return;
  if (d) cudaFree(d);
  size_t want = std::max(bytes, cap * 2);
  want = (want + 255) & ~(size_t)255;
  cudaError_t e = cudaMalloc(&d, want);
  if (e != cudaSuccess) {
    d = nullptr;
    cap = 0;
    throw std::runtime_error(std::string("cudaMalloc failed: ") + cudaGetErrorString(e));
  }
  cap = want;
}
void DevBuf::release() {
  if (d) cudaFree(d);
  d = nullptr;
  cap = 0;
}

void PinnedBuf::ensure(size_t bytes) {
  if (bytes <= cap) return;
  if (p) cudaFreeHost(p);
  size_t want = std::max(bytes, cap * 2);
  cudaError_t e = cudaMallocHost(&p, want);
  if (e != cudaSuccess) {
    p = nullptr;
    cap = 0;
    throw std::runtime_error(std::string("cudaMallocHost failed: ") + cudaGetErrorString(e));
  }
  cap = want;
}
void PinnedBuf::release() {
  if (p) cudaFreeHost(p);
  p = nullptr;
  cap = 0;
}

static std::string join3(const std::string& a, const std::string& b, const std::string& c) {
  std::string s = a;
  s.push_back('\0');
  s += b;
  s.push_back('\0');
  s += c;
  return s;
}

static bool is_scalar_resource(const std::string& n) {  // pkg/apis/core/v1/helper/helpers.go:32-84
  bool default_ns = n.find('/') == std::string::npos || n.find("kubernetes.io/") != std::string::npos;
  return !default_ns || n.rfind("hugepages-", 0) == 0;
}

Engine::Engine(const ksb_config& c) : cfg(c) {
  W = cfg.label_words > 0 ? cfg.label_words : 4;
  if (W > 16) throw std::runtime_error("label_words must be <= 16");
  int count = 0;
  cudaError_t e = cudaGetDeviceCount(&count);
  if (e != cudaSuccess || count == 0)
    throw std::runtime_error("no usable CUDA device: the engine has no CPU fallback");
  CUDA_OK(cudaSetDevice(cfg.device));
  CUDA_OK(cudaStreamCreateWithFlags(&stream, cudaStreamNonBlocking));
  CUDA_OK(cudaEventCreate(&ev_start));
  CUDA_OK(cudaEventCreate(&ev_stop));
  topo_values.reserve(KSB_MAX_TOPO);
}

Engine::~Engine() {
  cudaSetDevice(cfg.device);
  for (void* m : ipc_mappings) cudaIpcCloseMemHandle(m);
  for (DevBuf* b : {&b_i64, &b_u32, &b_u64, &b_dev_class, &b_topo, &b_spread, &b_tables, &b_headers, &b_pool,
                    &b_results, &b_scratch, &b_snapshot, &b_one_header, &b_one_pool, &b_flush, &b_dep, &b_fiterr})
    b->release();
  pin_chunk[0].release();
  pin_chunk[1].release();
  pin_ctl.release();
  pin_out.release();
  pin_col.release();
  b_fill_rows.release();
  b_pre_pods.release();
  b_pre_rows.release();
  b_pre_out.release();
  b_ext_keep.release();
  b_ext_score.release();
  for (auto& kv : numcols) kv.second.buf.release();
  if (ev_start) cudaEventDestroy(ev_start);
  if (ev_stop) cudaEventDestroy(ev_stop);
  if (stream) cudaStreamDestroy(stream);
}

// ------------------------------------------------------------------ dictionaries

bool Engine::node_has_label_bit(const Node& nd, const LabelBitDef& d) const {
  auto it = nd.labels.find(d.key);
  if (it == nd.labels.end()) return false;
  switch (d.kind) {
    case 'P':
      return it->second == d.value;
    case 'K':
      return true;
    default: {  // 'G' / 'L' (selector.go:214-233)
      int64_t v;
      if (!sel::to_int64(it->second, &v)) return false;
      return d.kind == 'G' ? v > d.num : v < d.num;
    }
  }
}

int Engine::label_bit(char kind, const std::string& key, const std::string& value, int64_t num) {
  std::string k(1, kind);
  k += join3(key, value, kind == 'G' || kind == 'L' ? std::to_string(num) : std::string());
  auto it = label_bits.find(k);
  if (it != label_bits.end()) return it->second;
  will_mutate();
  int bit = (int)label_defs.size();
  if (bit >= W * 64) {
    if (cfg.shard_count > 1 && shard_connected)
      throw std::runtime_error("label universe exceeds label_words*64 bits (a sharded engine keeps the width it was connected with)");
    // Nothing outside the pod being encoded refers to a label bit by number (encoded pods are not kept across a
    // rebuild), so a full universe at the widest columns starts afresh instead of failing every later pod that
    // mentions a new pair: a long-running scheduler sees many selector values come and go. Only a single call that
    // itself needs more than 1,024 distinct pairs cannot be served (schedule_batch gives up after three rounds).
    if (W >= 16) label_reset_pending = true;
    else pending_W = std::min(16, W * 2);
    throw GrowLabels();
  }
  LabelBitDef d;
  d.kind = kind;
  d.key = key;
  d.value = value;
  d.num = num;
  label_defs.push_back(d);
  label_bits[k] = bit;
  if (!layout_dirty) {  // fill the new bit column from the retained node labels
    if (!label_index_valid) build_label_index();
    // on one GPU with the table already uploaded the device sets the bit itself (run_pending_fills)
    const bool dev_fill = cfg.shard_count <= 1 && !full_upload && b_u64.d != nullptr && dirty_rows.size() <= 64;
    if (dev_fill && (kind == 'G' || kind == 'L') && ensure_numcol(key)) {
      pending_fills.push_back(PendingFill{bit, {}});
      stale_host_bits.push_back(bit);
    } else {
      host_fill_bit(bit);
      if (dev_fill) {
        // the rows holding the pair / the key, as a list for the device (however long: a column re-sent from the host
        // mirror would first need every kernel-set bit of the mirror brought up to date)
        PendingFill f{bit, {}};
        if (kind == 'P') {
          std::string pk = key;
          pk.push_back('\0');
          pk += value;
          auto it2 = rows_by_pair.find(pk);
          if (it2 != rows_by_pair.end()) f.rows.assign(it2->second.begin(), it2->second.end());
        } else {
          auto it2 = rows_by_key.find(key);
          if (it2 != rows_by_key.end())
            for (auto& rv : it2->second)
              if (kind == 'K' || (rv.is_num && (kind == 'G' ? rv.num > num : rv.num < num))) f.rows.push_back(rv.row);
        }
        if (!f.rows.empty()) pending_fills.push_back(std::move(f));
      } else {
        dirty_label_words.insert(bit >> 6);
      }
    }
  }
  return bit;
}

// the host mirror's copy of one label bit, from the label index
void Engine::host_fill_bit(int bit) {
  const LabelBitDef& d = label_defs[bit];
  const uint64_t m = 1ull << (bit & 63);
  if (d.kind == 'P') {
    std::string pk = d.key;
    pk.push_back('\0');
    pk += d.value;
    auto it2 = rows_by_pair.find(pk);
    if (it2 != rows_by_pair.end())
      for (int r : it2->second) u64(bit >> 6, r) |= m;
  } else {
    auto it2 = rows_by_key.find(d.key);
    if (it2 != rows_by_key.end())
      for (auto& rv : it2->second) {
        bool on = true;
        if (d.kind != 'K') on = rv.is_num && (d.kind == 'G' ? rv.num > d.num : rv.num < d.num);  // 'G' / 'L' (selector.go:214-233)
        if (on) u64(bit >> 6, rv.row) |= m;
      }
  }
}

void Engine::settle_host_bits() {
  if (stale_host_bits.empty()) return;
  if (!label_index_valid) build_label_index();
  for (int b : stale_host_bits) host_fill_bit(b);
  stale_host_bits.clear();
}

void Engine::invalidate_numcols() {
  for (auto& kv : numcols) kv.second.valid = false;
}

// device column of the key's parsed numeric values; false when no row has a numeric value under the key
bool Engine::ensure_numcol(const std::string& key) {
  auto it = numcols.find(key);
  if (it != numcols.end() && it->second.valid) return true;
  if (!label_index_valid) build_label_index();
  auto rk = rows_by_key.find(key);
  if (rk == rows_by_key.end()) return false;
  std::vector<int64_t> num((size_t)n_pad, 0);
  std::vector<uint8_t> ok((size_t)n_pad, 0);
  bool any = false;
  for (auto& rv : rk->second)
    if (rv.is_num) {
      num[rv.row] = rv.num;
      ok[rv.row] = 1;
      any = true;
    }
  if (!any) return false;
  NumCol& c = numcols[key];
  c.buf.ensure((size_t)n_pad * 9);
  CUDA_OK(cudaMemcpyAsync(c.buf.d, num.data(), (size_t)n_pad * 8, cudaMemcpyHostToDevice, stream));
  CUDA_OK(cudaMemcpyAsync((char*)c.buf.d + (size_t)n_pad * 8, ok.data(), (size_t)n_pad, cudaMemcpyHostToDevice, stream));
  CUDA_OK(cudaStreamSynchronize(stream));  // the sources are locals
  stats.h2d_bytes += (size_t)n_pad * 9;
  c.valid = true;
  return true;
}

// the device's copy of the bits interned since the last upload (words re-sent as a whole already carry them)
void Engine::run_pending_fills(const std::set<int>& uploaded_words) {
  for (auto& f : pending_fills) {
    const int w = f.bit >> 6;
    if (uploaded_words.count(w)) continue;
    uint64_t* col = (uint64_t*)b_u64.d + (size_t)w * ln_pad;
    const uint64_t m = 1ull << (f.bit & 63);
    int rc;
    if (f.rows.empty()) {  // (a list is never empty: this is a threshold fill)
      const LabelBitDef& d = label_defs[f.bit];
      NumCol& c = numcols[d.key];
      rc = ksb_launch_label_threshold(col, (const int64_t*)c.buf.d, (const uint8_t*)c.buf.d + (size_t)n_pad * 8, n, m, d.kind == 'G' ? 1 : 0,
                                      d.num, stream);
    } else {
      b_fill_rows.ensure(std::max(f.rows.size(), (size_t)n_pad) * 4);
      pin_col.ensure(std::max(f.rows.size() * 4, (size_t)ln_pad * 8));
      memcpy(pin_col.p, f.rows.data(), f.rows.size() * 4);
      CUDA_OK(cudaMemcpyAsync(b_fill_rows.d, pin_col.p, f.rows.size() * 4, cudaMemcpyHostToDevice, stream));
      rc = ksb_launch_label_scatter(col, (const int32_t*)b_fill_rows.d, (int)f.rows.size(), m, stream);
      CUDA_OK(cudaStreamSynchronize(stream));  // one staging buffer, pageable source
      stats.h2d_bytes += f.rows.size() * 4;
    }
    if (rc != 0) throw std::runtime_error(std::string("label fill kernel failed: ") + cudaGetErrorString((cudaError_t)rc));
    stats.kernel_launches += 1;
  }
  pending_fills.clear();
}

void Engine::build_label_index() {
  rows_by_pair.clear();
  rows_by_key.clear();
  for (int r = 0; r < n; ++r)
    for (auto& kv : rows[r]->node.labels) {
      std::string pk = kv.first;
      pk.push_back('\0');
      pk += kv.second;
      rows_by_pair[pk].push_back(r);
      KeyRow kr;
      kr.row = r;
      kr.num = 0;
      kr.is_num = sel::to_int64(kv.second, &kr.num);
      rows_by_key[kv.first].push_back(kr);
    }
  label_index_valid = true;
}

int Engine::topo_key(const std::string& key) {
  auto it = topo_idx.find(key);
  if (it != topo_idx.end()) return it->second;
  will_mutate();
  int k = (int)topo_keys.size();
  if (k >= KSB_MAX_TOPO) throw std::runtime_error("more than KSB_MAX_TOPO distinct topology keys");
  topo_keys.push_back(key);
  topo_values.emplace_back();
  topo_idx[key] = k;
  if (!layout_dirty) {
    for (int r = 0; r < n; ++r) h_topo[(size_t)k * n_pad + r] = (uint32_t)topo_domain(k, rows[r]->node);
    dirty_topo_cols.insert(k);
  }
  return k;
}

int Engine::topo_domain(int k, const Node& nd) {  // priorities/util/topologies.go:51-71
  if (!nd.has_labels) return 0;
  auto it = nd.labels.find(topo_keys[k]);
  if (it == nd.labels.end()) return 0;
  auto& dict = topo_values[k];
  auto d = dict.find(it->second);
  if (d != dict.end()) return d->second;
  int id = (int)dict.size() + 1;
  dict[it->second] = id;
  // tables over this key grow by one domain
  for (auto& t : tables)
    if (t.topo == k) t.vals.resize(id + 1, 0);
  arena_dirty = true;
  return id;
}

int Engine::scalar_slot(const std::string& name) {
  auto it = scalar_idx.find(name);
  if (it != scalar_idx.end()) return it->second;
  will_mutate();
  int s = (int)scalar_names.size();
  if (s >= KSB_MAX_SCALARS) throw std::runtime_error("more than KSB_MAX_SCALARS distinct scalar resource names");
  scalar_names.push_back(name);
  scalar_idx[name] = s;
  return s;
}

int Engine::port_bit(const std::string& proto_port) {
  auto it = port_slot.find(proto_port);
  if (it != port_slot.end()) return it->second;
  will_mutate();
  int b = (int)port_slot.size();
  if (b >= KSB_PORT_WORDS * 64) throw std::runtime_error("more than 256 distinct protocol/hostPort pairs");
  port_slot[proto_port] = b;
  return b;
}
int Engine::port_exact_bit(const std::string& triple) {
  auto it = port_exact_slot.find(triple);
  if (it != port_exact_slot.end()) return it->second;
  will_mutate();
  int b = (int)port_exact_slot.size();
  if (b >= KSB_PORT_WORDS * 64) throw std::runtime_error("more than 256 distinct protocol/hostIP/hostPort triples");
  port_exact_slot[triple] = b;
  return b;
}

// ------------------------------------------------------------------ signatures and classes

SigInfo* Engine::sig_of(const Pod& p) {
  std::string k = p.ns;
  for (auto& kv : p.labels) {
    k.push_back('\0');
    k += kv.first;
    k.push_back('\1');
    k += kv.second;
  }
  auto it = sigs.find(k);
  if (it != sigs.end()) return it->second.get();
  will_mutate();
  auto s = std::make_unique<SigInfo>();
  s->ns = p.ns;
  s->labels = p.labels;
  SigInfo* raw = s.get();
  sigs.emplace(std::move(k), std::move(s));
  return raw;
}

bool Engine::sig_matches_sel(SigInfo* s, int c) {
  if ((int)s->sel_match.size() <= c) {
    will_mutate();
    s->sel_match.resize(selclasses.size(), -1);
  }
  if (s->sel_match[c] < 0) {
    will_mutate();
    const SelClass& sc = selclasses[c];
    s->sel_match[c] = sc.nss.count(s->ns) && sel::matches(sc.selector, s->labels);  // topologies.go:41-50
  }
  return s->sel_match[c] != 0;
}

bool Engine::sig_matches_spread(SigInfo* s, int c) {
  if ((int)s->spread_match.size() <= c) {
    will_mutate();
    s->spread_match.resize(spread_classes.size(), -1);
  }
  if (s->spread_match[c] < 0) {
    will_mutate();
    const SpreadClass& sc = spread_classes[c];
    bool m = false;
    if (!sc.image_class && sc.ns == s->ns)
      for (auto& selx : sc.selectors)
        if (sel::matches(selx, s->labels)) {
          m = true;
          break;
        }
    s->spread_match[c] = m;
  }
  return s->spread_match[c] != 0;
}

int Engine::selclass_of(const Pod& owner, const PodAffinityTerm& t, bool* ok) {
  sel::Selector sx;
  *ok = sel::from_label_selector(t.selector, &sx);
  if (!*ok) return -1;
  std::set<std::string> nss;  // topologies.go:28-37
  if (t.namespaces.empty()) nss.insert(owner.ns);
  else nss.insert(t.namespaces.begin(), t.namespaces.end());
  std::string key;
  for (auto& x : nss) {
    key += x;
    key.push_back('\x04');
  }
  key.push_back('\x05');
  key += sel::canonical(sx);
  auto it = selclass_idx.find(key);
  if (it != selclass_idx.end()) return it->second;
  will_mutate();
  int id = (int)selclasses.size();
  SelClass c;
  c.nss = std::move(nss);
  c.selector = std::move(sx);
  selclasses.push_back(std::move(c));
  selclass_idx[key] = id;
  return id;
}

int Engine::table_of(int kind, int selclass, int topo) {
  std::string key = std::to_string(kind) + "|" + std::to_string(selclass) + "|" + std::to_string(topo);
  auto it = table_idx.find(key);
  if (it != table_idx.end()) return it->second;
  will_mutate();
  int id = (int)tables.size();
  Table t;
  t.kind = kind;
  t.selclass = selclass;
  t.topo = topo;
  t.vals.assign(topo_values[topo].size() + 1, 0);
  if (kind == TK_MATCH && !layout_dirty) {
    // count the existing pods that match this class, per topology domain of their node
    for (auto& kv : sigs) {  // through the signatures, as for a new SelectorSpread class
      SigInfo* es = kv.second.get();
      if (es->placed.empty() || !sig_matches_sel(es, selclass)) continue;
      for (auto& pl : es->placed) {
        auto nit = nodes.find(pl.first);
        if (nit == nodes.end() || nit->second.row < 0) continue;
        uint32_t d = h_topo[(size_t)topo * n_pad + nit->second.row];
        if (d) t.vals[d] += pl.second.first;
        t.vals[0] += pl.second.first;
      }
    }
  }
  tables.push_back(std::move(t));
  table_idx[key] = id;
  if (kind == TK_MATCH) selclasses[selclass].match_tables.push_back(id);
  else if (kind == TK_HELD_ANTI) held_anti_tables.push_back(id);
  else held_w_tables.push_back(id);
  // A new table goes to the end of the arena (layout_arena numbers the tables in creation order), so only its own
  // counters need to reach the device: re-sending the whole arena for every first use of an affinity term cost a
  // Schedule call about a millisecond with hostname-keyed tables.
  if (!arena_dirty && !layout_dirty && cfg.shard_count <= 1) appended_tables.push_back(id);
  else arena_dirty = true;
  return id;
}

void Engine::layout_arena() {
  uint32_t off = 0;
  for (auto& t : tables) {
    t.off = off;
    off += (uint32_t)t.vals.size();
  }
  arena_words = off;
}

// priorities/metadata.go:69-96 over the in-memory listers (testing/fake_lister.go:80-178)
int Engine::spread_class_of(const Pod& p, SigInfo* s) {
  if (s->spread_class != -2) return s->spread_class;
  will_mutate();
  std::vector<sel::Selector> sels;
  for (int kind = 0; kind < 2; ++kind)
    for (auto& o : owners[kind])
      if (o.ns == p.ns && sel::set_matches(o.map_selector, p.labels)) sels.push_back(sel::from_set(o.map_selector));
  for (int kind = 2; kind < 4; ++kind) {
    std::vector<sel::Selector> tmp;
    bool ok = true;
    for (auto& o : owners[kind]) {
      if (o.ns != p.ns) continue;
      sel::Selector sx;
      if (!sel::from_label_selector(o.selector, &sx)) {
        ok = false;
        break;
      }
      if (sel::matches(sx, p.labels)) tmp.push_back(std::move(sx));
    }
    if (ok)
      for (auto& x : tmp) sels.push_back(std::move(x));
  }
  if (sels.empty()) {
    s->spread_class = -1;
    return -1;
  }
  std::vector<std::string> canon;
  for (auto& x : sels) canon.push_back(sel::canonical(x));
  std::sort(canon.begin(), canon.end());
  canon.erase(std::unique(canon.begin(), canon.end()), canon.end());
  std::string key = p.ns;
  for (auto& c : canon) {
    key.push_back('\x06');
    key += c;
  }
  auto it = spread_idx.find(key);
  if (it != spread_idx.end()) {
    s->spread_class = it->second;
    return it->second;
  }
  int id = (int)spread_classes.size();
  SpreadClass sc;
  sc.ns = p.ns;
  sc.selectors = std::move(sels);
  spread_classes.push_back(std::move(sc));
  spread_idx[key] = id;
  h_spread.emplace_back((size_t)n_pad, 0u);
  if (!layout_dirty) {  // selector_spreading.go:85-108 over every pod already on a node
    auto& col = h_spread.back();
    for (auto& kv : sigs) {  // every cached pod, through its signature
      SigInfo* es = kv.second.get();
      if (es->placed.empty() || !sig_matches_spread(es, id)) continue;
      for (auto& pl : es->placed) {
        if (pl.second.second <= 0) continue;
        auto nit = nodes.find(pl.first);
        if (nit == nodes.end() || nit->second.row < 0) continue;
        col[nit->second.row] += (uint32_t)pl.second.second;
      }
    }
    dirty_spread_cols.insert(id);
  }
  s->spread_class = id;
  return id;
}

// priorities/image_locality.go:32-82: total size of the listed images present on the node, bucketed into 0..10
uint32_t Engine::image_score(const HostNode& hn, const std::vector<std::string>& images) const {
  constexpr int64_t mb = 1024 * 1024, min_size = 23 * mb, max_size = 1000 * mb;  // balanced_resource_allocation.go:32-36
  int64_t total = 0;
  for (auto& want : images) {
    int64_t size = 0;
    bool have = false;
    for (auto& im : hn.node.images)  // totalImageSize :66-82: a later entry carrying the same name replaces an earlier one
      for (auto& name : im.names)
        if (name == want) {
          size = im.size_bytes;
          have = true;
        }
    if (have) total += size;
  }
  if (total == 0 || total < min_size) return 0;
  if (total >= max_size) return 10;
  return (uint32_t)(10 * (total - min_size) / (max_size - min_size) + 1);
}

void Engine::fill_image_scores(int row) {
  for (size_t c = 0; c < spread_classes.size(); ++c)
    if (spread_classes[c].image_class) h_spread[c][row] = image_score(*rows[row], spread_classes[c].images);
}

// The score column of a pod's container image list (one per distinct list, created on first use)
int Engine::image_class_of(const Pod& p) {
  if (p.images.empty()) return -1;
  std::string key = "\x07img";
  for (auto& im : p.images) {
    key.push_back('\x06');
    key += im;
  }
  auto it = spread_idx.find(key);
  if (it != spread_idx.end()) return it->second;
  will_mutate();
  int id = (int)spread_classes.size();
  if (id >= 32767) throw std::runtime_error("more than 32,767 SelectorSpread / ImageLocality classes");
  SpreadClass sc;
  sc.image_class = true;
  sc.images = p.images;
  spread_classes.push_back(std::move(sc));
  spread_idx[key] = id;
  h_spread.emplace_back((size_t)n_pad, 0u);
  if (!layout_dirty) {
    for (int r = 0; r < n; ++r) h_spread[id][r] = image_score(*rows[r], spread_classes[id].images);
    dirty_spread_cols.insert(id);
  }
  return id;
}

// ------------------------------------------------------------------ node / pod state

static void reslist_to_aggregates(const ResourceList& rl, int64_t* cpu, int64_t* mem, int64_t* agpu, int64_t* eph,
                                  int64_t* pods, std::map<std::string, int64_t>* scalars) {  // node_info.go:90-111
  for (auto& kv : rl) {
    if (kv.first == kResCPU) *cpu += kv.second;
    else if (kv.first == kResMemory) *mem += kv.second;
    else if (kv.first == kResAlphaGPU) *agpu += kv.second;
    else if (kv.first == kResPods) { if (pods) *pods += kv.second; }
    else if (kv.first == kResEphemeral) *eph += kv.second;
    else if (is_scalar_resource(kv.first)) (*scalars)[kv.first] += kv.second;
  }
}

void Engine::node_upsert(const Node& nd) {
  std::lock_guard<std::mutex> g(mu);
  have_snapshot = false;  // the replay entry points are only valid right after the batch
  HostNode& hn = nodes[nd.name];
  bool was_listed = hn.has_node;
  hn.node = nd;
  hn.has_node = true;
  // NodeInfo.SetNode (node_info.go:469-491): pressure conditions are sticky
  for (auto& c : nd.conditions) {
    if (c.type == "MemoryPressure") hn.mem_pressure = c.status;
    else if (c.type == "DiskPressure") hn.disk_pressure = c.status;
  }
  // fork: extendedResourceManager.SetNode (schedulercache/extended_resources.go:154-207)
  for (auto& kv : hn.ext) {
    DeviceState& ds = kv.second;
    auto nres = nd.extended.find(kv.first);
    if (nres == nd.extended.end()) {
      ds.available.clear();
      ds.domain_present = true;
      continue;
    }
    for (auto& a : ds.allocatable) {
      bool present = false;
      for (auto& d : nres->second)
        if (d.id == a.first) present = true;
      if (!present) ds.available.erase(a.first);
    }
  }
  for (auto& kv : nd.extended) {
    auto it = hn.ext.find(kv.first);
    if (it == hn.ext.end()) {
      DeviceState ds;
      for (auto& d : kv.second) {
        ds.allocatable[d.id] = d;
        ds.available.insert(d.id);
      }
      ds.domain_present = true;
      hn.ext[kv.first] = std::move(ds);
      continue;
    }
    for (auto& d : kv.second) {
      if (it->second.used.count(d.id)) continue;
      it->second.available.insert(d.id);
      it->second.allocatable[d.id] = d;
    }
  }
  if (!was_listed || layout_dirty) {
    layout_dirty = true;
  } else {
    // topology domains may be new; tables are resized inside topo_domain
    settle_host_bits();  // (with the index as it stood: this row is recomputed from its labels just below)
    fill_row(hn.row);
    dirty_rows.insert(hn.row);
    label_index_valid = false;
    invalidate_numcols();
  }
}

void Engine::node_remove(const std::string& name) {
  std::lock_guard<std::mutex> g(mu);
  have_snapshot = false;  // the replay entry points are only valid right after the batch
  auto it = nodes.find(name);
  if (it == nodes.end()) return;
  it->second.has_node = false;  // cache.go:373-387: the NodeInfo survives while it still holds pods
  if (it->second.pod_keys.empty()) nodes.erase(it);
  layout_dirty = true;
}

void Engine::owner_upsert(int kind, const SelectorOwner& o) {
  std::lock_guard<std::mutex> g(mu);
  have_snapshot = false;  // the replay entry points are only valid right after the batch
  if (kind < 0 || kind > 3) throw std::runtime_error("selector owner kind out of range");
  bool found = false;
  for (auto& x : owners[kind])
    if (x.ns == o.ns && x.name == o.name) {
      x = o;
      found = true;
    }
  if (!found) owners[kind].push_back(o);
  for (auto& kv : sigs) kv.second->spread_class = -2;
}

void Engine::owner_remove(int kind, const std::string& ns, const std::string& name) {
  std::lock_guard<std::mutex> g(mu);
  have_snapshot = false;  // the replay entry points are only valid right after the batch
  if (kind < 0 || kind > 3) return;
  auto& v = owners[kind];
  v.erase(std::remove_if(v.begin(), v.end(), [&](const SelectorOwner& x) { return x.ns == ns && x.name == name; }),
          v.end());
  for (auto& kv : sigs) kv.second->spread_class = -2;
}

static std::string port_string(const ContainerPort& p) {  // node_info.go:444-456
  std::string proto = p.protocol.empty() ? "TCP" : p.protocol;
  std::string ip = p.host_ip.empty() ? "0.0.0.0" : p.host_ip;
  return proto + "/" + ip + "/" + std::to_string(p.host_port);
}

// NodeInfo.AddPod (node_info.go:329-356) on the host-side record
void Engine::host_add_pod(const Pod& p, HostNode& hn, const std::string* key, SigInfo* sig) {
  int64_t dummy_pods = 0;
  for (auto& c : p.containers) {
    reslist_to_aggregates(c.requests, &hn.req_cpu, &hn.req_mem, &hn.req_agpu, &hn.req_eph, &dummy_pods, &hn.req_scalar);
    auto cpu = c.requests.find(kResCPU);
    auto mem = c.requests.find(kResMemory);
    hn.nz_cpu += cpu == c.requests.end() ? 100 : cpu->second;                    // non_zero.go:29-50
    hn.nz_mem += mem == c.requests.end() ? 200LL * 1024 * 1024 : mem->second;
    for (auto& port : c.ports)
      if (port.host_port != 0) hn.used_ports.insert(port_string(port));
  }
  {
    std::string own;
    if (!key) {
      own = p.key();
      key = &own;
    }
    hn.pod_keys.insert(*key);
    hn.pod_order.push_back(*key);
  }
  {
    auto& pl = (sig ? sig : sig_of(p))->placed[p.node_name];
    pl.first += 1;
    if (!p.deleting) pl.second += 1;
  }
  for (auto& e : p.extended) {  // extended_resources.go:86-110
    if (e.limits.size() != 1) continue;
    auto ds = hn.ext.find(e.limits.begin()->first);
    if (ds == hn.ext.end()) continue;
    for (auto& id : e.assigned) {
      if (!ds->second.available.count(id)) continue;
      ds->second.available.erase(id);
      ds->second.used.insert(id);
    }
  }
}

// NodeInfo.RemovePod (node_info.go:359-419), including the missing EphemeralStorage subtraction
void Engine::host_remove_pod(const Pod& p, HostNode& hn) {
  int64_t dummy_pods = 0, eph_ignored = 0;
  int64_t cpu = 0, mem = 0, agpu = 0;
  std::map<std::string, int64_t> sc;
  for (auto& c : p.containers) {
    reslist_to_aggregates(c.requests, &cpu, &mem, &agpu, &eph_ignored, &dummy_pods, &sc);
    auto ccpu = c.requests.find(kResCPU);
    auto cmem = c.requests.find(kResMemory);
    hn.nz_cpu -= ccpu == c.requests.end() ? 100 : ccpu->second;
    hn.nz_mem -= cmem == c.requests.end() ? 200LL * 1024 * 1024 : cmem->second;
    for (auto& port : c.ports)
      if (port.host_port != 0) hn.used_ports.erase(port_string(port));
  }
  hn.req_cpu -= cpu;
  hn.req_mem -= mem;
  hn.req_agpu -= agpu;
  for (auto& kv : sc) hn.req_scalar[kv.first] -= kv.second;
  hn.pod_keys.erase(p.key());
  for (size_t i = 0; i < hn.pod_order.size(); ++i)
    if (hn.pod_order[i] == p.key()) {
      hn.pod_order[i] = hn.pod_order.back();
      hn.pod_order.pop_back();
      break;
    }
  {
    SigInfo* s = sig_of(p);
    auto pit = s->placed.find(p.node_name);
    if (pit != s->placed.end()) {
      pit->second.first -= 1;
      if (!p.deleting) pit->second.second -= 1;
      if (pit->second.first <= 0) s->placed.erase(pit);
    }
  }
  for (auto& e : p.extended) {  // extended_resources.go:114-152
    if (e.limits.size() != 1) continue;
    auto ds = hn.ext.find(e.limits.begin()->first);
    if (ds == hn.ext.end()) continue;
    for (auto& id : e.assigned) {
      if (!ds->second.used.count(id)) continue;
      ds->second.used.erase(id);
      if (ds->second.allocatable.count(id)) ds->second.available.insert(id);
    }
  }
}

void reslist_to_aggregates_public(const ResourceList& rl, int64_t* cpu, int64_t* mem, int64_t* agpu, int64_t* eph, int64_t* pods,
                                  std::map<std::string, int64_t>* scalars) {
  reslist_to_aggregates(rl, cpu, mem, agpu, eph, pods, scalars);
}

// ------------------------------------------------------------------ layout

void Engine::fill_ports(int row) {
  HostNode& hn = *rows[row];
  for (int w = 0; w < KSB_PORT_WORDS; ++w) {
    u64(U64_PORT_ANY() + w, row) = 0;
    u64(U64_PORT_WILD() + w, row) = 0;
    u64(U64_PORT_EXACT() + w, row) = 0;
  }
  for (auto& s : hn.used_ports) {  // predicates/utils.go:161-202 turned into three bit tests
    size_t a = s.find('/');
    size_t b = s.find('/', a + 1);
    std::string proto = s.substr(0, a), ip = s.substr(a + 1, b - a - 1), port = s.substr(b + 1);
    int pb = port_bit(proto + "/" + port);
    int eb = port_exact_bit(s);
    u64(U64_PORT_ANY() + (pb >> 6), row) |= 1ull << (pb & 63);
    if (ip == "0.0.0.0") u64(U64_PORT_WILD() + (pb >> 6), row) |= 1ull << (pb & 63);
    u64(U64_PORT_EXACT() + (eb >> 6), row) |= 1ull << (eb & 63);
  }
}

void Engine::fill_devices(int row) {
  HostNode& hn = *rows[row];
  for (int rs = 0; rs < KSB_MAX_RES; ++rs) {
    u32(U32_DEV_AVAIL + rs, row) = 0;
    for (int d = 0; d < KSB_MAX_DEV; ++d) h_dev_class[((size_t)rs * n_pad + row) * KSB_MAX_DEV + d] = 0;
  }
  uint32_t& flags = u32(U32_FLAGS, row);
  for (auto& kv : hn.ext) {
    auto it = res_idx.find(kv.first);
    int rs;
    if (it == res_idx.end()) {
      rs = (int)res_names.size();
      if (rs >= KSB_MAX_RES) throw std::runtime_error("more than KSB_MAX_RES distinct extended resource names");
      res_names.push_back(kv.first);
      res_idx[kv.first] = rs;
    } else {
      rs = it->second;
    }
    if (kv.second.domain_present) flags |= 256u << rs;
    if (kv.second.allocatable.size() > KSB_MAX_DEV)
      throw std::runtime_error("node " + hn.node.name + " has more than KSB_MAX_DEV devices of " + kv.first);
    int bit = 0;
    for (auto& dev : kv.second.allocatable) {  // ascending id = bit order
      std::string ck;
      for (auto& a : dev.second.attributes) {
        ck += a.first;
        ck.push_back('\1');
        ck += a.second;
        ck.push_back('\2');
      }
      auto ci = class_idx.find(ck);
      int cls;
      if (ci == class_idx.end()) {
        cls = (int)class_attrs.size();
        if (cls >= KSB_CLASS_WORDS * 64) throw std::runtime_error("more than 256 distinct device attribute classes");
        class_attrs.push_back(dev.second.attributes);
        class_idx[ck] = cls;
      } else {
        cls = ci->second;
      }
      h_dev_class[((size_t)rs * n_pad + row) * KSB_MAX_DEV + bit] = (uint16_t)cls;
      if (kv.second.available.count(dev.first)) u32(U32_DEV_AVAIL + rs, row) |= 1u << bit;
      ++bit;
    }
  }
}

// Everything about one row that derives from its Node object and its aggregates (not the
// per-class counters, which follow the pods).
void Engine::policy_set(int kind, const std::string& key, bool presence, int32_t weight) {
  std::lock_guard<std::mutex> g(mu);
  have_snapshot = false;
  switch (kind) {
    case 0:
      if (key.empty()) throw std::runtime_error("NodeLabelPriority needs a label");
      pol_node_label.push_back(NodeLabelPrio{key, presence, weight});
      break;
    case 1:
      if (key.empty()) throw std::runtime_error("CheckNodeLabelPresence needs a label");
      pol_presence_labels.push_back(key);
      pol_presence = presence;
      break;
    case 2:
      w_most_requested = weight;
      break;
    case 3:
      w_resource_limits = weight;
      break;
    case 4:
      pol_node_label.clear();
      pol_presence_labels.clear();
      pol_presence = true;
      w_most_requested = w_resource_limits = w_image_locality = w_service_spreading = w_equal = 0;
      break;
    case 5:
      w_image_locality = weight;
      break;
    case 6:
      w_service_spreading = weight;
      break;
    case 7:
      w_equal = weight;
      break;
    default:
      throw std::runtime_error("unknown policy plugin kind");
  }
  layout_dirty = true;  // per-node facts are recomputed with the table
}

// Bit index of a new entry of a taint / preferAvoid set that already holds `have`. When the set outgrows its words, the
// word count grows (up to KSB_MAX_TAINT_WORDS) and the table is re-laid on the next sync: the row being filled keeps
// going (its bits beyond the current words are dropped here and set by the rebuild, which recomputes every row).
int Engine::set_bit_slot(int have, const char* what) {
  if (have >= 64 * KSB_MAX_TAINT_WORDS)
    throw std::runtime_error("more than " + std::to_string(64 * KSB_MAX_TAINT_WORDS) + " distinct " + what);
  if (have >= 64 * TW) {
    if (cfg.shard_count > 1 && shard_connected)
      throw std::runtime_error(std::string("sharded engine: more than ") + std::to_string(64 * TW) + " distinct " + what +
                               " after ksb_shard_connect (shared buffers keep their layout)");
    pending_TW = std::max(pending_TW, have / 64 + 1);
    layout_dirty = true;
  }
  return have;
}

void Engine::fill_row(int row) {
  HostNode& hn = *rows[row];
  const Node& nd = hn.node;
  int64_t cpu = 0, mem = 0, agpu = 0, eph = 0, npods = 0;
  std::map<std::string, int64_t> sc;
  reslist_to_aggregates(nd.allocatable, &cpu, &mem, &agpu, &eph, &npods, &sc);
  i64(I64_ALLOC_CPU, row) = cpu;
  i64(I64_ALLOC_MEM, row) = mem;
  i64(I64_ALLOC_AGPU, row) = agpu;
  i64(I64_ALLOC_EPH, row) = eph;
  i64(I64_REQ_CPU, row) = hn.req_cpu;
  i64(I64_REQ_MEM, row) = hn.req_mem;
  i64(I64_REQ_AGPU, row) = hn.req_agpu;
  i64(I64_REQ_EPH, row) = hn.req_eph;
  i64(I64_NZ_CPU, row) = hn.nz_cpu;
  i64(I64_NZ_MEM, row) = hn.nz_mem;
  for (int s = 0; s < KSB_MAX_SCALARS; ++s) {
    i64(I64_ALLOC_SCALAR + s, row) = 0;
    i64(I64_REQ_SCALAR + s, row) = 0;
  }
  for (auto& kv : sc) i64(I64_ALLOC_SCALAR + scalar_slot(kv.first), row) = kv.second;
  for (auto& kv : hn.req_scalar) i64(I64_REQ_SCALAR + scalar_slot(kv.first), row) = kv.second;
  u32(U32_ALLOC_PODS, row) = (uint32_t)(int32_t)npods;
  u32(U32_NUM_PODS, row) = (uint32_t)hn.pod_keys.size();
  uint32_t flags = 0;  // predicates.go:1382-1441
  for (auto& c : nd.conditions) {
    if (c.type == "Ready" && c.status != "True") flags |= NF_NOT_READY;
    else if (c.type == "OutOfDisk" && c.status != "False") flags |= NF_OUT_OF_DISK;
    else if (c.type == "NetworkUnavailable" && c.status != "False") flags |= NF_NET_UNAVAILABLE;
  }
  if (nd.unschedulable) flags |= NF_UNSCHEDULABLE;
  if (hn.mem_pressure == "True") flags |= NF_MEM_PRESSURE;
  if (hn.disk_pressure == "True") flags |= NF_DISK_PRESSURE;
  u32(U32_FLAGS, row) = flags;
  // zone key (pkg/util/node/node.go:116-133)
  uint32_t zone = 0;
  if (nd.has_labels) {
    auto r = nd.labels.find(kLabelRegion);
    auto z = nd.labels.find(kLabelZone);
    std::string region = r == nd.labels.end() ? "" : r->second;
    std::string zn = z == nd.labels.end() ? "" : z->second;
    if (!region.empty() || !zn.empty()) {
      std::string key = region + std::string(":\x00:", 3) + zn;
      auto zi = zone_idx.find(key);
      if (zi == zone_idx.end()) {
        zone = (uint32_t)zone_idx.size() + 1;
        zone_idx[key] = (int)zone;
      } else {
        zone = (uint32_t)zi->second;
      }
    }
  }
  u32(U32_ZONE, row) = zone;
  {  // Policy plugins that depend on the node alone (node_label.go:46-60, predicates.go:800-826)
    int32_t sscore = w_equal;  // EqualPriorityMap: 1 on every node
    for (auto& a : pol_node_label) {
      const bool exists = nd.has_labels && nd.labels.count(a.label) != 0;
      if ((exists && a.presence) || (!exists && !a.presence)) sscore += 10 * a.weight;
    }
    u32(U32_STATIC, row) = (uint32_t)sscore;
    fill_image_scores(row);
    for (auto& l : pol_presence_labels) {
      const bool exists = nd.has_labels && nd.labels.count(l) != 0;
      if ((exists && !pol_presence) || (!exists && pol_presence)) {
        u32(U32_FLAGS, row) |= NF_LABEL_PRESENCE;
        break;
      }
    }
  }
  for (int w = 0; w < W; ++w) u64(w, row) = 0;
  for (size_t b = 0; b < label_defs.size(); ++b)
    if (node_has_label_bit(nd, label_defs[b])) u64((int)(b >> 6), row) |= 1ull << (b & 63);
  for (int w = 0; w < 3 * TW; ++w) u64(U64_HARD() + w, row) = 0;
  for (auto& t : nd.taints) {
    std::string k = join3(t.key, t.value, t.effect);
    if (t.effect == "NoSchedule" || t.effect == "NoExecute") {
      auto it = hard_taint_idx.find(k);
      int b;
      if (it == hard_taint_idx.end()) {
        b = set_bit_slot((int)hard_taints.size(), "NoSchedule/NoExecute taints");
        hard_taints.push_back(t);
        hard_taint_idx[k] = b;
      } else {
        b = it->second;
      }
      if (b < 64 * TW) u64(U64_HARD(b >> 6), row) |= 1ull << (b & 63);
    } else if (t.effect == "PreferNoSchedule") {
      auto it = soft_taint_idx.find(k);
      int b;
      if (it == soft_taint_idx.end()) {
        b = set_bit_slot((int)soft_taints.size(), "PreferNoSchedule taints");
        soft_taints.push_back(t);
        soft_taint_idx[k] = b;
      } else {
        b = it->second;
      }
      if (b < 64 * TW) u64(U64_SOFT(b >> 6), row) |= 1ull << (b & 63);
    }
  }
  if (!nd.avoid_annotation_invalid)
    for (auto& a : nd.avoid) {
      std::string k = a.kind;
      k.push_back('\0');
      k += a.uid;
      auto it = avoid_idx.find(k);
      int b;
      if (it == avoid_idx.end()) {
        b = set_bit_slot((int)avoid_idx.size(), "preferAvoidPods controller references");
        avoid_idx[k] = b;
      } else {
        b = it->second;
      }
      if (b < 64 * TW) u64(U64_AVOID(b >> 6), row) |= 1ull << (b & 63);
    }
  fill_ports(row);
  fill_devices(row);
  for (int k = 0; k < (int)topo_keys.size(); ++k) h_topo[(size_t)k * n_pad + row] = (uint32_t)topo_domain(k, nd);
}

// Rebuild every column from the host records (node set changed).
void Engine::rebuild_layout() {
  have_snapshot = false;  // rows move: a table snapshot taken before no longer lines up
  rows.clear();
  for (auto& kv : nodes) {
    kv.second.row = -1;
    if (kv.second.has_node) {
      kv.second.row = (int)rows.size();
      rows.push_back(&kv.second);
    }
  }
  n = (int)rows.size();
  n_pad = std::max(256, (n + 255) & ~255);
  if (pending_W > W) W = pending_W;  // wider label columns (GrowLabels)
  if (label_reset_pending) {
    label_defs.clear();
    label_bits.clear();
    label_reset_pending = false;
  }
  // The taint and preferAvoid dictionaries follow the nodes that exist now: entries of nodes long gone do not pile up
  // (every row is recomputed below, and encoded pods do not outlive a rebuild).
  if (!(cfg.shard_count > 1 && shard_connected)) {  // (shared buffers keep the layout they were connected with)
    hard_taints.clear();
    soft_taints.clear();
    hard_taint_idx.clear();
    soft_taint_idx.clear();
    avoid_idx.clear();
    TW = 1;
    pending_TW = 1;
  }
  for (;;) {
    TW = std::max(TW, pending_TW);  // a taint / preferAvoid set outgrew its words: wider columns from here on
    h_i64.assign((size_t)I64_COLS * n_pad, 0);
    h_u32.assign((size_t)U32_COLS * n_pad, 0);
    h_u64.assign((size_t)u64_cols() * n_pad, 0);
    h_dev_class.assign((size_t)KSB_MAX_RES * n_pad * KSB_MAX_DEV, 0);
    h_topo.assign((size_t)KSB_MAX_TOPO * n_pad, 0);
    for (auto& c : h_spread) c.assign((size_t)n_pad, 0u);
    layout_dirty = false;
    for (int r = 0; r < n; ++r) fill_row(r);
    if (pending_TW <= TW) break;  // else a row met the 65th (129th) entry of a set: lay the table out again
  }
  layout_dirty = false;
  build_label_index();  // here, not inside the first Schedule that mentions a new label
  stale_host_bits.clear();  // every row was just recomputed from its labels
  pending_fills.clear();
  invalidate_numcols();
  // per-class counters follow the pods: recount them all
  // (a table first registered here is created with the existing pods already counted: zero AFTER registering, or
  // those pods are counted twice by the recount below)
  for (auto& kv : pods) register_pod_classes(kv.second.pod);
  for (auto& t : tables) {
    t.vals.resize(topo_values[t.topo].size() + 1, 0);
    std::fill(t.vals.begin(), t.vals.end(), 0);
  }
  for (auto& c : h_spread) c.assign((size_t)n_pad, 0u);  // the same for SelectorSpread columns created just now
  for (auto& kv : pods) {
    const Pod& p = kv.second.pod;
    auto nit = nodes.find(p.node_name);
    if (nit == nodes.end() || nit->second.row < 0) continue;
    std::vector<CommitEntry> commits;
    std::vector<int> inc;
    commit_lists(p, sig_of(p), &commits, &inc);
    apply_class_part_host(nit->second.row, commits, inc, +1);
  }
  for (int r = 0; r < n; ++r) fill_image_scores(r);  // (zeroed with the SelectorSpread counters above)
  dirty_rows.clear();
  dirty_label_words.clear();
  dirty_topo_cols.clear();
  dirty_spread_cols.clear();
  arena_dirty = true;
  full_upload = true;
}

}  // namespace ksb
