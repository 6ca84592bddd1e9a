// Engine, part 2: the per-pod requirement-block encoder, host mirror of the assume step,
// device synchronisation and the batch driver. Reference paths are under
// /root/reference/plugin/pkg/scheduler unless noted.
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

static bool tolerates(const Toleration& t, const Taint& taint) {  // api/core/v1/toleration.go:37-56
  if (!t.effect.empty() && t.effect != taint.effect) return false;
  if (!t.key.empty() && t.key != taint.key) return false;
  if (t.op.empty() || t.op == "Equal") return t.value == taint.value;
  return t.op == "Exists";
}

static bool is_scalar_resource(const std::string& n) {
  bool default_ns = n.find('/') == std::string::npos || n.find("kubernetes.io/") != std::string::npos;
  return !default_ns || n.rfind("hugepages-", 0) == 0;
}

// ------------------------------------------------------------------ classes a pod touches

// Pass 1 over a batch: make sure every class / table this pod reads or will bump exists, so that
// the commit lists of EARLIER pods in the same batch already include them.
void Engine::register_pod_classes(const Pod& p) {
  SigInfo* s = sig_of(p);
  spread_class_of(p, s);
  if (!p.affinity.present) return;
  bool ok;
  if (p.affinity.pod.present) {
    for (auto& t : p.affinity.pod.required) {
      if (t.topology_key.empty()) continue;
      int c = selclass_of(p, t, &ok);
      if (!ok) continue;
      int k = topo_key(t.topology_key);
      table_of(TK_MATCH, c, k);                               // own check (predicates.go:1300-1331)
      if (cfg.hard_pod_affinity_weight > 0) table_of(TK_HELD_W, c, k);  // symmetry (interpod_affinity.go:184-196)
    }
    for (auto& wt : p.affinity.pod.preferred) {
      if (wt.term.topology_key.empty()) continue;
      int c = selclass_of(p, wt.term, &ok);
      if (!ok) continue;
      int k = topo_key(wt.term.topology_key);
      table_of(TK_MATCH, c, k);
      table_of(TK_HELD_W, c, k);
    }
  }
  if (p.affinity.anti.present) {
    const auto& req = p.affinity.anti.required;
    for (size_t i = 0; i < req.size(); ++i) {
      if (!req[i].topology_key.empty()) {
        int c = selclass_of(p, req[i], &ok);
        if (ok) table_of(TK_MATCH, c, topo_key(req[i].topology_key));  // own check (predicates.go:1334-1341)
      }
      // what OTHER pods will see: the reference records &term of a range variable, so every
      // matching term of this pod is checked with the LAST term's topology key (predicates.go:1166-1177)
      const std::string& eff = req.back().topology_key;
      if (eff.empty()) continue;
      int c = selclass_of(p, req[i], &ok);
      if (ok) table_of(TK_HELD_ANTI, c, topo_key(eff));
    }
    for (auto& wt : p.affinity.anti.preferred) {
      if (wt.term.topology_key.empty()) continue;
      int c = selclass_of(p, wt.term, &ok);
      if (!ok) continue;
      int k = topo_key(wt.term.topology_key);
      table_of(TK_MATCH, c, k);
      table_of(TK_HELD_W, c, k);
    }
  }
}

// What the assume step of this pod bumps besides its node's own columns.
void Engine::commit_lists(const Pod& p, SigInfo* s, std::vector<CommitEntry>* commits, std::vector<int>* spread_inc) {
  if (!p.deleting) {
    // the spread classes this signature matches, kept with the signature while the set of classes stands
    if (s->spread_inc_classes != spread_classes.size()) {
      will_mutate();
      s->spread_inc.clear();
      for (int c = 0; c < (int)spread_classes.size(); ++c)
        if (sig_matches_spread(s, c)) s->spread_inc.push_back(c);
      s->spread_inc_classes = spread_classes.size();
    }
    *spread_inc = s->spread_inc;
  }
  // the pod becomes an "existing pod": every MATCH table whose class it belongs to
  for (int c = 0; c < (int)selclasses.size(); ++c) {
    if (selclasses[c].match_tables.empty()) continue;
    if (!sig_matches_sel(s, c)) continue;
    for (int t : selclasses[c].match_tables) commits->push_back({t, 1});
  }
  if (!p.affinity.present) return;
  bool ok;
  auto held = [&](int kind, const PodAffinityTerm& t, const std::string& key, int32_t delta) {
    if (key.empty() || delta == 0) return;
    int c = selclass_of(p, t, &ok);
    if (!ok) return;
    commits->push_back({table_of(kind, c, topo_key(key)), delta});
  };
  if (p.affinity.pod.present) {
    if (cfg.hard_pod_affinity_weight > 0)
      for (auto& t : p.affinity.pod.required) held(TK_HELD_W, t, t.topology_key, cfg.hard_pod_affinity_weight);
    for (auto& wt : p.affinity.pod.preferred) held(TK_HELD_W, wt.term, wt.term.topology_key, wt.weight);
  }
  if (p.affinity.anti.present) {
    const auto& req = p.affinity.anti.required;
    for (auto& t : req) held(TK_HELD_ANTI, t, req.back().topology_key, 1);
    for (auto& wt : p.affinity.anti.preferred) held(TK_HELD_W, wt.term, wt.term.topology_key, -wt.weight);
  }
  // one entry per table: the device applies the list lane-parallel, so a table may appear only once
  std::sort(commits->begin(), commits->end(), [](const CommitEntry& a, const CommitEntry& b) { return a.table < b.table; });
  size_t w = 0;
  for (size_t i = 0; i < commits->size(); ++i) {
    if (w > 0 && (*commits)[w - 1].table == (*commits)[i].table) (*commits)[w - 1].delta += (*commits)[i].delta;
    else (*commits)[w++] = (*commits)[i];
  }
  commits->resize(w);
}

// ------------------------------------------------------------------ encoder

void Engine::compile_expr(const SelectorRequirement& r, bool* ok, std::vector<uint64_t>* mask, uint32_t* kind) {
  mask->assign(W, 0);
  *ok = false;
  sel::Op op;
  sel::Req req;
  if (!sel::op_from_node_string(r.op, &op) || !sel::make_req(r.key, op, r.values, &req)) return;
  auto set = [&](int bit) { (*mask)[bit >> 6] |= 1ull << (bit & 63); };
  switch (op) {
    case sel::OP_IN:
    case sel::OP_NOT_IN:
      for (auto& v : req.values) set(label_bit('P', r.key, v, 0));
      *kind = op == sel::OP_IN ? EK_ANY : EK_NONE;
      break;
    case sel::OP_EXISTS:
    case sel::OP_NOT_EXISTS:
      set(label_bit('K', r.key, "", 0));
      *kind = op == sel::OP_EXISTS ? EK_ANY : EK_NONE;
      break;
    case sel::OP_GT:
    case sel::OP_LT: {
      int64_t num = 0;
      sel::to_int64(req.values[0], &num);
      set(label_bit(op == sel::OP_GT ? 'G' : 'L', r.key, "", num));
      *kind = EK_ANY;
      break;
    }
  }
  *ok = true;
}

void Engine::encode_pod(const Pod& p, EncodedPod* out) {
  KsbPodHeader& h = out->h;
  memset(&h, 0, sizeof h);
  out->pool.clear();
  out->commits.clear();
  out->spread_inc.clear();
  out->ext_reasons.assign(KSB_MAX_EXT + 1, std::string());
  SigInfo* sig = sig_of(p);
  out->sig = sig;

  // --- resources: GetResourceRequest (predicates.go:583-621), calculateResource (node_info.go:421-433),
  //     getNonZeroRequests (priorities/metadata.go / non_zero.go:29-50)
  std::map<std::string, int64_t> req_sc, add_sc;
  int64_t dummy = 0;
  for (auto& c : p.containers) {
    for (auto& kv : c.requests) {
      if (kv.first == kResCPU) h.add_cpu += kv.second;
      else if (kv.first == kResMemory) h.add_mem += kv.second;
      else if (kv.first == kResAlphaGPU) h.add_agpu += kv.second;
      else if (kv.first == kResEphemeral) h.add_eph += kv.second;
      else if (kv.first == kResPods) dummy += kv.second;
      else if (is_scalar_resource(kv.first)) add_sc[kv.first] += kv.second;
    }
    auto cpu = c.requests.find(kResCPU);
    auto mem = c.requests.find(kResMemory);
    h.nz_cpu += cpu == c.requests.end() ? 100 : cpu->second;
    h.nz_mem += mem == c.requests.end() ? 200LL * 1024 * 1024 : mem->second;
  }
  h.req_cpu = h.add_cpu;
  h.req_mem = h.add_mem;
  h.req_agpu = h.add_agpu;
  h.req_eph = h.add_eph;
  req_sc = add_sc;
  for (auto& c : p.init_containers)
    for (auto& kv : c.requests) {
      if (kv.first == kResCPU) h.req_cpu = std::max(h.req_cpu, kv.second);
      else if (kv.first == kResMemory) h.req_mem = std::max(h.req_mem, kv.second);
      else if (kv.first == kResAlphaGPU) h.req_agpu = std::max(h.req_agpu, kv.second);
      else if (kv.first == kResEphemeral) h.req_eph = std::max(h.req_eph, kv.second);
      else if (is_scalar_resource(kv.first)) {
        auto it = req_sc.find(kv.first);
        if (kv.second > (it == req_sc.end() ? 0 : it->second)) req_sc[kv.first] = kv.second;
      }
    }
  for (auto& kv : req_sc) {
    int s = scalar_slot(kv.first);
    h.scalar_req[s] = kv.second;
    h.scalar_mask |= 1u << s;
  }
  for (auto& kv : add_sc) h.scalar_add[scalar_slot(kv.first)] = kv.second;
  if (!(h.req_cpu == 0 && h.req_mem == 0 && h.req_agpu == 0 && h.req_eph == 0 && req_sc.empty())) h.flags |= PF_HAS_REQUEST;
  // --- QoS BestEffort (pkg/apis/core/v1/helper/qos/qos.go:38-100)
  {
    bool best_effort = true;
    auto counted = [](const std::string& nme) { return nme == kResCPU || nme == kResMemory || nme.rfind("hugepages-", 0) == 0; };
    for (auto& c : p.containers) {
      for (auto& kv : c.requests)
        if (counted(kv.first) && kv.second > 0) best_effort = false;
      for (auto& kv : c.limits)
        if (counted(kv.first) && kv.second > 0) best_effort = false;
    }
    if (best_effort) h.flags |= PF_BEST_EFFORT;
  }
  // --- PodFitsHost
  h.pinned_row = -1;
  if (!p.node_name.empty()) {
    auto it = nodes.find(p.node_name);
    h.pinned_row = (it == nodes.end() || it->second.row < 0) ? -2 : it->second.row;
  }

  // --- label expressions: nodeSelector, required node affinity, preferred node affinity
  std::vector<uint64_t> exprs;
  int n_exprs = 0;
  auto push_expr = [&](const std::vector<uint64_t>& m, uint32_t kind) {
    if (n_exprs >= KSB_MAX_EXPRS) throw std::runtime_error("pod " + p.key() + " has more than KSB_MAX_EXPRS label expressions");
    exprs.insert(exprs.end(), m.begin(), m.end());
    h.expr_kind[n_exprs >> 4] |= kind << ((n_exprs & 15) * 2);
    for (int w = 0; w < 4 && w < (int)m.size(); ++w)
      if (m[w]) h.expr_wm[n_exprs] |= (uint8_t)(1u << w);
    ++n_exprs;
  };
  if (!p.node_selector.empty()) {  // predicates.go:708-713 with labels.SelectorFromSet's degrade-to-everything
    bool valid = true;
    for (auto& kv : p.node_selector)
      if (!sel::label_key_ok(kv.first) || !sel::label_value_ok(kv.second)) valid = false;
    if (valid) {
      std::vector<uint64_t> m(W, 0);
      for (auto& kv : p.node_selector) {
        int b = label_bit('P', kv.first, kv.second, 0);
        m[b >> 6] |= 1ull << (b & 63);
      }
      push_expr(m, EK_ALL);
      h.flags |= PF_NODE_SELECTOR;
    }
  }
  std::vector<uint64_t> m;
  uint32_t kind;
  bool ok;
  if (p.affinity.present && p.affinity.node.present && p.affinity.node.has_required) {
    h.flags |= PF_HAS_REQUIRED;
    for (auto& term : p.affinity.node.required_terms) {
      // a term that fails to convert makes nodeMatchesNodeSelectorTerms return false on the spot
      // (predicates.go:693-697): later terms are unreachable
      bool term_ok = true;
      std::vector<std::pair<std::vector<uint64_t>, uint32_t>> compiled;
      for (auto& e : term.match_expressions) {
        compile_expr(e, &ok, &m, &kind);
        if (!ok) {
          term_ok = false;
          break;
        }
        compiled.push_back({m, kind});
      }
      if (!term_ok) break;
      if (compiled.empty()) continue;  // empty MatchExpressions = labels.Nothing() (helpers.go:213-215)
      if (h.n_req_terms >= KSB_MAX_TERMS) throw std::runtime_error("pod " + p.key() + " has more than KSB_MAX_TERMS node selector terms");
      for (auto& c : compiled) push_expr(c.first, c.second);
      h.term_end[h.n_req_terms++] = (uint8_t)n_exprs;
    }
  }
  if (p.affinity.present && p.affinity.node.present) {
    for (auto& pt : p.affinity.node.preferred) {  // node_affinity.go:52-68
      if (pt.weight == 0) continue;
      bool term_ok = true;
      std::vector<std::pair<std::vector<uint64_t>, uint32_t>> compiled;
      for (auto& e : pt.preference.match_expressions) {
        compile_expr(e, &ok, &m, &kind);
        if (!ok) {
          term_ok = false;
          break;
        }
        compiled.push_back({m, kind});
      }
      if (!term_ok) {
        h.flags |= PF_PRIO_ERROR;  // the priority returns the conversion error: Schedule fails
        continue;
      }
      if (compiled.empty()) continue;
      if (h.n_pref >= KSB_MAX_TERMS) throw std::runtime_error("pod " + p.key() + " has more than KSB_MAX_TERMS preferred terms");
      for (auto& c : compiled) push_expr(c.first, c.second);
      h.pref_weight[h.n_pref] = pt.weight;
      h.pref_end[h.n_pref++] = (uint8_t)n_exprs;
    }
  }
  h.n_exprs = (uint8_t)n_exprs;

  // --- pod limits (resource_limits.go:96-126): sum over the containers, raised to any init container's
  {
    int64_t lc = 0, lm = 0;
    for (auto& c : p.containers)
      for (auto& kv : c.limits) {
        if (kv.first == kResCPU) lc += kv.second;
        else if (kv.first == kResMemory) lm += kv.second;
      }
    for (auto& c : p.init_containers)
      for (auto& kv : c.limits) {
        if (kv.first == kResCPU) lc = std::max(lc, kv.second);
        else if (kv.first == kResMemory) lm = std::max(lm, kv.second);
      }
    h.lim_cpu = lc;
    h.lim_mem = lm;
  }
  // --- taints (predicates.go:1349-1373, taint_toleration.go:29-53)
  // word 0 of each set travels in the header, the words above it (engines with more than 64 distinct entries) behind
  // the pool block: [TW-1] hard, [TW-1] soft, [TW-1] preferAvoid
  std::vector<uint64_t> tw_ext((size_t)3 * (TW - 1), 0);
  auto set_mask = [&](int set, size_t i) {
    if (i < 64) {
      (set == 0 ? h.hard_untol : set == 1 ? h.soft_untol : h.avoid_mask) |= 1ull << i;
    } else if ((int)(i >> 6) < TW) {
      tw_ext[(size_t)set * (TW - 1) + (i >> 6) - 1] |= 1ull << (i & 63);
    }
  };
  for (size_t i = 0; i < hard_taints.size(); ++i) {
    bool tol = false;
    for (auto& t : p.tolerations)
      if (tolerates(t, hard_taints[i])) tol = true;
    if (!tol) set_mask(0, i);
  }
  for (size_t i = 0; i < soft_taints.size(); ++i) {
    bool tol = false;
    for (auto& t : p.tolerations)
      if ((t.effect.empty() || t.effect == "PreferNoSchedule") && tolerates(t, soft_taints[i])) tol = true;
    if (!tol) set_mask(1, i);
  }
  // --- NodePreferAvoidPods (node_prefer_avoid_pods.go:35-57)
  if (p.controller.present && (p.controller.kind == "ReplicationController" || p.controller.kind == "ReplicaSet")) {
    std::string k = p.controller.kind;
    k.push_back('\0');
    k += p.controller.uid;
    auto it = avoid_idx.find(k);
    if (it != avoid_idx.end()) set_mask(2, (size_t)it->second);
  }
  // --- host ports (util/utils.go:33-62)
  std::vector<uint64_t> ports(6 * KSB_PORT_WORDS, 0);
  for (auto& c : p.containers)
    for (auto& port : c.ports) {
      if (port.host_port == 0) continue;
      std::string proto = port.protocol.empty() ? "TCP" : port.protocol;
      std::string ip = port.host_ip.empty() ? "0.0.0.0" : port.host_ip;
      int pb = port_bit(proto + "/" + std::to_string(port.host_port));
      int eb = port_exact_bit(proto + "/" + ip + "/" + std::to_string(port.host_port));
      bool wild = ip == "0.0.0.0";
      auto setw = [&](int fam, int bit) { ports[fam * KSB_PORT_WORDS + (bit >> 6)] |= 1ull << (bit & 63); };
      if (wild) setw(0, pb);  // conflicts with any use of proto/port
      setw(1, pb);            // conflicts with a wildcard-ip use of proto/port
      setw(2, eb);            // conflicts with the identical triple
      setw(3, pb);
      if (wild) setw(4, pb);
      setw(5, eb);
      h.flags |= PF_HAS_PORTS;
    }
  // --- fork: pod.Spec.ExtendedResources (core/extended_resources.go:83-150)
  std::vector<uint64_t> extcls;
  if (p.extended.size() > KSB_MAX_EXT) throw std::runtime_error("pod " + p.key() + " has more than KSB_MAX_EXT extended resource entries");
  h.n_ext = (uint8_t)p.extended.size();
  for (size_t i = 0; i < p.extended.size(); ++i) {
    const PodExtendedResource& e = p.extended[i];
    uint64_t cm[KSB_CLASS_WORDS] = {0, 0, 0, 0};
    h.ext[i].rslot = 255;
    h.ext[i].num = 0;
    if (e.limits.size() != 1) {
      if (!(h.flags & PF_EXT_LIMITS_ERROR)) {
        h.flags |= PF_EXT_LIMITS_ERROR;
        h.ext_err_entry = (uint8_t)i;
        out->ext_reasons[KSB_MAX_EXT] = "unexpected limits length: " + std::to_string(e.limits.size()) + " != 1";
      }
    } else {
      const std::string& rname = e.limits.begin()->first;
      out->ext_reasons[i] = "Insufficient " + rname;
      int64_t num = e.limits.begin()->second;
      h.ext[i].num = (int32_t)std::max<int64_t>(-1, std::min<int64_t>(num, 1 << 20));
      auto it = res_idx.find(rname);
      if (it != res_idx.end()) h.ext[i].rslot = (uint8_t)it->second;
      sel::Selector sx;
      bool conv = e.required.empty() ? true : sel::from_node_requirements(e.required, &sx);
      for (size_t c = 0; c < class_attrs.size(); ++c) {
        bool match = e.required.empty() ? true : (conv && sel::matches(sx, class_attrs[c]));  // isDeviceAMatch :152-167
        if (match) cm[c >> 6] |= 1ull << (c & 63);
      }
    }
    extcls.insert(extcls.end(), cm, cm + KSB_CLASS_WORDS);
    // equal (slot, mask) pairs get one id: the kernel keeps per-row match masks under it while a group commits
    std::array<uint64_t, KSB_CLASS_WORDS + 1> key;
    for (int w = 0; w < KSB_CLASS_WORDS; ++w) key[w] = cm[w];
    key[KSB_CLASS_WORDS] = h.ext[i].rslot;
    auto kid = ext_mask_ids.find(key);
    if (kid == ext_mask_ids.end() && ext_mask_ids.size() < 65535) {
      will_mutate();
      kid = ext_mask_ids.emplace(key, (uint16_t)(ext_mask_ids.size() + 1)).first;
    }
    h.ext[i].cls_id = kid == ext_mask_ids.end() ? 0 : kid->second;
    h.ext[i].pad = 0;
  }
  // --- SelectorSpread class (selector_spreading.go:64-113)
  h.spread_col = spread_class_of(p, sig);
  h.img_col = w_image_locality ? (int16_t)image_class_of(p) : (int16_t)-1;

  // --- inter-pod affinity: checks (predicates.go:1238-1346) and score terms (interpod_affinity.go:121-216)
  std::vector<uint64_t> checks, score;
  auto pack = [](uint32_t off, int topo, uint32_t mode, uint32_t self) {
    return (uint64_t)off | ((uint64_t)(uint32_t)topo << 32) | ((uint64_t)mode << 40) | ((uint64_t)self << 48);
  };
  for (int t : held_anti_tables)
    if (sig_matches_sel(sig, tables[t].selclass)) checks.push_back(pack(tables[t].off, tables[t].topo, CM_FORBID_EXISTING, 0));
  for (int t : held_w_tables)
    if (sig_matches_sel(sig, tables[t].selclass)) {
      score.push_back(pack(tables[t].off, tables[t].topo, 0, 0));
      score.push_back((uint64_t)(int64_t)1);
    }
  if (p.affinity.present && (p.affinity.pod.present || p.affinity.anti.present)) {
    if (p.affinity.pod.present) {
      for (auto& t : p.affinity.pod.required) {
        int c = t.topology_key.empty() ? -1 : selclass_of(p, t, &ok);
        if (c < 0) {
          h.flags |= PF_SCHED_ERROR;  // anyPodMatchesPodAffinityTerm errors on every node (predicates.go:1081-1088)
          continue;
        }
        int tb = table_of(TK_MATCH, c, topo_key(t.topology_key));
        checks.push_back(pack(tables[tb].off, tables[tb].topo, CM_REQUIRE, sig_matches_sel(sig, c) ? 1 : 0));
      }
      for (auto& wt : p.affinity.pod.preferred) {
        if (wt.term.topology_key.empty() || wt.weight == 0) continue;
        int c = selclass_of(p, wt.term, &ok);
        if (!ok) {
          h.flags |= PF_PRIO_ERROR;
          continue;
        }
        int tb = table_of(TK_MATCH, c, topo_key(wt.term.topology_key));
        score.push_back(pack(tables[tb].off, tables[tb].topo, 0, 0));
        score.push_back((uint64_t)(int64_t)wt.weight);
      }
    }
    if (p.affinity.anti.present) {
      for (auto& t : p.affinity.anti.required) {
        int c = t.topology_key.empty() ? -1 : selclass_of(p, t, &ok);
        if (c < 0) {
          checks.push_back(pack(0, 0, CM_FAIL_ALWAYS, 0));  // `err != nil || termMatches` fails the node (predicates.go:1336)
          continue;
        }
        int tb = table_of(TK_MATCH, c, topo_key(t.topology_key));
        checks.push_back(pack(tables[tb].off, tables[tb].topo, CM_FORBID_OWN, 0));
      }
      for (auto& wt : p.affinity.anti.preferred) {
        if (wt.term.topology_key.empty() || wt.weight == 0) continue;
        int c = selclass_of(p, wt.term, &ok);
        if (!ok) {
          h.flags |= PF_PRIO_ERROR;
          continue;
        }
        int tb = table_of(TK_MATCH, c, topo_key(wt.term.topology_key));
        score.push_back(pack(tables[tb].off, tables[tb].topo, 0, 0));
        score.push_back((uint64_t)(int64_t)(-wt.weight));
      }
    }
  }
  commit_lists(p, sig, &out->commits, &out->spread_inc);
  if (checks.size() > 255 || score.size() / 2 > 255 || out->commits.size() > 255 || out->spread_inc.size() > 255)
    throw std::runtime_error("pod " + p.key() + " touches more than 255 affinity/spread classes");
  h.n_checks = (uint8_t)checks.size();
  h.n_score = (uint8_t)(score.size() / 2);
  h.n_commit = (uint8_t)out->commits.size();
  h.n_spread_inc = (uint8_t)out->spread_inc.size();

  // --- pool
  auto& pool = out->pool;
  pool.insert(pool.end(), exprs.begin(), exprs.end());
  if (h.flags & PF_HAS_PORTS) pool.insert(pool.end(), ports.begin(), ports.end());
  pool.insert(pool.end(), extcls.begin(), extcls.end());
  pool.insert(pool.end(), checks.begin(), checks.end());
  pool.insert(pool.end(), score.begin(), score.end());
  for (auto& c : out->commits) {
    const Table& t = tables[c.table];
    pool.push_back((uint64_t)t.off | ((uint64_t)(uint32_t)t.topo << 32) | ((uint64_t)(t.kind == TK_MATCH ? 1 : 0) << 40));
    pool.push_back((uint64_t)(int64_t)c.delta);
  }
  for (int c : out->spread_inc) pool.push_back((uint64_t)c);
  pool.insert(pool.end(), tw_ext.begin(), tw_ext.end());
  if ((int)pool.size() > ksb_max_pool_words()) throw std::runtime_error("pod " + p.key() + " requirement block too large");
}

// ------------------------------------------------------------------ host mirror of commit_pod (kernels.cu)

void Engine::apply_node_part_host(int row, const EncodedPod& e, const uint32_t* dev_masks) {
  const KsbPodHeader& h = e.h;
  i64(I64_REQ_CPU, row) += h.add_cpu;
  i64(I64_REQ_MEM, row) += h.add_mem;
  i64(I64_REQ_AGPU, row) += h.add_agpu;
  i64(I64_REQ_EPH, row) += h.add_eph;
  i64(I64_NZ_CPU, row) += h.nz_cpu;
  i64(I64_NZ_MEM, row) += h.nz_mem;
  u32(U32_NUM_PODS, row) += 1;
  for (int s = 0; s < KSB_MAX_SCALARS; ++s) i64(I64_REQ_SCALAR + s, row) += h.scalar_add[s];
  if (h.flags & PF_HAS_PORTS) {
    const uint64_t* ports = e.pool.data() + (size_t)h.n_exprs * W;
    for (int w = 0; w < KSB_PORT_WORDS; ++w) {
      u64(U64_PORT_ANY() + w, row) |= ports[3 * KSB_PORT_WORDS + w];
      u64(U64_PORT_WILD() + w, row) |= ports[4 * KSB_PORT_WORDS + w];
      u64(U64_PORT_EXACT() + w, row) |= ports[5 * KSB_PORT_WORDS + w];
    }
  }
  for (int i = 0; i < h.n_ext; ++i)
    if (h.ext[i].rslot < KSB_MAX_RES) u32(U32_DEV_AVAIL + h.ext[i].rslot, row) &= ~dev_masks[i];
}

void Engine::apply_class_part_host(int row, const std::vector<CommitEntry>& commits, const std::vector<int>& spread_inc,
                                   int sign) {
  for (int c : spread_inc) h_spread[c][row] += (uint32_t)sign;
  for (auto& c : commits) {
    Table& t = tables[c.table];
    uint32_t d = h_topo[(size_t)t.topo * n_pad + row];
    if (d) t.vals[d] += sign * c.delta;
    if (t.kind == TK_MATCH) t.vals[0] += sign * c.delta;
  }
}

// ------------------------------------------------------------------ device synchronisation
//
// The host mirror always holds every row. The device holds this engine's shard: rows
// [lbase, lbase + ln) with column stride ln_pad (the whole table when shard_count == 1).

void Engine::plan_shard() {
  const int R = std::max(1, cfg.shard_count);
  if (R == 1) {
    lbase = 0;
    ln = n;
    ln_pad = n_pad;
    rows_per_shard = n_pad;
    eval_base = 0;
    eval_n = n;
    return;
  }
  rows_per_shard = std::max(256, (((n + R - 1) / R) + 255) & ~255);  // whole segments per shard
  eval_base = std::min(n, cfg.shard_rank * rows_per_shard);
  eval_n = std::max(0, std::min(n - eval_base, rows_per_shard));
  if (cfg.shard_rank == 0) {
    // rank 0 commits: it holds EVERY row (and receives every rank's evaluation results), so that the sequential part
    // never reads another GPU's memory; it still evaluates only its own range of rows
    lbase = 0;
    ln = n;
    ln_pad = n_pad;
  } else {
    lbase = eval_base;
    ln = eval_n;
    ln_pad = rows_per_shard;
  }
}

void Engine::setup_table_pointers() {
  KsbNodeTable& T = dT;
  const size_t np = ln_pad;
  T.n = ln;
  T.n_pad = ln_pad;
  T.row_base = lbase;
  T.n_total = n;
  T.W = W;
  T.n_zones = (int)zone_idx.size() + 1;
  T.w_lr = cfg.w_least_requested;
  T.w_mr = w_most_requested;
  T.w_rl = w_resource_limits;
  T.w_il = w_image_locality;
  T.has_static = (pol_node_label.empty() && w_equal == 0) ? 0 : 1;
  T.has_presence = pol_presence_labels.empty() ? 0 : 1;
  T.w_bra = cfg.w_balanced;
  T.w_ss = cfg.w_selector_spread + w_service_spreading;
  T.w_na = cfg.w_node_affinity;
  T.w_tt = cfg.w_taint_toleration;
  T.w_ipa = cfg.w_inter_pod_affinity;
  T.w_avoid = cfg.w_prefer_avoid;
  fill_table_pointers(&T, b_i64.d, b_u32.d, b_u64.d, b_dev_class.d, b_topo.d, b_spread.d, b_tables.d, np);
}

// Column pointers of a shard from the base addresses of its buffers (also used for peers' mapped buffers).
void Engine::fill_table_pointers(KsbNodeTable* T, void* i64b, void* u32b, void* u64b, void* devb, void* topob,
                                 void* spreadb, void* tablesb, size_t np) const {
  int64_t* a = (int64_t*)i64b;
  auto c64 = [&](int col) { return a + (size_t)col * np; };
  T->alloc_cpu = c64(I64_ALLOC_CPU);
  T->alloc_mem = c64(I64_ALLOC_MEM);
  T->alloc_agpu = c64(I64_ALLOC_AGPU);
  T->alloc_eph = c64(I64_ALLOC_EPH);
  T->req_cpu = c64(I64_REQ_CPU);
  T->req_mem = c64(I64_REQ_MEM);
  T->req_agpu = c64(I64_REQ_AGPU);
  T->req_eph = c64(I64_REQ_EPH);
  T->nz_cpu = c64(I64_NZ_CPU);
  T->nz_mem = c64(I64_NZ_MEM);
  T->alloc_scalar = c64(I64_ALLOC_SCALAR);
  T->req_scalar = c64(I64_REQ_SCALAR);
  uint32_t* b = (uint32_t*)u32b;
  T->alloc_pods = (int32_t*)(b + (size_t)U32_ALLOC_PODS * np);
  T->num_pods = (int32_t*)(b + (size_t)U32_NUM_PODS * np);
  T->flags = b + (size_t)U32_FLAGS * np;
  T->zone = b + (size_t)U32_ZONE * np;
  T->dev_avail = b + (size_t)U32_DEV_AVAIL * np;
  T->static_score = reinterpret_cast<int32_t*>(b + (size_t)U32_STATIC * np);
  uint64_t* c = (uint64_t*)u64b;
  T->labels = c;
  T->tw = TW;
  T->taint_hard = c + (size_t)U64_HARD() * np;
  T->taint_soft = c + (size_t)U64_SOFT() * np;
  T->avoid = c + (size_t)U64_AVOID() * np;
  T->port_any = c + (size_t)U64_PORT_ANY() * np;
  T->port_wild = c + (size_t)U64_PORT_WILD() * np;
  T->port_exact = c + (size_t)U64_PORT_EXACT() * np;
  T->dev_class = (uint16_t*)devb;
  T->topo = (uint32_t*)topob;
  T->spread = (uint32_t*)spreadb;
  T->tables = (int32_t*)tablesb;
}

// Copy `cols` columns of `elem`-byte entries: host column stride n_pad, device column stride ln_pad, rows
// [lbase, lbase + ln). The padding rows of the device columns are zeroed first.
void Engine::upload_family(DevBuf& dst, const void* src, size_t elem, int cols) {
  const size_t bytes = (size_t)cols * ln_pad * elem;
  dst.ensure(std::max<size_t>(bytes, 256));
  CUDA_OK(cudaMemsetAsync(dst.d, 0, bytes, stream));
  if (ln > 0)
    CUDA_OK(cudaMemcpy2DAsync(dst.d, (size_t)ln_pad * elem, (const char*)src + (size_t)lbase * elem, (size_t)n_pad * elem,
                              (size_t)ln * elem, cols, cudaMemcpyHostToDevice, stream));
  stats.h2d_bytes += (size_t)cols * ln * elem;
}

void Engine::upload_all() {
  settle_host_bits();
  pending_fills.clear();
  plan_shard();
  upload_family(b_i64, h_i64.data(), 8, I64_COLS);
  upload_family(b_u32, h_u32.data(), 4, U32_COLS);
  upload_family(b_u64, h_u64.data(), 8, u64_cols());
  upload_family(b_dev_class, h_dev_class.data(), (size_t)KSB_MAX_DEV * 2, KSB_MAX_RES);
  upload_family(b_topo, h_topo.data(), 4, KSB_MAX_TOPO);
  spread_cap_cols = 0;  // force the spread block to be re-uploaded below
  full_upload = false;
  dirty_rows.clear();
  dirty_label_words.clear();
  dirty_topo_cols.clear();
}

void Engine::upload_column(void* dev_base, const void* host_col, size_t elem, int dev_col) {
  if (ln <= 0) return;
  const size_t bytes = (size_t)ln * elem;
  static const bool trace = getenv("KSB_TRACE") != nullptr;
  const auto tc0 = std::chrono::steady_clock::now();
  pin_col.ensure(std::max(bytes, (size_t)ln_pad * 8));  // (sized for the widest column at the first use, i.e. ksb_flush)
  memcpy(pin_col.p, (const char*)host_col + (size_t)lbase * elem, bytes);
  CUDA_OK(cudaMemcpyAsync((char*)dev_base + (size_t)dev_col * ln_pad * elem, pin_col.p, bytes, cudaMemcpyHostToDevice, stream));
  const auto tc1 = std::chrono::steady_clock::now();
  CUDA_OK(cudaStreamSynchronize(stream));  // one staging buffer
  stats.h2d_bytes += bytes;
  if (trace) {
    const auto tc2 = std::chrono::steady_clock::now();
    const double a = std::chrono::duration<double, std::milli>(tc1 - tc0).count(), b = std::chrono::duration<double, std::milli>(tc2 - tc1).count();
    if (a + b > 0.15) fprintf(stderr, "[ksb] column upload %zu bytes: stage+issue %.3f ms, wait %.3f ms\n", bytes, a, b);
  }
}

void Engine::upload_dirty() {
  if (dirty_rows.size() > 64) {
    upload_all();
    return;
  }
  if (!dirty_label_words.empty()) settle_host_bits();
  for (int w : dirty_label_words) upload_column(b_u64.d, &h_u64[(size_t)w * n_pad], 8, w);
  run_pending_fills(dirty_label_words);
  dirty_label_words.clear();
  for (int k : dirty_topo_cols) upload_column(b_topo.d, &h_topo[(size_t)k * n_pad], 4, k);
  dirty_topo_cols.clear();
  for (int gr : dirty_rows) {  // rare path (node update, pod removal): one element per column
    const int r = gr - lbase;
    if (r < 0 || r >= ln) continue;  // another shard's row
    const size_t np = ln_pad;
    for (int c = 0; c < I64_COLS; ++c)
      CUDA_OK(cudaMemcpyAsync((int64_t*)b_i64.d + (size_t)c * np + r, &i64(c, gr), 8, cudaMemcpyHostToDevice, stream));
    for (int c = 0; c < U32_COLS; ++c)
      CUDA_OK(cudaMemcpyAsync((uint32_t*)b_u32.d + (size_t)c * np + r, &u32(c, gr), 4, cudaMemcpyHostToDevice, stream));
    for (int c = 0; c < u64_cols(); ++c)
      CUDA_OK(cudaMemcpyAsync((uint64_t*)b_u64.d + (size_t)c * np + r, &u64(c, gr), 8, cudaMemcpyHostToDevice, stream));
    for (int rs = 0; rs < KSB_MAX_RES; ++rs)
      CUDA_OK(cudaMemcpyAsync((uint16_t*)b_dev_class.d + ((size_t)rs * np + r) * KSB_MAX_DEV,
                              &h_dev_class[((size_t)rs * n_pad + gr) * KSB_MAX_DEV], KSB_MAX_DEV * 2, cudaMemcpyHostToDevice,
                              stream));
    for (int k = 0; k < KSB_MAX_TOPO; ++k)
      CUDA_OK(cudaMemcpyAsync((uint32_t*)b_topo.d + (size_t)k * np + r, &h_topo[(size_t)k * n_pad + gr], 4,
                              cudaMemcpyHostToDevice, stream));
    for (size_t c = 0; c < h_spread.size() && c < spread_cap_cols; ++c)
      CUDA_OK(cudaMemcpyAsync((uint32_t*)b_spread.d + c * np + r, &h_spread[c][gr], 4, cudaMemcpyHostToDevice, stream));
    stats.h2d_bytes += 8 * I64_COLS + 4 * U32_COLS + 8 * u64_cols() + KSB_MAX_RES * KSB_MAX_DEV * 2;
  }
  dirty_rows.clear();
}

void Engine::sync_device() {
  static const bool trace = getenv("KSB_TRACE") != nullptr;  // what a slow table sync was made of, on stderr
  const auto t_sync0 = std::chrono::steady_clock::now();
  struct Report {
    const std::chrono::steady_clock::time_point t0;
    Engine* e;
    bool on;
    size_t lw, tc, dr, sp, pf, at;
    bool ld, fu, ad;
    ~Report() {
      if (!on) return;
      const double ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
      if (ms > 0.3)
        fprintf(stderr, "[ksb] table sync %.3f ms: layout %d full %d arena %d label words %zu topo cols %zu rows %zu spread cols %zu fills %zu appended %zu\n",
                ms, (int)ld, (int)fu, (int)ad, lw, tc, dr, sp, pf, at);
    }
  } report{t_sync0, this, trace, dirty_label_words.size(), dirty_topo_cols.size(), dirty_rows.size(), dirty_spread_cols.size(),
           pending_fills.size(), appended_tables.size(), layout_dirty, full_upload, arena_dirty};
  if (layout_dirty) rebuild_layout();
  if (full_upload) upload_all();
  else upload_dirty();
  // spread columns: capacity grows geometrically (fixed when the shard's buffers are shared with peers);
  // new columns are uploaded individually
  if (h_spread.size() > spread_cap_cols || b_spread.d == nullptr) {
    // 256 columns up front (51 MB at 50,000 rows): growing means a free + malloc + re-upload inside a Schedule call
    size_t cap = std::max<size_t>(std::max<size_t>(256, min_spread_cols), h_spread.size() * 2);
    if (shard_connected && cap * ln_pad * 4 > b_spread.cap)
      throw std::runtime_error("sharded engine: more SelectorSpread classes than the shared buffers were sized for");
    b_spread.ensure(cap * ln_pad * 4);
    CUDA_OK(cudaMemsetAsync(b_spread.d, 0, cap * ln_pad * 4, stream));
    for (size_t c = 0; c < h_spread.size(); ++c) upload_column(b_spread.d, h_spread[c].data(), 4, (int)c);
    spread_cap_cols = cap;
    dirty_spread_cols.clear();
  }
  for (int c : dirty_spread_cols) upload_column(b_spread.d, h_spread[c].data(), 4, c);
  dirty_spread_cols.clear();
  if (!arena_dirty && !appended_tables.empty()) {
    layout_arena();
    grow_keep(b_tables, std::max<size_t>((size_t)arena_words * 4, min_table_bytes));
    for (int id : appended_tables) {
      const Table& t = tables[id];
      CUDA_OK(cudaMemcpyAsync((int32_t*)b_tables.d + t.off, t.vals.data(), t.vals.size() * 4, cudaMemcpyHostToDevice, stream));
      stats.h2d_bytes += t.vals.size() * 4;
    }
    CUDA_OK(cudaStreamSynchronize(stream));  // pageable sources that may change right after
  }
  appended_tables.clear();
  if (arena_dirty) {
    layout_arena();
    std::vector<int32_t> arena(std::max<uint32_t>(arena_words, 1), 0);
    for (auto& t : tables) std::copy(t.vals.begin(), t.vals.end(), arena.begin() + t.off);
    if (shard_connected && arena.size() * 4 > b_tables.cap)
      throw std::runtime_error("sharded engine: affinity tables outgrew the shared buffer");
    b_tables.ensure(std::max<size_t>(arena.size() * 4, min_table_bytes));
    CUDA_OK(cudaMemcpyAsync(b_tables.d, arena.data(), arena.size() * 4, cudaMemcpyHostToDevice, stream));
    CUDA_OK(cudaStreamSynchronize(stream));  // `arena` is a local
    stats.h2d_bytes += arena.size() * 4;
    arena_dirty = false;
  }
  setup_table_pointers();
}

}  // namespace ksb
