// Wire layout of the records that cross the C-ABI of include/ksched_b200.h.
//
// All integers little-endian. str = u32 length + bytes (no terminator). list<T> = u32 count + items.
// map<str,str> = list of (str key, str value). reslist = list of (str name, i64 quantity) with "cpu"
// in milli-units and every other resource as Quantity.Value() (schedulercache/node_info.go:90-111).
//
// Node record (v1.Node fields NodeInfo.SetNode reads, schedulercache/node_info.go:469-491):
//   str name; u8 has_labels; map labels; reslist allocatable;
//   list<taint: str key, str value, str effect>; list<condition: str type, str status>;
//   u8 unschedulable; u8 avoid_annotation_invalid; list<avoid: str kind, str uid>;
//   list<resource: str name; list<device: str id, str health, map attributes>>;
//   list<image: list<str name>, i64 size_bytes>   (Status.Images; read by ImageLocalityPriority only)
//
// Pod record (v1.Pod fields the path reads):
//   str namespace; str name; map labels; u8 deleting; str node_name;
//   u8 has_controller; str kind; str uid;
//   list<container: reslist requests; reslist limits; list<port: i32 host_port, str protocol, str host_ip>>;
//   list<init container: same shape>;
//   map node_selector;
//   u8 has_affinity;
//     u8 has_node_affinity; u8 has_required; list<term: list<req>>; list<preferred: i32 weight, list<req>>;
//     pod_affinity; pod_anti_affinity   (each: u8 present; list<term>; list<weighted: i32 weight, term>)
//       term = u8 selector_present; map match_labels; list<req>; list<str namespace>; str topology_key
//   list<toleration: str key, str operator, str value, str effect>;
//   list<extended: str name; reslist limits; list<req>; list<str assigned>>;
//   i32 num_volumes
//   i32 priority   (util.GetPodPriority: *Spec.Priority, 0 when unset; read by preemption only)
//   list<str image>   (Spec.Containers[i].Image in container order; read by ImageLocalityPriority only)
//   req = str key; str op; list<str value>
//
// SelectorOwner record: str namespace; str name; map map_selector; u8 selector_present; map match_labels; list<req>
#pragma once
#include <cstdint>
#include <cstring>
#include <stdexcept>
#include <string>
#include <vector>

#include "model.hpp"

namespace ksb {
namespace wire {

class Writer {
 public:
  std::string buf;
  void u8(uint8_t v) { buf.push_back((char)v); }
  void u32(uint32_t v) { buf.append((const char*)&v, 4); }
  void i32(int32_t v) { buf.append((const char*)&v, 4); }
  void i64(int64_t v) { buf.append((const char*)&v, 8); }
  void str(const std::string& s) {
    u32((uint32_t)s.size());
    buf.append(s);
  }
  void map(const StrMap& m) {
    u32((uint32_t)m.size());
    for (auto& kv : m) {
      str(kv.first);
      str(kv.second);
    }
  }
  void reslist(const ResourceList& r) {
    u32((uint32_t)r.size());
    for (auto& kv : r) {
      str(kv.first);
      i64(kv.second);
    }
  }
  void strs(const std::vector<std::string>& v) {
    u32((uint32_t)v.size());
    for (auto& s : v) str(s);
  }
  void reqs(const std::vector<SelectorRequirement>& v) {
    u32((uint32_t)v.size());
    for (auto& r : v) {
      str(r.key);
      str(r.op);
      strs(r.values);
    }
  }
};

class Reader {
 public:
  Reader(const uint8_t* p, size_t n) : p_(p), end_(p + n) {}
  uint8_t u8() {
    need(1);
    return *p_++;
  }
  uint32_t u32() {
    need(4);
    uint32_t v;
    memcpy(&v, p_, 4);
    p_ += 4;
    return v;
  }
  int32_t i32() { return (int32_t)u32(); }
  int64_t i64() {
    need(8);
    int64_t v;
    memcpy(&v, p_, 8);
    p_ += 8;
    return v;
  }
  std::string str() {
    uint32_t n = u32();
    need(n);
    std::string s((const char*)p_, n);
    p_ += n;
    return s;
  }
  StrMap map() {
    StrMap m;
    uint32_t n = u32();
    for (uint32_t i = 0; i < n; ++i) {
      std::string k = str();
      m[k] = str();
    }
    return m;
  }
  ResourceList reslist() {
    ResourceList r;
    uint32_t n = u32();
    for (uint32_t i = 0; i < n; ++i) {
      std::string k = str();
      r[k] = i64();
    }
    return r;
  }
  std::vector<std::string> strs() {
    std::vector<std::string> v;
    uint32_t n = u32();
    for (uint32_t i = 0; i < n; ++i) v.push_back(str());
    return v;
  }
  std::vector<SelectorRequirement> reqs() {
    std::vector<SelectorRequirement> v;
    uint32_t n = u32();
    for (uint32_t i = 0; i < n; ++i) {
      SelectorRequirement r;
      r.key = str();
      r.op = str();
      r.values = strs();
      v.push_back(std::move(r));
    }
    return v;
  }
  bool done() const { return p_ == end_; }

 private:
  const uint8_t* p_;
  const uint8_t* end_;
  void need(size_t n) {
    if ((size_t)(end_ - p_) < n) throw std::runtime_error("wire: truncated record");
  }
};

// ---------------------------------------------------------------- Node

inline void put(Writer& w, const Node& n) {
  w.str(n.name);
  w.u8(n.has_labels);
  w.map(n.labels);
  w.reslist(n.allocatable);
  w.u32((uint32_t)n.taints.size());
  for (auto& t : n.taints) {
    w.str(t.key);
    w.str(t.value);
    w.str(t.effect);
  }
  w.u32((uint32_t)n.conditions.size());
  for (auto& c : n.conditions) {
    w.str(c.type);
    w.str(c.status);
  }
  w.u8(n.unschedulable);
  w.u8(n.avoid_annotation_invalid);
  w.u32((uint32_t)n.avoid.size());
  for (auto& a : n.avoid) {
    w.str(a.kind);
    w.str(a.uid);
  }
  w.u32((uint32_t)n.extended.size());
  for (auto& kv : n.extended) {
    w.str(kv.first);
    w.u32((uint32_t)kv.second.size());
    for (auto& d : kv.second) {
      w.str(d.id);
      w.str(d.health);
      w.map(d.attributes);
    }
  }
  w.u32((uint32_t)n.images.size());
  for (auto& im : n.images) {
    w.strs(im.names);
    w.i64(im.size_bytes);
  }
}

inline Node get_node(Reader& r) {
  Node n;
  n.name = r.str();
  n.has_labels = r.u8() != 0;
  n.labels = r.map();
  n.allocatable = r.reslist();
  uint32_t c = r.u32();
  for (uint32_t i = 0; i < c; ++i) {
    Taint t;
    t.key = r.str();
    t.value = r.str();
    t.effect = r.str();
    n.taints.push_back(std::move(t));
  }
  c = r.u32();
  for (uint32_t i = 0; i < c; ++i) {
    NodeCondition nc;
    nc.type = r.str();
    nc.status = r.str();
    n.conditions.push_back(std::move(nc));
  }
  n.unschedulable = r.u8() != 0;
  n.avoid_annotation_invalid = r.u8() != 0;
  c = r.u32();
  for (uint32_t i = 0; i < c; ++i) {
    AvoidEntry a;
    a.kind = r.str();
    a.uid = r.str();
    n.avoid.push_back(std::move(a));
  }
  c = r.u32();
  for (uint32_t i = 0; i < c; ++i) {
    std::string rname = r.str();
    uint32_t nd = r.u32();
    std::vector<ExtendedDevice> devs;
    for (uint32_t j = 0; j < nd; ++j) {
      ExtendedDevice d;
      d.id = r.str();
      d.health = r.str();
      d.attributes = r.map();
      devs.push_back(std::move(d));
    }
    n.extended[rname] = std::move(devs);
  }
  uint32_t ni = r.u32();
  for (uint32_t i = 0; i < ni; ++i) {
    NodeImage im;
    im.names = r.strs();
    im.size_bytes = r.i64();
    n.images.push_back(std::move(im));
  }
  return n;
}

// ---------------------------------------------------------------- Pod

inline void put_container(Writer& w, const Container& c) {
  w.reslist(c.requests);
  w.reslist(c.limits);
  w.u32((uint32_t)c.ports.size());
  for (auto& p : c.ports) {
    w.i32(p.host_port);
    w.str(p.protocol);
    w.str(p.host_ip);
  }
}
inline Container get_container(Reader& r) {
  Container c;
  c.requests = r.reslist();
  c.limits = r.reslist();
  uint32_t n = r.u32();
  for (uint32_t i = 0; i < n; ++i) {
    ContainerPort p;
    p.host_port = r.i32();
    p.protocol = r.str();
    p.host_ip = r.str();
    c.ports.push_back(std::move(p));
  }
  return c;
}
inline void put_term(Writer& w, const PodAffinityTerm& t) {
  w.u8(t.selector.present);
  w.map(t.selector.match_labels);
  w.reqs(t.selector.match_expressions);
  w.strs(t.namespaces);
  w.str(t.topology_key);
}
inline PodAffinityTerm get_term(Reader& r) {
  PodAffinityTerm t;
  t.selector.present = r.u8() != 0;
  t.selector.match_labels = r.map();
  t.selector.match_expressions = r.reqs();
  t.namespaces = r.strs();
  t.topology_key = r.str();
  return t;
}
inline void put_pod_affinity(Writer& w, const PodAffinity& a) {
  w.u8(a.present);
  w.u32((uint32_t)a.required.size());
  for (auto& t : a.required) put_term(w, t);
  w.u32((uint32_t)a.preferred.size());
  for (auto& t : a.preferred) {
    w.i32(t.weight);
    put_term(w, t.term);
  }
}
inline PodAffinity get_pod_affinity(Reader& r) {
  PodAffinity a;
  a.present = r.u8() != 0;
  uint32_t n = r.u32();
  for (uint32_t i = 0; i < n; ++i) a.required.push_back(get_term(r));
  n = r.u32();
  for (uint32_t i = 0; i < n; ++i) {
    WeightedPodAffinityTerm t;
    t.weight = r.i32();
    t.term = get_term(r);
    a.preferred.push_back(std::move(t));
  }
  return a;
}

inline void put(Writer& w, const Pod& p) {
  w.str(p.ns);
  w.str(p.name);
  w.map(p.labels);
  w.u8(p.deleting);
  w.str(p.node_name);
  w.u8(p.controller.present);
  w.str(p.controller.kind);
  w.str(p.controller.uid);
  w.u32((uint32_t)p.containers.size());
  for (auto& c : p.containers) put_container(w, c);
  w.u32((uint32_t)p.init_containers.size());
  for (auto& c : p.init_containers) put_container(w, c);
  w.map(p.node_selector);
  w.u8(p.affinity.present);
  w.u8(p.affinity.node.present);
  w.u8(p.affinity.node.has_required);
  w.u32((uint32_t)p.affinity.node.required_terms.size());
  for (auto& t : p.affinity.node.required_terms) w.reqs(t.match_expressions);
  w.u32((uint32_t)p.affinity.node.preferred.size());
  for (auto& t : p.affinity.node.preferred) {
    w.i32(t.weight);
    w.reqs(t.preference.match_expressions);
  }
  put_pod_affinity(w, p.affinity.pod);
  put_pod_affinity(w, p.affinity.anti);
  w.u32((uint32_t)p.tolerations.size());
  for (auto& t : p.tolerations) {
    w.str(t.key);
    w.str(t.op);
    w.str(t.value);
    w.str(t.effect);
  }
  w.u32((uint32_t)p.extended.size());
  for (auto& e : p.extended) {
    w.str(e.name);
    w.reslist(e.limits);
    w.reqs(e.required);
    w.strs(e.assigned);
  }
  w.i32(p.num_volumes);
  w.i32(p.priority);
  w.strs(p.images);
}

inline Pod get_pod(Reader& r) {
  Pod p;
  p.ns = r.str();
  p.name = r.str();
  p.labels = r.map();
  p.deleting = r.u8() != 0;
  p.node_name = r.str();
  p.controller.present = r.u8() != 0;
  p.controller.kind = r.str();
  p.controller.uid = r.str();
  uint32_t n = r.u32();
  for (uint32_t i = 0; i < n; ++i) p.containers.push_back(get_container(r));
  n = r.u32();
  for (uint32_t i = 0; i < n; ++i) p.init_containers.push_back(get_container(r));
  p.node_selector = r.map();
  p.affinity.present = r.u8() != 0;
  p.affinity.node.present = r.u8() != 0;
  p.affinity.node.has_required = r.u8() != 0;
  n = r.u32();
  for (uint32_t i = 0; i < n; ++i) {
    NodeSelectorTerm t;
    t.match_expressions = r.reqs();
    p.affinity.node.required_terms.push_back(std::move(t));
  }
  n = r.u32();
  for (uint32_t i = 0; i < n; ++i) {
    PreferredSchedulingTerm t;
    t.weight = r.i32();
    t.preference.match_expressions = r.reqs();
    p.affinity.node.preferred.push_back(std::move(t));
  }
  p.affinity.pod = get_pod_affinity(r);
  p.affinity.anti = get_pod_affinity(r);
  n = r.u32();
  for (uint32_t i = 0; i < n; ++i) {
    Toleration t;
    t.key = r.str();
    t.op = r.str();
    t.value = r.str();
    t.effect = r.str();
    p.tolerations.push_back(std::move(t));
  }
  n = r.u32();
  for (uint32_t i = 0; i < n; ++i) {
    PodExtendedResource e;
    e.name = r.str();
    e.limits = r.reslist();
    e.required = r.reqs();
    e.assigned = r.strs();
    p.extended.push_back(std::move(e));
  }
  p.num_volumes = r.i32();
  p.priority = r.i32();
  p.images = r.strs();
  return p;
}

// ---------------------------------------------------------------- SelectorOwner

inline void put(Writer& w, const SelectorOwner& s) {
  w.str(s.ns);
  w.str(s.name);
  w.map(s.map_selector);
  w.u8(s.selector.present);
  w.map(s.selector.match_labels);
  w.reqs(s.selector.match_expressions);
}
inline SelectorOwner get_owner(Reader& r) {
  SelectorOwner s;
  s.ns = r.str();
  s.name = r.str();
  s.map_selector = r.map();
  s.selector.present = r.u8() != 0;
  s.selector.match_labels = r.map();
  s.selector.match_expressions = r.reqs();
  return s;
}

}  // namespace wire
}  // namespace ksb
