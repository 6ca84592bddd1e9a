// Host-side label-selector semantics the encoder needs to turn string requirements into
// bit masks and class memberships. Behaviour follows the reference's labels package
// (staging/src/k8s.io/apimachinery/pkg/labels/selector.go:134-237, :837-857), its validators
// (pkg/util/validation/validation.go:39-101) and the conversions in
// pkg/apis/core/v1/helper/helpers.go:212-242, :465-498 and apis/meta/v1/helpers.go:31-67.
#pragma once
#include <algorithm>
#include <cerrno>
#include <cstdlib>
#include <string>
#include <vector>

#include "model.hpp"

namespace ksb {
namespace sel {

enum Op : uint8_t { OP_IN, OP_NOT_IN, OP_EXISTS, OP_NOT_EXISTS, OP_GT, OP_LT };

struct Req {
  std::string key;
  Op op;
  std::vector<std::string> values;  // sorted
};

struct Selector {
  bool nothing = false;  // labels.Nothing()
  std::vector<Req> reqs; // empty + !nothing = labels.Everything()
};

inline bool alnum(unsigned char c) { return (c >= '0' && c <= '9') || (c >= 'a' && c <= 'z') || (c >= 'A' && c <= 'Z'); }

inline bool name_chars_ok(const std::string& s) {  // ([A-Za-z0-9][-A-Za-z0-9_.]*)?[A-Za-z0-9]
  if (s.empty() || !alnum(s.front()) || !alnum(s.back())) return false;
  for (unsigned char c : s)
    if (!alnum(c) && c != '-' && c != '_' && c != '.') return false;
  return true;
}

inline bool dns_subdomain_ok(const std::string& s) {
  if (s.empty() || s.size() > 253) return false;
  size_t i = 0;
  while (i <= s.size()) {
    size_t j = s.find('.', i);
    if (j == std::string::npos) j = s.size();
    if (j == i) return false;
    for (size_t k = i; k < j; ++k) {
      unsigned char c = s[k];
      bool lower = (c >= 'a' && c <= 'z') || (c >= '0' && c <= '9');
      if (!lower && !(c == '-' && k != i && k != j - 1)) return false;
    }
    i = j + 1;
  }
  return true;
}

inline bool label_key_ok(const std::string& key) {
  size_t slash = key.find('/');
  std::string name = key;
  if (slash != std::string::npos) {
    if (key.find('/', slash + 1) != std::string::npos) return false;
    if (!dns_subdomain_ok(key.substr(0, slash))) return false;
    name = key.substr(slash + 1);
  }
  return name.size() <= 63 && name_chars_ok(name);
}

inline bool label_value_ok(const std::string& v) { return v.size() <= 63 && (v.empty() || name_chars_ok(v)); }

inline bool to_int64(const std::string& s, int64_t* out) {
  size_t i = (!s.empty() && (s[0] == '+' || s[0] == '-')) ? 1 : 0;
  if (i >= s.size()) return false;
  for (size_t k = i; k < s.size(); ++k)
    if (s[k] < '0' || s[k] > '9') return false;
  errno = 0;
  long long v = strtoll(s.c_str(), nullptr, 10);
  if (errno == ERANGE) return false;
  *out = v;
  return true;
}

// labels.NewRequirement. Returns false where the reference returns an error.
inline bool make_req(const std::string& key, Op op, const std::vector<std::string>& values, Req* out) {
  if (!label_key_ok(key)) return false;
  if ((op == OP_IN || op == OP_NOT_IN) && values.empty()) return false;
  if ((op == OP_EXISTS || op == OP_NOT_EXISTS) && !values.empty()) return false;
  if (op == OP_GT || op == OP_LT) {
    int64_t tmp;
    if (values.size() != 1 || !to_int64(values[0], &tmp)) return false;
  }
  for (auto& v : values)
    if (!label_value_ok(v)) return false;
  out->key = key;
  out->op = op;
  out->values = values;
  std::sort(out->values.begin(), out->values.end());
  return true;
}

inline bool op_from_node_string(const std::string& s, Op* op) {
  if (s == "In") *op = OP_IN;
  else if (s == "NotIn") *op = OP_NOT_IN;
  else if (s == "Exists") *op = OP_EXISTS;
  else if (s == "DoesNotExist") *op = OP_NOT_EXISTS;
  else if (s == "Gt") *op = OP_GT;
  else if (s == "Lt") *op = OP_LT;
  else return false;
  return true;
}

// NodeSelectorRequirementsAsSelector / ExtendedRequirementsAsSelector. false = conversion error.
inline bool from_node_requirements(const std::vector<SelectorRequirement>& in, Selector* out) {
  *out = Selector();
  if (in.empty()) {
    out->nothing = true;
    return true;
  }
  for (auto& e : in) {
    Op op;
    Req r;
    if (!op_from_node_string(e.op, &op) || !make_req(e.key, op, e.values, &r)) return false;
    out->reqs.push_back(std::move(r));
  }
  return true;
}

// metav1.LabelSelectorAsSelector. false = conversion error.
inline bool from_label_selector(const LabelSelector& ps, Selector* out) {
  *out = Selector();
  if (!ps.present) {
    out->nothing = true;
    return true;
  }
  for (auto& kv : ps.match_labels) {
    Req r;
    if (!make_req(kv.first, OP_IN, {kv.second}, &r)) return false;
    out->reqs.push_back(std::move(r));
  }
  for (auto& e : ps.match_expressions) {
    Op op;
    if (!op_from_node_string(e.op, &op) || op == OP_GT || op == OP_LT) return false;
    Req r;
    if (!make_req(e.key, op, e.values, &r)) return false;
    out->reqs.push_back(std::move(r));
  }
  return true;
}

// labels.SelectorFromSet: one invalid pair turns the whole selector into Everything().
inline Selector from_set(const StrMap& m) {
  Selector s;
  for (auto& kv : m) {
    Req r;
    if (!make_req(kv.first, OP_IN, {kv.second}, &r)) return Selector();
    s.reqs.push_back(std::move(r));
  }
  return s;
}

inline bool req_matches(const Req& r, const StrMap& labels) {
  auto it = labels.find(r.key);
  bool has = it != labels.end();
  switch (r.op) {
    case OP_IN:
      return has && std::binary_search(r.values.begin(), r.values.end(), it->second);
    case OP_NOT_IN:
      return !has || !std::binary_search(r.values.begin(), r.values.end(), it->second);
    case OP_EXISTS:
      return has;
    case OP_NOT_EXISTS:
      return !has;
    case OP_GT:
    case OP_LT: {
      int64_t lv, rv;
      if (!has || !to_int64(it->second, &lv) || r.values.size() != 1 || !to_int64(r.values[0], &rv)) return false;
      return r.op == OP_GT ? lv > rv : lv < rv;
    }
  }
  return false;
}

inline bool matches(const Selector& s, const StrMap& labels) {
  if (s.nothing) return false;
  for (auto& r : s.reqs)
    if (!req_matches(r, labels)) return false;
  return true;
}

// plain equality on every pair, no validation: labels.Set.AsSelectorPreValidated (fake listers)
inline bool set_matches(const StrMap& want, const StrMap& labels) {
  for (auto& kv : want) {
    auto it = labels.find(kv.first);
    if (it == labels.end() || it->second != kv.second) return false;
  }
  return true;
}

// canonical text of a selector, used as a dictionary key for classes
inline std::string canonical(const Selector& s) {
  if (s.nothing) return "<nothing>";
  std::vector<std::string> parts;
  for (auto& r : s.reqs) {
    std::string t = r.key;
    t.push_back('\x01');
    t.push_back((char)('0' + r.op));
    for (auto& v : r.values) {
      t.push_back('\x02');
      t += v;
    }
    parts.push_back(std::move(t));
  }
  std::sort(parts.begin(), parts.end());
  std::string out;
  for (auto& p : parts) {
    out += p;
    out.push_back('\x03');
  }
  return out;
}

}  // namespace sel
}  // namespace ksb
