// ORACLE — TEST INFRASTRUCTURE ONLY. CPU restatement of the reference's label
// selector machinery. Nothing under kubernetes_b200/ may include or link this.
//
// Follows (paths relative to /root/reference):
//   staging/src/k8s.io/apimachinery/pkg/labels/selector.go:134-168 (NewRequirement),
//     :193-237 (Requirement.Matches), :313-335 (internalSelector.Matches),
//     :837-857 (SelectorFromSet)
//   staging/src/k8s.io/apimachinery/pkg/util/validation/validation.go:28-66 (IsQualifiedName),
//     :86-101 (IsValidLabelValue), IsDNS1123Subdomain
//   staging/src/k8s.io/apimachinery/pkg/apis/meta/v1/helpers.go:31-67 (LabelSelectorAsSelector)
//   pkg/apis/core/v1/helper/helpers.go:212-242 (NodeSelectorRequirementsAsSelector),
//     :465-498 (ExtendedRequirementsAsSelector)
#pragma once
#include <algorithm>
#include <cerrno>
#include <climits>
#include <cstdlib>
#include <string>
#include <vector>

#include "model.hpp"

namespace oracle {

using ksb::LabelSelector;
using ksb::SelectorRequirement;
using ksb::StrMap;

enum class Op { In, NotIn, Equals, Exists, DoesNotExist, Gt, Lt };

struct Requirement {
  std::string key;
  Op op;
  std::vector<std::string> values;
};

// labels.Selector: Nothing() matches nothing; an empty requirement list matches everything.
struct Selector {
  bool nothing = false;
  std::vector<Requirement> reqs;
};

inline bool is_alnum(char c) {
  return (c >= 'A' && c <= 'Z') || (c >= 'a' && c <= 'z') || (c >= '0' && c <= '9');
}
inline bool is_qname_ext(char c) { return is_alnum(c) || c == '-' || c == '_' || c == '.'; }

// regexp ^([A-Za-z0-9][-A-Za-z0-9_.]*)?[A-Za-z0-9]$
inline bool match_qualified_name_fmt(const std::string& s) {
  if (s.empty()) return false;
  if (!is_alnum(s.back())) return false;
  if (s.size() == 1) return true;
  if (!is_alnum(s.front())) return false;
  for (char c : s)
    if (!is_qname_ext(c)) return false;
  return true;
}

// validation.go: dns1123SubdomainFmt = label(\.label)*, label = [a-z0-9]([-a-z0-9]*[a-z0-9])?, max 253
inline bool is_dns1123_subdomain(const std::string& s) {
  if (s.empty() || s.size() > 253) return false;
  size_t start = 0;
  while (true) {
    size_t dot = s.find('.', start);
    std::string lab = s.substr(start, dot == std::string::npos ? std::string::npos : dot - start);
    if (lab.empty()) return false;
    auto lower_alnum = [](char c) { return (c >= 'a' && c <= 'z') || (c >= '0' && c <= '9'); };
    if (!lower_alnum(lab.front()) || !lower_alnum(lab.back())) return false;
    for (char c : lab)
      if (!lower_alnum(c) && c != '-') return false;
    if (dot == std::string::npos) break;
    start = dot + 1;
  }
  return true;
}

// validation.go:39-66
inline bool is_qualified_name(const std::string& value) {
  size_t slashes = std::count(value.begin(), value.end(), '/');
  std::string name;
  if (slashes == 0) {
    name = value;
  } else if (slashes == 1) {
    size_t p = value.find('/');
    std::string prefix = value.substr(0, p);
    name = value.substr(p + 1);
    if (prefix.empty()) return false;
    if (!is_dns1123_subdomain(prefix)) return false;
  } else {
    return false;
  }
  if (name.empty() || name.size() > 63) return false;
  return match_qualified_name_fmt(name);
}

// validation.go:86-101: ^(qualifiedNameFmt)?$, max 63
inline bool is_valid_label_value(const std::string& v) {
  if (v.size() > 63) return false;
  if (v.empty()) return true;
  return match_qualified_name_fmt(v);
}

// strconv.ParseInt(s, 10, 64): optional sign, digits only, range-checked.
inline bool parse_int64(const std::string& s, int64_t* out) {
  if (s.empty()) return false;
  size_t i = 0;
  if (s[0] == '+' || s[0] == '-') i = 1;
  if (i == s.size()) return false;
  for (size_t j = i; j < s.size(); ++j)
    if (s[j] < '0' || s[j] > '9') return false;
  errno = 0;
  long long v = strtoll(s.c_str(), nullptr, 10);
  if (errno == ERANGE) return false;
  *out = (int64_t)v;
  return true;
}

// selector.go:134-168
inline bool new_requirement(const std::string& key, Op op, std::vector<std::string> vals, Requirement* out,
                            std::string* err) {
  if (!is_qualified_name(key)) {
    *err = "invalid label key " + key;
    return false;
  }
  switch (op) {
    case Op::In:
    case Op::NotIn:
      if (vals.empty()) {
        *err = "for 'in', 'notin' operators, values set can't be empty";
        return false;
      }
      break;
    case Op::Equals:
      if (vals.size() != 1) {
        *err = "exact-match compatibility requires one single value";
        return false;
      }
      break;
    case Op::Exists:
    case Op::DoesNotExist:
      if (!vals.empty()) {
        *err = "values set must be empty for exists and does not exist";
        return false;
      }
      break;
    case Op::Gt:
    case Op::Lt: {
      if (vals.size() != 1) {
        *err = "for 'Gt', 'Lt' operators, exactly one value is required";
        return false;
      }
      int64_t tmp;
      if (!parse_int64(vals[0], &tmp)) {
        *err = "for 'Gt', 'Lt' operators, the value must be an integer";
        return false;
      }
      break;
    }
  }
  for (auto& v : vals)
    if (!is_valid_label_value(v)) {
      *err = "invalid label value " + v;
      return false;
    }
  std::sort(vals.begin(), vals.end());
  out->key = key;
  out->op = op;
  out->values = std::move(vals);
  return true;
}

// selector.go:193-237
inline bool req_matches(const Requirement& r, const StrMap& ls) {
  auto it = ls.find(r.key);
  bool has = it != ls.end();
  auto has_value = [&](const std::string& v) {
    for (auto& x : r.values)
      if (x == v) return true;
    return false;
  };
  switch (r.op) {
    case Op::In:
    case Op::Equals:
      if (!has) return false;
      return has_value(it->second);
    case Op::NotIn:
      if (!has) return true;
      return !has_value(it->second);
    case Op::Exists:
      return has;
    case Op::DoesNotExist:
      return !has;
    case Op::Gt:
    case Op::Lt: {
      if (!has) return false;
      int64_t lv;
      if (!parse_int64(it->second, &lv)) return false;
      if (r.values.size() != 1) return false;
      int64_t rv;
      if (!parse_int64(r.values[0], &rv)) return false;
      return (r.op == Op::Gt && lv > rv) || (r.op == Op::Lt && lv < rv);
    }
  }
  return false;
}

// selector.go:313-335
inline bool sel_matches(const Selector& s, const StrMap& ls) {
  if (s.nothing) return false;
  for (auto& r : s.reqs)
    if (!req_matches(r, ls)) return false;
  return true;
}

// selector.go:837-857: any invalid pair degrades to the empty (match-everything) selector.
inline Selector selector_from_set(const StrMap& set) {
  Selector s;
  for (auto& kv : set) {
    Requirement r;
    std::string err;
    if (!new_requirement(kv.first, Op::Equals, {kv.second}, &r, &err)) return Selector();
    s.reqs.push_back(std::move(r));
  }
  return s;
}

// labels.Set(x).AsSelectorPreValidated(): no validation at all (used by the fake listers).
inline Selector selector_from_validated_set(const StrMap& set) {
  Selector s;
  for (auto& kv : set) {
    Requirement r;
    r.key = kv.first;
    r.op = Op::Equals;
    r.values = {kv.second};
    s.reqs.push_back(std::move(r));
  }
  return s;
}

inline bool node_op_from_string(const std::string& op, Op* out) {
  if (op == "In") *out = Op::In;
  else if (op == "NotIn") *out = Op::NotIn;
  else if (op == "Exists") *out = Op::Exists;
  else if (op == "DoesNotExist") *out = Op::DoesNotExist;
  else if (op == "Gt") *out = Op::Gt;
  else if (op == "Lt") *out = Op::Lt;
  else return false;
  return true;
}

// helpers.go:212-242 and :465-498 (same body for node and device-attribute requirements)
inline bool requirements_as_selector(const std::vector<SelectorRequirement>& nsm, Selector* out, std::string* err) {
  *out = Selector();
  if (nsm.empty()) {
    out->nothing = true;
    return true;
  }
  for (auto& e : nsm) {
    Op op;
    if (!node_op_from_string(e.op, &op)) {
      *err = "\"" + e.op + "\" is not a valid node selector operator";
      return false;
    }
    Requirement r;
    if (!new_requirement(e.key, op, e.values, &r, err)) return false;
    out->reqs.push_back(std::move(r));
  }
  return true;
}

// meta/v1/helpers.go:31-67
inline bool label_selector_as_selector(const LabelSelector& ps, Selector* out, std::string* err) {
  *out = Selector();
  if (!ps.present) {
    out->nothing = true;
    return true;
  }
  if (ps.match_labels.size() + ps.match_expressions.size() == 0) return true;  // Everything
  for (auto& kv : ps.match_labels) {
    Requirement r;
    if (!new_requirement(kv.first, Op::Equals, {kv.second}, &r, err)) return false;
    out->reqs.push_back(std::move(r));
  }
  for (auto& e : ps.match_expressions) {
    Op op;
    if (e.op == "In") op = Op::In;
    else if (e.op == "NotIn") op = Op::NotIn;
    else if (e.op == "Exists") op = Op::Exists;
    else if (e.op == "DoesNotExist") op = Op::DoesNotExist;
    else {
      *err = "\"" + e.op + "\" is not a valid pod selector operator";
      return false;
    }
    Requirement r;
    if (!new_requirement(e.key, op, e.values, &r, err)) return false;
    out->reqs.push_back(std::move(r));
  }
  return true;
}

}  // namespace oracle
