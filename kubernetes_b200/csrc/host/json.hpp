// Minimal JSON value tree + recursive-descent parser + writer.
// Used only for the interchange files (synthetic clusters, pod streams, golden
// fixtures) that the test harness feeds to both the oracle and the host mirror.
// Not on the scheduling hot path.
#pragma once
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <map>
#include <memory>
#include <stdexcept>
#include <string>
#include <utility>
#include <vector>

namespace ksb {
namespace json {

struct Value;
using Array = std::vector<Value>;
// Insertion-ordered object: lists of pairs keep the file's order, which matters
// for ordered things (containers, terms); lookups are linear but objects are tiny.
using Object = std::vector<std::pair<std::string, Value>>;

struct Value {
  enum Kind { Null, Bool, Int, Double, String, Arr, Obj } kind = Null;
  bool b = false;
  int64_t i = 0;
  double d = 0.0;
  std::string s;
  std::shared_ptr<Array> a;
  std::shared_ptr<Object> o;

  bool is_null() const { return kind == Null; }
  bool is_obj() const { return kind == Obj; }
  bool is_arr() const { return kind == Arr; }
  bool is_str() const { return kind == String; }

  const Value* find(const char* key) const {
    if (kind != Obj) return nullptr;
    for (auto& kv : *o)
      if (kv.first == key) return &kv.second;
    return nullptr;
  }
  bool has(const char* key) const {
    const Value* v = find(key);
    return v != nullptr;
  }
  int64_t get_int(const char* key, int64_t dflt = 0) const {
    const Value* v = find(key);
    if (!v || v->is_null()) return dflt;
    if (v->kind == Int) return v->i;
    if (v->kind == Double) return (int64_t)v->d;
    if (v->kind == Bool) return v->b ? 1 : 0;
    throw std::runtime_error(std::string("json: key '") + key + "' is not a number");
  }
  double get_num(const char* key, double dflt = 0.0) const {
    const Value* v = find(key);
    if (!v || v->is_null()) return dflt;
    if (v->kind == Int) return (double)v->i;
    if (v->kind == Double) return v->d;
    throw std::runtime_error(std::string("json: key '") + key + "' is not a number");
  }
  bool get_bool(const char* key, bool dflt = false) const {
    const Value* v = find(key);
    if (!v || v->is_null()) return dflt;
    if (v->kind == Bool) return v->b;
    if (v->kind == Int) return v->i != 0;
    throw std::runtime_error(std::string("json: key '") + key + "' is not a bool");
  }
  std::string get_str(const char* key, const std::string& dflt = "") const {
    const Value* v = find(key);
    if (!v || v->is_null()) return dflt;
    if (v->kind != String) throw std::runtime_error(std::string("json: key '") + key + "' is not a string");
    return v->s;
  }
  const Array& arr() const {
    static const Array empty;
    if (kind == Null) return empty;
    if (kind != Arr) throw std::runtime_error("json: not an array");
    return *a;
  }
  const Object& obj() const {
    static const Object empty;
    if (kind == Null) return empty;
    if (kind != Obj) throw std::runtime_error("json: not an object");
    return *o;
  }
  const Array& get_arr(const char* key) const {
    static const Array empty;
    const Value* v = find(key);
    if (!v) return empty;
    return v->arr();
  }
  const Object& get_obj(const char* key) const {
    static const Object empty;
    const Value* v = find(key);
    if (!v) return empty;
    return v->obj();
  }
};

class Parser {
 public:
  Parser(const char* p, size_t n) : p_(p), end_(p + n) {}
  Value parse() {
    Value v = value();
    ws();
    if (p_ != end_) fail("trailing characters");
    return v;
  }

 private:
  const char* p_;
  const char* end_;
  [[noreturn]] void fail(const char* msg) {
    throw std::runtime_error(std::string("json parse error: ") + msg);
  }
  void ws() {
    while (p_ < end_ && (*p_ == ' ' || *p_ == '\n' || *p_ == '\t' || *p_ == '\r')) ++p_;
  }
  Value value() {
    ws();
    if (p_ >= end_) fail("unexpected end");
    char c = *p_;
    if (c == '{') return object();
    if (c == '[') return array();
    if (c == '"') {
      Value v;
      v.kind = Value::String;
      v.s = string();
      return v;
    }
    if (c == 't' && end_ - p_ >= 4 && !memcmp(p_, "true", 4)) {
      p_ += 4;
      Value v;
      v.kind = Value::Bool;
      v.b = true;
      return v;
    }
    if (c == 'f' && end_ - p_ >= 5 && !memcmp(p_, "false", 5)) {
      p_ += 5;
      Value v;
      v.kind = Value::Bool;
      v.b = false;
      return v;
    }
    if (c == 'n' && end_ - p_ >= 4 && !memcmp(p_, "null", 4)) {
      p_ += 4;
      return Value();
    }
    return number();
  }
  Value number() {
    const char* s = p_;
    bool isd = false;
    if (p_ < end_ && (*p_ == '-' || *p_ == '+')) ++p_;
    while (p_ < end_ && ((*p_ >= '0' && *p_ <= '9') || *p_ == '.' || *p_ == 'e' || *p_ == 'E' || *p_ == '-' || *p_ == '+')) {
      if (*p_ == '.' || *p_ == 'e' || *p_ == 'E') isd = true;
      ++p_;
    }
    if (s == p_) fail("bad value");
    std::string t(s, p_ - s);
    Value v;
    if (isd) {
      v.kind = Value::Double;
      v.d = strtod(t.c_str(), nullptr);
    } else {
      v.kind = Value::Int;
      v.i = strtoll(t.c_str(), nullptr, 10);
    }
    return v;
  }
  std::string string() {
    ++p_;  // opening quote
    std::string out;
    while (p_ < end_ && *p_ != '"') {
      char c = *p_++;
      if (c == '\\') {
        if (p_ >= end_) fail("bad escape");
        char e = *p_++;
        switch (e) {
          case 'n': out.push_back('\n'); break;
          case 't': out.push_back('\t'); break;
          case 'r': out.push_back('\r'); break;
          case 'b': out.push_back('\b'); break;
          case 'f': out.push_back('\f'); break;
          case 'u': {
            if (end_ - p_ < 4) fail("bad \\u");
            unsigned cp = (unsigned)strtoul(std::string(p_, 4).c_str(), nullptr, 16);
            p_ += 4;
            if (cp < 0x80) out.push_back((char)cp);
            else if (cp < 0x800) {
              out.push_back((char)(0xC0 | (cp >> 6)));
              out.push_back((char)(0x80 | (cp & 0x3F)));
            } else {
              out.push_back((char)(0xE0 | (cp >> 12)));
              out.push_back((char)(0x80 | ((cp >> 6) & 0x3F)));
              out.push_back((char)(0x80 | (cp & 0x3F)));
            }
            break;
          }
          default: out.push_back(e);
        }
      } else {
        out.push_back(c);
      }
    }
    if (p_ >= end_) fail("unterminated string");
    ++p_;
    return out;
  }
  Value array() {
    ++p_;
    Value v;
    v.kind = Value::Arr;
    v.a = std::make_shared<Array>();
    ws();
    if (p_ < end_ && *p_ == ']') {
      ++p_;
      return v;
    }
    for (;;) {
      v.a->push_back(value());
      ws();
      if (p_ >= end_) fail("unterminated array");
      if (*p_ == ',') {
        ++p_;
        continue;
      }
      if (*p_ == ']') {
        ++p_;
        return v;
      }
      fail("expected , or ]");
    }
  }
  Value object() {
    ++p_;
    Value v;
    v.kind = Value::Obj;
    v.o = std::make_shared<Object>();
    ws();
    if (p_ < end_ && *p_ == '}') {
      ++p_;
      return v;
    }
    for (;;) {
      ws();
      if (p_ >= end_ || *p_ != '"') fail("expected key");
      std::string k = string();
      ws();
      if (p_ >= end_ || *p_ != ':') fail("expected :");
      ++p_;
      v.o->emplace_back(std::move(k), value());
      ws();
      if (p_ >= end_) fail("unterminated object");
      if (*p_ == ',') {
        ++p_;
        continue;
      }
      if (*p_ == '}') {
        ++p_;
        return v;
      }
      fail("expected , or }");
    }
  }
};

inline Value parse(const std::string& text) { return Parser(text.data(), text.size()).parse(); }

inline std::string read_file(const std::string& path) {
  FILE* f = fopen(path.c_str(), "rb");
  if (!f) throw std::runtime_error("cannot open " + path);
  std::string out;
  char buf[1 << 16];
  size_t n;
  while ((n = fread(buf, 1, sizeof buf, f)) > 0) out.append(buf, n);
  fclose(f);
  return out;
}

inline Value parse_file(const std::string& path) { return parse(read_file(path)); }

inline void escape_to(std::string& out, const std::string& s) {
  out.push_back('"');
  for (unsigned char c : s) {
    switch (c) {
      case '"': out += "\\\""; break;
      case '\\': out += "\\\\"; break;
      case '\n': out += "\\n"; break;
      case '\t': out += "\\t"; break;
      case '\r': out += "\\r"; break;
      default:
        if (c < 0x20) {
          char b[8];
          snprintf(b, sizeof b, "\\u%04x", c);
          out += b;
        } else {
          out.push_back((char)c);
        }
    }
  }
  out.push_back('"');
}

}  // namespace json
}  // namespace ksb
