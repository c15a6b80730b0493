// json.hpp — a small self-contained JSON value (parse / serialise / build) for the control plane.
// Objects keep keys sorted (std::map), which also gives the deterministic "hyperparameters visited in
// sorted-name order" the searchers rely on.
#pragma once
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <map>
#include <memory>
#include <sstream>
#include <stdexcept>
#include <string>
#include <vector>

namespace det {

class Json {
 public:
  enum Type { Null, Bool, Int, Double, String, Array, Object };
  using array_t = std::vector<Json>;
  using object_t = std::map<std::string, Json>;

  Json() : type_(Null) {}
  Json(std::nullptr_t) : type_(Null) {}
  Json(bool b) : type_(Bool), b_(b) {}
  Json(int v) : type_(Int), i_(v) {}
  Json(long v) : type_(Int), i_(v) {}
  Json(long long v) : type_(Int), i_(v) {}
  Json(unsigned v) : type_(Int), i_(v) {}
  Json(unsigned long v) : type_(Int), i_(static_cast<long long>(v)) {}
  Json(double v) : type_(Double), d_(v) {}
  Json(const char* s) : type_(String), s_(s) {}
  Json(const std::string& s) : type_(String), s_(s) {}
  Json(const array_t& a) : type_(Array), a_(std::make_shared<array_t>(a)) {}
  Json(const object_t& o) : type_(Object), o_(std::make_shared<object_t>(o)) {}

  static Json array() { return Json(array_t{}); }
  static Json object() { return Json(object_t{}); }

  Type type() const { return type_; }
  bool is_null() const { return type_ == Null; }
  bool is_bool() const { return type_ == Bool; }
  bool is_int() const { return type_ == Int; }
  bool is_number() const { return type_ == Int || type_ == Double; }
  bool is_string() const { return type_ == String; }
  bool is_array() const { return type_ == Array; }
  bool is_object() const { return type_ == Object; }

  bool as_bool(bool def = false) const { return type_ == Bool ? b_ : (is_number() ? as_double() != 0 : def); }
  long long as_int(long long def = 0) const {
    if (type_ == Int) return i_;
    if (type_ == Double) return static_cast<long long>(d_);
    return def;
  }
  double as_double(double def = 0) const {
    if (type_ == Double) return d_;
    if (type_ == Int) return static_cast<double>(i_);
    return def;
  }
  const std::string& as_string() const {
    static const std::string empty;
    return type_ == String ? s_ : empty;
  }
  std::string str_or(const std::string& def) const { return type_ == String ? s_ : def; }

  // containers (copy-on-write not needed: values are mutated only while being built)
  array_t& arr() {
    if (type_ != Array) { *this = array(); }
    return *a_;
  }
  const array_t& arr() const {
    static const array_t empty;
    return type_ == Array ? *a_ : empty;
  }
  object_t& obj() {
    if (type_ != Object) { *this = object(); }
    return *o_;
  }
  const object_t& obj() const {
    static const object_t empty;
    return type_ == Object ? *o_ : empty;
  }
  size_t size() const { return type_ == Array ? a_->size() : (type_ == Object ? o_->size() : 0); }
  bool contains(const std::string& k) const { return type_ == Object && o_->count(k) > 0; }
  Json& operator[](const std::string& k) { return obj()[k]; }
  const Json& operator[](const std::string& k) const {
    static const Json null;
    if (type_ != Object) return null;
    auto it = o_->find(k);
    return it == o_->end() ? null : it->second;
  }
  const Json& at(size_t i) const {
    static const Json null;
    return (type_ == Array && i < a_->size()) ? (*a_)[i] : null;
  }
  void push_back(const Json& v) { arr().push_back(v); }

  Json deep_copy() const {
    if (type_ == Array) {
      Json r = array();
      for (auto& v : *a_) r.push_back(v.deep_copy());
      return r;
    }
    if (type_ == Object) {
      Json r = object();
      for (auto& kv : *o_) r[kv.first] = kv.second.deep_copy();
      return r;
    }
    return *this;
  }

  bool operator==(const Json& o) const {
    if (is_number() && o.is_number()) return as_double() == o.as_double();
    if (type_ != o.type_) return false;
    switch (type_) {
      case Null: return true;
      case Bool: return b_ == o.b_;
      case String: return s_ == o.s_;
      case Array: return *a_ == *o.a_;
      case Object: return *o_ == *o.o_;
      default: return false;
    }
  }
  bool operator!=(const Json& o) const { return !(*this == o); }

  // ---- serialise ----
  std::string dump(int indent = -1) const {
    std::string out;
    dump_to(out, indent, 0);
    return out;
  }

  // ---- parse ----
  static Json parse(const std::string& text) {
    size_t pos = 0;
    Json v = parse_value(text, pos);
    skip_ws(text, pos);
    if (pos != text.size()) throw std::runtime_error("json: trailing characters at offset " + std::to_string(pos));
    return v;
  }

 private:
  Type type_;
  bool b_ = false;
  long long i_ = 0;
  double d_ = 0;
  std::string s_;
  std::shared_ptr<array_t> a_;
  std::shared_ptr<object_t> o_;

  static void escape_to(std::string& out, const std::string& s) {
    out.push_back('"');
    for (unsigned char c : s) {
      switch (c) {
        case '"': out += "\\\""; break;
        case '\\': out += "\\\\"; break;
        case '\n': out += "\\n"; break;
        case '\r': out += "\\r"; break;
        case '\t': out += "\\t"; break;
        default:
          if (c < 0x20) {
            char buf[8];
            snprintf(buf, sizeof buf, "\\u%04x", c);
            out += buf;
          } else {
            out.push_back(static_cast<char>(c));
          }
      }
    }
    out.push_back('"');
  }

  void dump_to(std::string& out, int indent, int depth) const {
    auto nl = [&](int d) {
      if (indent >= 0) {
        out.push_back('\n');
        out.append(static_cast<size_t>(indent * d), ' ');
      }
    };
    switch (type_) {
      case Null: out += "null"; break;
      case Bool: out += b_ ? "true" : "false"; break;
      case Int: out += std::to_string(i_); break;
      case Double: {
        if (std::isnan(d_) || std::isinf(d_)) { out += "null"; break; }
        char buf[40];
        snprintf(buf, sizeof buf, "%.17g", d_);
        std::string s(buf);
        if (s.find_first_of(".eE") == std::string::npos) s += ".0";
        out += s;
        break;
      }
      case String: escape_to(out, s_); break;
      case Array: {
        out.push_back('[');
        bool first = true;
        for (auto& v : *a_) {
          if (!first) out.push_back(',');
          first = false;
          nl(depth + 1);
          v.dump_to(out, indent, depth + 1);
        }
        if (!a_->empty()) nl(depth);
        out.push_back(']');
        break;
      }
      case Object: {
        out.push_back('{');
        bool first = true;
        for (auto& kv : *o_) {
          if (!first) out.push_back(',');
          first = false;
          nl(depth + 1);
          escape_to(out, kv.first);
          out += indent >= 0 ? ": " : ":";
          kv.second.dump_to(out, indent, depth + 1);
        }
        if (!o_->empty()) nl(depth);
        out.push_back('}');
        break;
      }
    }
  }

  static void skip_ws(const std::string& t, size_t& p) {
    while (p < t.size() && (t[p] == ' ' || t[p] == '\n' || t[p] == '\r' || t[p] == '\t')) ++p;
  }
  [[noreturn]] static void fail(const std::string& what, size_t p) {
    throw std::runtime_error("json: " + what + " at offset " + std::to_string(p));
  }
  static void append_utf8(std::string& out, unsigned cp) {
    if (cp < 0x80) out.push_back(static_cast<char>(cp));
    else if (cp < 0x800) { out.push_back(static_cast<char>(0xC0 | (cp >> 6))); out.push_back(static_cast<char>(0x80 | (cp & 0x3F))); }
    else if (cp < 0x10000) {
      out.push_back(static_cast<char>(0xE0 | (cp >> 12)));
      out.push_back(static_cast<char>(0x80 | ((cp >> 6) & 0x3F)));
      out.push_back(static_cast<char>(0x80 | (cp & 0x3F)));
    } else {
      out.push_back(static_cast<char>(0xF0 | (cp >> 18)));
      out.push_back(static_cast<char>(0x80 | ((cp >> 12) & 0x3F)));
      out.push_back(static_cast<char>(0x80 | ((cp >> 6) & 0x3F)));
      out.push_back(static_cast<char>(0x80 | (cp & 0x3F)));
    }
  }
  static std::string parse_string(const std::string& t, size_t& p) {
    if (t[p] != '"') fail("expected string", p);
    ++p;
    std::string out;
    while (p < t.size() && t[p] != '"') {
      char c = t[p++];
      if (c != '\\') { out.push_back(c); continue; }
      if (p >= t.size()) fail("bad escape", p);
      char e = t[p++];
      switch (e) {
        case '"': out.push_back('"'); break;
        case '\\': out.push_back('\\'); break;
        case '/': out.push_back('/'); break;
        case 'b': out.push_back('\b'); break;
        case 'f': out.push_back('\f'); break;
        case 'n': out.push_back('\n'); break;
        case 'r': out.push_back('\r'); break;
        case 't': out.push_back('\t'); break;
        case 'u': {
          if (p + 4 > t.size()) fail("bad \\u escape", p);
          unsigned cp = static_cast<unsigned>(strtoul(t.substr(p, 4).c_str(), nullptr, 16));
          p += 4;
          if (cp >= 0xD800 && cp <= 0xDBFF && p + 6 <= t.size() && t[p] == '\\' && t[p + 1] == 'u') {
            unsigned lo = static_cast<unsigned>(strtoul(t.substr(p + 2, 4).c_str(), nullptr, 16));
            p += 6;
            cp = 0x10000 + ((cp - 0xD800) << 10) + (lo - 0xDC00);
          }
          append_utf8(out, cp);
          break;
        }
        default: fail("bad escape", p);
      }
    }
    if (p >= t.size()) fail("unterminated string", p);
    ++p;
    return out;
  }
  static Json parse_value(const std::string& t, size_t& p) {
    skip_ws(t, p);
    if (p >= t.size()) fail("unexpected end", p);
    char c = t[p];
    if (c == '{') {
      ++p;
      Json o = object();
      skip_ws(t, p);
      if (p < t.size() && t[p] == '}') { ++p; return o; }
      while (true) {
        skip_ws(t, p);
        std::string k = parse_string(t, p);
        skip_ws(t, p);
        if (p >= t.size() || t[p] != ':') fail("expected ':'", p);
        ++p;
        o[k] = parse_value(t, p);
        skip_ws(t, p);
        if (p < t.size() && t[p] == ',') { ++p; continue; }
        if (p < t.size() && t[p] == '}') { ++p; return o; }
        fail("expected ',' or '}'", p);
      }
    }
    if (c == '[') {
      ++p;
      Json a = array();
      skip_ws(t, p);
      if (p < t.size() && t[p] == ']') { ++p; return a; }
      while (true) {
        a.push_back(parse_value(t, p));
        skip_ws(t, p);
        if (p < t.size() && t[p] == ',') { ++p; continue; }
        if (p < t.size() && t[p] == ']') { ++p; return a; }
        fail("expected ',' or ']'", p);
      }
    }
    if (c == '"') return Json(parse_string(t, p));
    if (t.compare(p, 4, "true") == 0) { p += 4; return Json(true); }
    if (t.compare(p, 5, "false") == 0) { p += 5; return Json(false); }
    if (t.compare(p, 4, "null") == 0) { p += 4; return Json(); }
    if (t.compare(p, 3, "NaN") == 0) { p += 3; return Json(std::nan("")); }
    size_t start = p;
    if (t[p] == '-' || t[p] == '+') ++p;
    bool is_float = false;
    while (p < t.size() && (isdigit(static_cast<unsigned char>(t[p])) || t[p] == '.' || t[p] == 'e' || t[p] == 'E' || t[p] == '-' || t[p] == '+')) {
      if (t[p] == '.' || t[p] == 'e' || t[p] == 'E') is_float = true;
      ++p;
    }
    if (p == start) fail("unexpected character", p);
    std::string num = t.substr(start, p - start);
    if (!is_float) {
      errno = 0;
      long long v = strtoll(num.c_str(), nullptr, 10);
      if (errno == 0) return Json(v);
    }
    return Json(strtod(num.c_str(), nullptr));
  }
};

}  // namespace det
