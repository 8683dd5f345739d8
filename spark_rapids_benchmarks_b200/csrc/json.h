// Minimal JSON reader for the plan IR / engine config (objects, arrays, strings, numbers,
// true/false/null). Numbers keep both an int64 and a double reading.
#pragma once
#include <map>
#include <memory>
#include "common.h"

namespace nds {

struct Json {
  enum Kind { Null, Bool, Num, Str, Arr, Obj } kind = Null;
  bool b = false;
  double d = 0;
  int64_t i = 0;
  bool is_int = false;
  std::string s;
  std::vector<Json> a;
  std::vector<std::pair<std::string, Json>> o;

  const Json* find(const std::string& k) const {
    for (auto& kv : o)
      if (kv.first == k) return &kv.second;
    return nullptr;
  }
  bool has(const std::string& k) const { return find(k) != nullptr; }
  const Json& at(const std::string& k) const {
    const Json* j = find(k);
    if (!j) fail("plan JSON: missing key '" + k + "'");
    return *j;
  }
  const std::string& str(const std::string& k) const {
    const Json& j = at(k);
    if (j.kind != Str) fail("plan JSON: key '" + k + "' is not a string");
    return j.s;
  }
  std::string str_or(const std::string& k, const std::string& dflt) const {
    const Json* j = find(k);
    return (j && j->kind == Str) ? j->s : dflt;
  }
  int64_t int_or(const std::string& k, int64_t dflt) const {
    const Json* j = find(k);
    return (j && j->kind == Num) ? j->i : dflt;
  }
  bool bool_or(const std::string& k, bool dflt) const {
    const Json* j = find(k);
    if (!j) return dflt;
    if (j->kind == Bool) return j->b;
    if (j->kind == Num) return j->i != 0;
    return dflt;
  }
};

class JsonParser {
 public:
  explicit JsonParser(const char* text) : p_(text) {}
  Json parse() {
    Json j = value();
    ws();
    if (*p_) fail("JSON: trailing characters");
    return j;
  }

 private:
  const char* p_;
  void ws() {
    while (*p_ == ' ' || *p_ == '\n' || *p_ == '\t' || *p_ == '\r') ++p_;
  }
  Json value() {
    ws();
    Json j;
    if (*p_ == '{') {
      ++p_;
      j.kind = Json::Obj;
      ws();
      if (*p_ == '}') { ++p_; return j; }
      for (;;) {
        ws();
        std::string k = string();
        ws();
        if (*p_ != ':') fail("JSON: expected ':'");
        ++p_;
        j.o.emplace_back(k, value());
        ws();
        if (*p_ == ',') { ++p_; continue; }
        if (*p_ == '}') { ++p_; return j; }
        fail("JSON: expected ',' or '}'");
      }
    }
    if (*p_ == '[') {
      ++p_;
      j.kind = Json::Arr;
      ws();
      if (*p_ == ']') { ++p_; return j; }
      for (;;) {
        j.a.push_back(value());
        ws();
        if (*p_ == ',') { ++p_; continue; }
        if (*p_ == ']') { ++p_; return j; }
        fail("JSON: expected ',' or ']'");
      }
    }
    if (*p_ == '"') {
      j.kind = Json::Str;
      j.s = string();
      return j;
    }
    if (!strncmp(p_, "true", 4)) { p_ += 4; j.kind = Json::Bool; j.b = true; return j; }
    if (!strncmp(p_, "false", 5)) { p_ += 5; j.kind = Json::Bool; j.b = false; return j; }
    if (!strncmp(p_, "null", 4)) { p_ += 4; return j; }
    // number
    const char* s = p_;
    bool isint = true;
    if (*p_ == '-' || *p_ == '+') ++p_;
    while ((*p_ >= '0' && *p_ <= '9') || *p_ == '.' || *p_ == 'e' || *p_ == 'E' || *p_ == '-' || *p_ == '+') {
      if (*p_ == '.' || *p_ == 'e' || *p_ == 'E') isint = false;
      ++p_;
    }
    if (s == p_) fail(std::string("JSON: unexpected character '") + *p_ + "'");
    std::string num(s, p_);
    j.kind = Json::Num;
    j.is_int = isint;
    j.d = strtod(num.c_str(), nullptr);
    j.i = isint ? strtoll(num.c_str(), nullptr, 10) : (int64_t)j.d;
    return j;
  }
  std::string string() {
    if (*p_ != '"') fail("JSON: expected string");
    ++p_;
    std::string out;
    while (*p_ && *p_ != '"') {
      if (*p_ == '\\') {
        ++p_;
        switch (*p_) {
          case 'n': out += '\n'; break;
          case 't': out += '\t'; break;
          case 'r': out += '\r'; break;
          case 'b': out += '\b'; break;
          case 'f': out += '\f'; break;
          case 'u': {
            unsigned cp = 0;
            for (int k = 1; k <= 4; ++k) {
              char c = p_[k];
              cp = cp * 16 + (c >= '0' && c <= '9' ? c - '0' : (c | 32) - 'a' + 10);
            }
            p_ += 4;
            if (cp < 0x80) out += (char)cp;
            else if (cp < 0x800) { out += (char)(0xC0 | (cp >> 6)); out += (char)(0x80 | (cp & 0x3F)); }
            else { out += (char)(0xE0 | (cp >> 12)); out += (char)(0x80 | ((cp >> 6) & 0x3F)); out += (char)(0x80 | (cp & 0x3F)); }
            break;
          }
          default: out += *p_;
        }
        ++p_;
      } else {
        out += *p_++;
      }
    }
    if (*p_ != '"') fail("JSON: unterminated string");
    ++p_;
    return out;
  }
};

inline std::string json_escape(const std::string& s) {
  std::string o;
  for (unsigned char c : s) {
    if (c == '"' || c == '\\') { o += '\\'; o += (char)c; }
    else if (c == '\n') o += "\\n";
    else if (c < 0x20) { char b[8]; snprintf(b, sizeof b, "\\u%04x", c); o += b; }
    else o += (char)c;
  }
  return o;
}

}  // namespace nds
