// Minimal JSON DOM (parse + serialise) for the query-spec wire format.
// The wire format is json4s output with ShortTypeHints ("jsonClass" keys):
//   /root/reference/src/main/scala/org/sparklinedata/druid/Utils.scala:38-105
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

namespace sdo {

struct JsonError : std::runtime_error {
  explicit JsonError(const std::string& m) : std::runtime_error(m) {}
};

struct Json {
  enum Type { Null, Bool, Num, Str, Arr, Obj } type = Null;
  bool b = false;
  double num = 0;
  bool is_int = false;
  int64_t inum = 0;
  std::string str;
  std::vector<Json> arr;
  std::vector<std::pair<std::string, Json>> obj;  // insertion order kept

  bool is_null() const { return type == Null; }
  bool is_str() const { return type == Str; }
  bool is_obj() const { return type == Obj; }
  bool is_arr() const { return type == Arr; }
  bool is_num() const { return type == Num; }
  bool is_bool() const { return type == Bool; }

  const Json* find(const char* key) const {
    if (type != Obj) return nullptr;
    for (auto& kv : obj)
      if (kv.first == key) return kv.second.is_null() ? nullptr : &kv.second;
    return nullptr;
  }
  bool has(const char* key) const { return find(key) != nullptr; }
  std::string get_str(const char* key, const std::string& dflt = "") const {
    const Json* j = find(key);
    if (!j) return dflt;
    if (j->type == Str) return j->str;
    if (j->type == Num) return j->is_int ? std::to_string(j->inum) : std::to_string(j->num);
    if (j->type == Bool) return j->b ? "true" : "false";
    throw JsonError(std::string("field '") + key + "' is not a string");
  }
  int64_t get_int(const char* key, int64_t dflt = 0) const {
    const Json* j = find(key);
    if (!j) return dflt;
    if (j->type == Num) return j->is_int ? j->inum : (int64_t)j->num;
    if (j->type == Str) return strtoll(j->str.c_str(), nullptr, 10);
    throw JsonError(std::string("field '") + key + "' is not a number");
  }
  double get_num(const char* key, double dflt = 0) const {
    const Json* j = find(key);
    if (!j) return dflt;
    if (j->type == Num) return j->num;
    if (j->type == Str) return strtod(j->str.c_str(), nullptr);
    throw JsonError(std::string("field '") + key + "' is not a number");
  }
  bool get_bool(const char* key, bool dflt = false) const {
    const Json* j = find(key);
    if (!j) return dflt;
    if (j->type == Bool) return j->b;
    if (j->type == Str) return j->str == "true";
    throw JsonError(std::string("field '") + key + "' is not a boolean");
  }
};

class JsonParser {
 public:
  JsonParser(const char* s, size_t n) : p_(s), e_(s + n) {}
  Json parse() {
    Json j = value(0);
    ws();
    if (p_ != e_) fail("trailing characters after JSON value");
    return j;
  }

 private:
  const char* p_;
  const char* e_;
  [[noreturn]] void fail(const std::string& m) { throw JsonError("JSON parse error: " + m); }
  void ws() {
    while (p_ < e_ && (*p_ == ' ' || *p_ == '\t' || *p_ == '\n' || *p_ == '\r')) p_++;
  }
  Json value(int depth) {
    if (depth > 200) fail("nesting too deep");
    ws();
    if (p_ >= e_) fail("unexpected end of input");
    char c = *p_;
    Json j;
    if (c == '{') {
      p_++;
      j.type = Json::Obj;
      ws();
      if (p_ < e_ && *p_ == '}') { p_++; return j; }
      for (;;) {
        ws();
        if (p_ >= e_ || *p_ != '"') fail("expected object key");
        std::string k = string();
        ws();
        if (p_ >= e_ || *p_ != ':') fail("expected ':'");
        p_++;
        j.obj.emplace_back(std::move(k), value(depth + 1));
        ws();
        if (p_ < e_ && *p_ == ',') { p_++; continue; }
        if (p_ < e_ && *p_ == '}') { p_++; break; }
        fail("expected ',' or '}'");
      }
      return j;
    }
    if (c == '[') {
      p_++;
      j.type = Json::Arr;
      ws();
      if (p_ < e_ && *p_ == ']') { p_++; return j; }
      for (;;) {
        j.arr.push_back(value(depth + 1));
        ws();
        if (p_ < e_ && *p_ == ',') { p_++; continue; }
        if (p_ < e_ && *p_ == ']') { p_++; break; }
        fail("expected ',' or ']'");
      }
      return j;
    }
    if (c == '"') {
      j.type = Json::Str;
      j.str = string();
      return j;
    }
    if (c == 't' && e_ - p_ >= 4 && !strncmp(p_, "true", 4)) { p_ += 4; j.type = Json::Bool; j.b = true; return j; }
    if (c == 'f' && e_ - p_ >= 5 && !strncmp(p_, "false", 5)) { p_ += 5; j.type = Json::Bool; j.b = false; return j; }
    if (c == 'n' && e_ - p_ >= 4 && !strncmp(p_, "null", 4)) { p_ += 4; return j; }
    if (c == '-' || (c >= '0' && c <= '9')) {
      const char* s = p_;
      bool isint = true;
      if (*p_ == '-') p_++;
      while (p_ < e_ && ((*p_ >= '0' && *p_ <= '9') || *p_ == '.' || *p_ == 'e' || *p_ == 'E' || *p_ == '+' || *p_ == '-')) {
        if (*p_ == '.' || *p_ == 'e' || *p_ == 'E') isint = false;
        p_++;
      }
      std::string t(s, p_);
      j.type = Json::Num;
      j.num = strtod(t.c_str(), nullptr);
      j.is_int = isint;
      if (isint) j.inum = strtoll(t.c_str(), nullptr, 10);
      return j;
    }
    fail(std::string("unexpected character '") + c + "'");
  }
  static void utf8(std::string& o, unsigned cp) {
    if (cp < 0x80) o += (char)cp;
    else if (cp < 0x800) { o += (char)(0xC0 | (cp >> 6)); o += (char)(0x80 | (cp & 0x3F)); }
    else if (cp < 0x10000) { o += (char)(0xE0 | (cp >> 12)); o += (char)(0x80 | ((cp >> 6) & 0x3F)); o += (char)(0x80 | (cp & 0x3F)); }
    else { o += (char)(0xF0 | (cp >> 18)); o += (char)(0x80 | ((cp >> 12) & 0x3F)); o += (char)(0x80 | ((cp >> 6) & 0x3F)); o += (char)(0x80 | (cp & 0x3F)); }
  }
  unsigned hex4() {
    if (e_ - p_ < 4) fail("bad \\u escape");
    unsigned v = 0;
    for (int i = 0; i < 4; i++) {
      char c = *p_++;
      v <<= 4;
      if (c >= '0' && c <= '9') v |= c - '0';
      else if (c >= 'a' && c <= 'f') v |= c - 'a' + 10;
      else if (c >= 'A' && c <= 'F') v |= c - 'A' + 10;
      else fail("bad \\u escape");
    }
    return v;
  }
  std::string string() {
    p_++;  // opening quote
    std::string o;
    while (p_ < e_ && *p_ != '"') {
      char c = *p_++;
      if (c != '\\') { o += c; continue; }
      if (p_ >= e_) fail("bad escape");
      char x = *p_++;
      switch (x) {
        case '"': o += '"'; break;
        case '\\': o += '\\'; break;
        case '/': o += '/'; break;
        case 'b': o += '\b'; break;
        case 'f': o += '\f'; break;
        case 'n': o += '\n'; break;
        case 'r': o += '\r'; break;
        case 't': o += '\t'; break;
        case 'u': {
          unsigned cp = hex4();
          if (cp >= 0xD800 && cp < 0xDC00 && e_ - p_ >= 6 && p_[0] == '\\' && p_[1] == 'u') {
            p_ += 2;
            unsigned lo = hex4();
            cp = 0x10000 + ((cp - 0xD800) << 10) + (lo - 0xDC00);
          }
          utf8(o, cp);
          break;
        }
        default: fail("bad escape");
      }
    }
    if (p_ >= e_) fail("unterminated string");
    p_++;
    return o;
  }
};

inline Json json_parse(const char* s, size_t n) { return JsonParser(s, n).parse(); }

inline void json_escape(std::string& o, const std::string& s) {
  o += '"';
  for (unsigned char c : s) {
    switch (c) {
      case '"': o += "\\\""; break;
      case '\\': o += "\\\\"; break;
      case '\n': o += "\\n"; break;
      case '\r': o += "\\r"; break;
      case '\t': o += "\\t"; break;
      default:
        if (c < 0x20) { char b[8]; snprintf(b, sizeof b, "\\u%04x", c); o += b; }
        else o += (char)c;
    }
  }
  o += '"';
}

}  // namespace sdo
