// json.h — a small JSON reader for the scene / renderer files of the reference.
//
// It keeps what the reference's loader depends on (src/core/loader.rs:413-470, built on serde_json): numbers remember
// whether they were written as an integer literal or not (Q23: an integer literal is NOT accepted where a float is
// expected), and objects keep their keys in file order.
#pragma once
#include <cstdint>
#include <cstdlib>
#include <memory>
#include <stdexcept>
#include <string>
#include <utility>
#include <vector>

namespace spt_host {

struct Json {
  enum Kind { Null, Bool, Int, Float, String, Array, Object } kind = Null;
  bool b = false;
  int64_t i = 0;
  double f = 0.0;
  std::string s;
  std::vector<Json> arr;
  std::vector<std::pair<std::string, Json>> obj;

  bool is_float() const { return kind == Float; }
  bool is_int() const { return kind == Int; }
  const Json* find(const std::string& k) const {
    for (const auto& kv : obj) if (kv.first == k) return &kv.second;
    return nullptr;
  }
};

class JsonParser {
 public:
  explicit JsonParser(const std::string& text) : t_(text) {}
  Json parse() {
    Json v = value();
    ws();
    if (p_ != t_.size()) fail("trailing characters");
    return v;
  }

 private:
  const std::string& t_;
  size_t p_ = 0;
  [[noreturn]] void fail(const char* m) const { throw std::runtime_error(std::string("json: ") + m + " at offset " + std::to_string(p_)); }
  void ws() { while (p_ < t_.size() && (t_[p_] == ' ' || t_[p_] == '\n' || t_[p_] == '\t' || t_[p_] == '\r')) p_++; }
  bool eat(char c) { ws(); if (p_ < t_.size() && t_[p_] == c) { p_++; return true; } return false; }
  Json value() {
    ws();
    if (p_ >= t_.size()) fail("unexpected end");
    const char c = t_[p_];
    Json v;
    if (c == '{') {
      p_++; v.kind = Json::Object;
      if (eat('}')) return v;
      do {
        ws();
        if (p_ >= t_.size() || t_[p_] != '"') fail("expected a key");
        std::string k = string();
        if (!eat(':')) fail("expected ':'");
        v.obj.emplace_back(std::move(k), value());
      } while (eat(','));
      if (!eat('}')) fail("expected '}'");
    } else if (c == '[') {
      p_++; v.kind = Json::Array;
      if (eat(']')) return v;
      do v.arr.push_back(value()); while (eat(','));
      if (!eat(']')) fail("expected ']'");
    } else if (c == '"') {
      v.kind = Json::String; v.s = string();
    } else if (t_.compare(p_, 4, "true") == 0) { p_ += 4; v.kind = Json::Bool; v.b = true; }
    else if (t_.compare(p_, 5, "false") == 0) { p_ += 5; v.kind = Json::Bool; v.b = false; }
    else if (t_.compare(p_, 4, "null") == 0) { p_ += 4; }
    else v = number();
    return v;
  }
  std::string string() {
    std::string out;
    p_++; // opening quote
    while (p_ < t_.size() && t_[p_] != '"') {
      char c = t_[p_++];
      if (c == '\\') {
        if (p_ >= t_.size()) fail("bad escape");
        const char e = t_[p_++];
        switch (e) {
          case 'n': out += '\n'; break; case 't': out += '\t'; break; case 'r': out += '\r'; break;
          case 'b': out += '\b'; break; case 'f': out += '\f'; break;
          case 'u': { // BMP code point -> UTF-8
            if (p_ + 4 > t_.size()) fail("bad \\u escape");
            const unsigned cp = (unsigned)std::strtoul(t_.substr(p_, 4).c_str(), nullptr, 16); p_ += 4;
            if (cp < 0x80) out += (char)cp;
            else if (cp < 0x800) { out += (char)(0xC0 | (cp >> 6)); out += (char)(0x80 | (cp & 0x3F)); }
            else { out += (char)(0xE0 | (cp >> 12)); out += (char)(0x80 | ((cp >> 6) & 0x3F)); out += (char)(0x80 | (cp & 0x3F)); }
            break;
          }
          default: out += e; // \" \\ \/
        }
      } else out += c;
    }
    if (p_ >= t_.size()) fail("unterminated string");
    p_++;
    return out;
  }
  Json number() {
    const size_t start = p_;
    bool is_float = false;
    if (p_ < t_.size() && (t_[p_] == '-' || t_[p_] == '+')) p_++;
    while (p_ < t_.size()) {
      const char c = t_[p_];
      if (c >= '0' && c <= '9') p_++;
      else if (c == '.' || c == 'e' || c == 'E' || c == '+' || c == '-') { is_float = true; p_++; }
      else break;
    }
    if (p_ == start) fail("unexpected character");
    const std::string tok = t_.substr(start, p_ - start);
    Json v;
    if (is_float) { v.kind = Json::Float; v.f = std::strtod(tok.c_str(), nullptr); }
    else { v.kind = Json::Int; v.i = std::strtoll(tok.c_str(), nullptr, 10); v.f = (double)v.i; }
    return v;
  }
};

} // namespace spt_host
