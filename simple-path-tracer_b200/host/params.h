// params.h — InputParams of the reference (src/core/loader.rs:12-17, 269-470): typed getters over one JSON object that
// remember which keys were visited (unused keys are reported, and drive TexInputModifier::load_with_tex).
#pragma once
#include <array>
#include <set>
#include <string>
#include <vector>

#include "host.h"
#include "json.h"

namespace spt_host {

inline std::string dir_of(const std::string& path) {
  const size_t p = path.find_last_of('/');
  return p == std::string::npos ? std::string() : path.substr(0, p);
}
inline std::string join_path(const std::string& dir, const std::string& rel) {
  if (!rel.empty() && rel[0] == '/') return rel;
  return dir.empty() ? rel : dir + "/" + rel;
}

class Params {
 public:
  std::string name;
  Params(const Json& obj, std::string name_, std::string base_path) : name(std::move(name_)), p_(obj), base_(std::move(base_path)) {
    if (obj.kind != Json::Object) throw HostError(name + " - expected an object");
  }

  bool contains_key(const std::string& k) const { return p_.find(k) != nullptr; }

  // core/loader.rs:422-427: an integer literal is an Int, everything else numeric a Float — and get_float does NOT
  // accept an Int (Q23).
  float get_float(const std::string& k) { return (float)get(k, Json::Float, "float").f; }
  float get_float_or(const std::string& k, float fb) { try { return get_float(k); } catch (const HostError&) { return fb; } }
  int64_t get_int(const std::string& k) { return get(k, Json::Int, "integer").i; }
  bool get_bool_or(const std::string& k, bool fb) { try { return get(k, Json::Bool, "boolean").b; } catch (const HostError&) { return fb; } }
  std::string get_str(const std::string& k) { return get(k, Json::String, "string").s; }
  const Json& get_any(const std::string& k) {
    const Json* v = p_.find(k);
    if (!v) throw HostError(name + " - there is no '" + k + "' field");
    visited_.insert(k);
    return *v;
  }

  std::array<float, 3> get_float3(const std::string& k) { auto v = get_vec(k, 3); return {v[0], v[1], v[2]}; }
  std::array<float, 3> get_float3_or(const std::string& k, std::array<float, 3> fb) { try { return get_float3(k); } catch (const HostError&) { return fb; } }
  std::array<float, 2> get_float2_or(const std::string& k, std::array<float, 2> fb) {
    try { auto v = get_vec(k, 2); return {v[0], v[1]}; } catch (const HostError&) { return fb; }
  }
  // core/loader.rs:307-335: 16 floats, column-major; entries that are not floats keep the identity's value.
  // Returned as m[row][col].
  std::array<std::array<float, 4>, 4> get_matrix(const std::string& k) {
    const Json& a = get(k, Json::Array, "an array");
    if (a.arr.size() != 16) throw HostError(name + " - '" + k + "' should be an array of 16 floats");
    std::array<std::array<float, 4>, 4> m{};
    for (int i = 0; i < 4; i++) m[i][i] = 1.f;
    for (int i = 0; i < 16; i++) if (a.arr[i].is_float()) m[i % 4][i / 4] = (float)a.arr[i].f;
    return m;
  }
  std::string get_file_path(const std::string& k) { return join_path(dir_of(base_), get_str(k)); }

  std::vector<std::string> unused_keys() const {
    std::vector<std::string> out;
    for (const auto& kv : p_.obj) if ((kv.first.empty() || kv.first[0] != '#') && !visited_.count(kv.first)) out.push_back(kv.first);
    return out;
  }

 private:
  const Json& p_;
  std::string base_;
  std::set<std::string> visited_;

  const Json& get(const std::string& k, Json::Kind kind, const char* hint) {
    const Json* v = p_.find(k);
    if (!v) throw HostError(name + " - there is no '" + k + "' field");
    if (v->kind != kind) throw HostError(name + " - '" + k + "' should be " + hint);
    visited_.insert(k);
    return *v;
  }
  std::vector<float> get_vec(const std::string& k, size_t n) {
    const Json* v = p_.find(k);
    if (!v) throw HostError(name + " - there is no '" + k + "' field");
    bool ok = v->kind == Json::Array && v->arr.size() == n;
    if (ok) for (const auto& e : v->arr) ok = ok && e.is_float();
    if (!ok) throw HostError(name + " - '" + k + "' should be array with " + std::to_string(n) + " floats");
    visited_.insert(k);
    std::vector<float> out;
    for (const auto& e : v->arr) out.push_back((float)e.f);
    return out;
  }
};

} // namespace spt_host
