// meshio.cpp — mesh ingestion of the C++ host: OBJ with the semantics of tobj::load_obj(triangulate, single_index) as
// TriMesh::load uses it (src/primitive/triangle.rs:57-108; the index-offset quirk Q22 is fixed, see below), TriMesh::calc_tangents
// (triangle.rs:339-388), and .npz meshes (positions / normals / texcoords / indices) as written by
// tools/import_reference_assets.py.
#include "host.h"
#include <zlib.h>
#include <cmath>
#include <cstring>
#include <fstream>
#include <map>
#include <sstream>
#include <tuple>

namespace spt_host {
namespace {

std::string read_file(const std::string& path) {
  std::ifstream fh(path, std::ios::binary);
  if (!fh) throw HostError("can't open '" + path + "'");
  std::ostringstream ss; ss << fh.rdbuf();
  return ss.str();
}

// ---- .npz: a zip archive of .npy members (stored or deflated) ------------------------------------------------------
struct NpyArray { std::string descr; std::vector<size_t> shape; std::string data; };

uint32_t rd32(const std::string& s, size_t o) { uint32_t v; memcpy(&v, s.data() + o, 4); return v; }
uint16_t rd16(const std::string& s, size_t o) { uint16_t v; memcpy(&v, s.data() + o, 2); return v; }

NpyArray parse_npy(const std::string& raw) {
  if (raw.size() < 10 || raw.compare(0, 6, "\x93NUMPY") != 0) throw HostError("npz member is not an .npy array");
  const int major = (unsigned char)raw[6];
  size_t hlen, hoff;
  if (major == 1) { hlen = rd16(raw, 8); hoff = 10; } else { hlen = rd32(raw, 8); hoff = 12; }
  const std::string hdr = raw.substr(hoff, hlen);
  NpyArray a;
  size_t p = hdr.find("'descr':");
  if (p == std::string::npos) throw HostError("npy header without descr");
  size_t q1 = hdr.find('\'', p + 8), q2 = hdr.find('\'', q1 + 1);
  a.descr = hdr.substr(q1 + 1, q2 - q1 - 1);
  if (hdr.find("'fortran_order': False") == std::string::npos) throw HostError("npy: fortran order is not supported");
  p = hdr.find("'shape':");
  q1 = hdr.find('(', p); q2 = hdr.find(')', q1);
  std::stringstream ss(hdr.substr(q1 + 1, q2 - q1 - 1));
  std::string tok;
  while (std::getline(ss, tok, ',')) { if (tok.find_first_of("0123456789") != std::string::npos) a.shape.push_back((size_t)std::stoull(tok)); }
  a.data = raw.substr(hoff + hlen);
  return a;
}

std::map<std::string, NpyArray> read_npz(const std::string& path) {
  const std::string z = read_file(path);
  // end of central directory
  size_t eocd = std::string::npos;
  for (size_t i = z.size() >= 22 ? z.size() - 22 : 0;; i--) { if (rd32(z, i) == 0x06054b50u) { eocd = i; break; } if (i == 0) break; }
  if (eocd == std::string::npos) throw HostError("'" + path + "' is not a zip archive");
  const size_t n = rd16(z, eocd + 10);
  size_t cd = rd32(z, eocd + 16);
  std::map<std::string, NpyArray> out;
  for (size_t e = 0; e < n; e++) {
    if (rd32(z, cd) != 0x02014b50u) throw HostError("npz: bad central directory");
    const int method = rd16(z, cd + 10);
    const size_t csize = rd32(z, cd + 20), usize = rd32(z, cd + 24), nlen = rd16(z, cd + 28), xlen = rd16(z, cd + 30), clen = rd16(z, cd + 32);
    const size_t lho = rd32(z, cd + 42);
    const std::string name = z.substr(cd + 46, nlen);
    const size_t data = lho + 30 + rd16(z, lho + 26) + rd16(z, lho + 28);
    std::string raw;
    if (method == 0) raw = z.substr(data, csize);
    else if (method == 8) {
      raw.resize(usize);
      z_stream zs; memset(&zs, 0, sizeof(zs));
      if (inflateInit2(&zs, -MAX_WBITS) != Z_OK) throw HostError("zlib init failed");
      zs.next_in = (Bytef*)(z.data() + data); zs.avail_in = (uInt)csize;
      zs.next_out = (Bytef*)&raw[0]; zs.avail_out = (uInt)usize;
      const int rc = inflate(&zs, Z_FINISH);
      inflateEnd(&zs);
      if (rc != Z_STREAM_END) throw HostError("npz: inflate failed for " + name);
    } else throw HostError("npz: unsupported compression method");
    std::string key = name;
    if (key.size() > 4 && key.compare(key.size() - 4, 4, ".npy") == 0) key.resize(key.size() - 4);
    out[key] = parse_npy(raw);
    cd += 46 + nlen + xlen + clen;
  }
  return out;
}

template <typename T>
std::vector<T> as_vec(const NpyArray& a, const char* descr) {
  if (a.descr != descr) throw HostError(std::string("npz: expected dtype ") + descr + ", found " + a.descr);
  std::vector<T> v(a.data.size() / sizeof(T));
  memcpy(v.data(), a.data.data(), v.size() * sizeof(T));
  return v;
}

Mesh load_npz(const std::string& path) {
  auto z = read_npz(path);
  if (!z.count("positions") || !z.count("indices")) throw HostError("npz mesh without positions / indices: " + path);
  Mesh m;
  m.positions = as_vec<float>(z["positions"], "<f4");
  const size_t n = m.positions.size() / 3;
  if (z.count("normals")) m.normals = as_vec<float>(z["normals"], "<f4");
  else { m.normals.resize(3 * n); for (size_t i = 0; i < n; i++) { m.normals[3 * i] = 0; m.normals[3 * i + 1] = 0; m.normals[3 * i + 2] = 1; } }
  if (z.count("texcoords")) m.texcoords = as_vec<float>(z["texcoords"], "<f4");
  else m.texcoords.assign(2 * n, 0.f);
  m.indices = as_vec<uint32_t>(z["indices"], "<u4");
  return m;
}

// ---- OBJ --------------------------------------------------------------------------------------------------------------
Mesh load_obj(const std::string& path) {
  std::ifstream fh(path);
  if (!fh) throw HostError("can't open '" + path + "'");
  std::vector<float> v, vt, vn;
  typedef std::tuple<int, int, int> Corner;
  std::vector<std::vector<std::vector<Corner>>> models(1);
  std::string line;
  while (std::getline(fh, line)) {
    std::istringstream ss(line);
    std::string k;
    if (!(ss >> k)) continue;
    if (k == "v") { float a, b, c; ss >> a >> b >> c; v.insert(v.end(), {a, b, c}); }
    else if (k == "vt") { float a = 0, b = 0; ss >> a >> b; vt.insert(vt.end(), {a, b}); }
    else if (k == "vn") { float a, b, c; ss >> a >> b >> c; vn.insert(vn.end(), {a, b, c}); }
    else if (k == "o" || k == "g") { if (!models.back().empty()) models.emplace_back(); }
    else if (k == "f") {
      std::vector<Corner> face;
      std::string tok;
      while (ss >> tok) {
        int idx[3] = {0, 0, 0}; int part = 0; std::string cur;
        for (size_t i = 0; i <= tok.size(); i++) {
          if (i == tok.size() || tok[i] == '/') { if (!cur.empty() && part < 3) idx[part] = std::stoi(cur); cur.clear(); part++; }
          else cur += tok[i];
        }
        const int nv = (int)(v.size() / 3), nt = (int)(vt.size() / 2), nn = (int)(vn.size() / 3);
        const int vi = idx[0] > 0 ? idx[0] - 1 : nv + idx[0];
        const int ti = idx[1] != 0 ? (idx[1] > 0 ? idx[1] - 1 : nt + idx[1]) : -1;
        const int ni = idx[2] != 0 ? (idx[2] > 0 ? idx[2] - 1 : nn + idx[2]) : -1;
        face.emplace_back(vi, ti, ni);
      }
      models.back().push_back(face);
    }
  }
  Mesh out;
  uint32_t n_vertices = 0;
  for (const auto& faces : models) {
    if (faces.empty()) continue;
    bool has_n = true, has_t = true;
    for (const auto& f : faces) for (const auto& c : f) { has_n = has_n && std::get<2>(c) >= 0; has_t = has_t && std::get<1>(c) >= 0; }
    std::map<Corner, uint32_t> remap;
    std::vector<float> mp, mn, mt; std::vector<uint32_t> mi;
    for (const auto& f : faces) {
      std::vector<uint32_t> corners;
      for (const auto& c : f) {
        auto it = remap.find(c);
        if (it == remap.end()) {
          it = remap.emplace(c, (uint32_t)(mp.size() / 3)).first;
          const int vi = std::get<0>(c);
          mp.insert(mp.end(), {v[3 * vi], v[3 * vi + 1], v[3 * vi + 2]});
          if (has_n) { const int ni = std::get<2>(c); mn.insert(mn.end(), {vn[3 * ni], vn[3 * ni + 1], vn[3 * ni + 2]}); }
          else mn.insert(mn.end(), {0.f, 0.f, 1.f}); // MeshVertex::default, triangle.rs:29-39
          if (has_t) { const int ti = std::get<1>(c); mt.insert(mt.end(), {vt[2 * ti], vt[2 * ti + 1]}); }
          else mt.insert(mt.end(), {0.f, 0.f});
        }
        corners.push_back(it->second);
      }
      for (size_t k = 1; k + 1 < corners.size(); k++) mi.insert(mi.end(), {corners[0], corners[k], corners[k + 1]}); // fan triangulation
    }
    out.positions.insert(out.positions.end(), mp.begin(), mp.end());
    out.normals.insert(out.normals.end(), mn.begin(), mn.end());
    out.texcoords.insert(out.texcoords.end(), mt.begin(), mt.end());
    // Q22 FIXED: the reference offsets by the number of INDICES loaded so far (triangle.rs:96-102), which panics or wires triangles to
    // the wrong vertices for any OBJ with more than one object; the evident intent is the running vertex count
    for (uint32_t i : mi) out.indices.push_back(i + n_vertices);
    n_vertices += (uint32_t)(mp.size() / 3);
  }
  for (uint32_t i : out.indices) if ((size_t)i >= out.positions.size() / 3) throw HostError("'" + path + "': vertex index out of range");
  return out;
}

void normalize3(float* a) {
  const float ln = std::sqrt((a[0] * a[0] + a[1] * a[1]) + a[2] * a[2]);
  a[0] /= ln; a[1] /= ln; a[2] /= ln;
}

} // namespace

// .npy file holding an f32 (or f64) array; used by the "npy" environment extension.
std::vector<float> load_npy_f32(const std::string& path, std::vector<size_t>& shape) {
  const NpyArray a = parse_npy(read_file(path));
  shape = a.shape;
  if (a.descr == "<f4") return as_vec<float>(a, "<f4");
  const std::vector<double> d = as_vec<double>(a, "<f8");
  return std::vector<float>(d.begin(), d.end());
}

// TriMesh::calc_tangents, triangle.rs:339-388.  Vertices no UV-nondegenerate triangle touches keep the MeshVertex
// defaults tangent = X, bitangent = Y.
void calc_tangents(Mesh& m) {
  const size_t n = m.positions.size() / 3, nt = m.indices.size() / 3;
  m.tangents.resize(3 * n); m.bitangents.resize(3 * n);
  for (size_t i = 0; i < n; i++) { m.tangents[3 * i] = 1; m.tangents[3 * i + 1] = 0; m.tangents[3 * i + 2] = 0; m.bitangents[3 * i] = 0; m.bitangents[3 * i + 1] = 1; m.bitangents[3 * i + 2] = 0; }
  std::vector<float> tsum(3 * n, 0.f), bsum(3 * n, 0.f);
  std::vector<int> deg(n, 0);
  for (size_t t = 0; t < nt; t++) {
    const uint32_t i0 = m.indices[3 * t], i1 = m.indices[3 * t + 1], i2 = m.indices[3 * t + 2];
    const float* p0 = &m.positions[3 * i0]; const float* p1 = &m.positions[3 * i1]; const float* p2 = &m.positions[3 * i2];
    const float* w0 = &m.texcoords[2 * i0]; const float* w1 = &m.texcoords[2 * i1]; const float* w2 = &m.texcoords[2 * i2];
    const float e1[3] = {p1[0] - p0[0], p1[1] - p0[1], p1[2] - p0[2]}, e2[3] = {p2[0] - p0[0], p2[1] - p0[1], p2[2] - p0[2]};
    const float u1[2] = {w1[0] - w0[0], w1[1] - w0[1]}, u2[2] = {w2[0] - w0[0], w2[1] - w0[1]};
    const float det = u1[0] * u2[1] - u1[1] * u2[0];
    if (det == 0.f) continue;
    const float inv = 1.0f / det;
    float tv[3], bv[3];
    for (int k = 0; k < 3; k++) { tv[k] = (e1[k] * u2[1] - e2[k] * u1[1]) * inv; bv[k] = (e2[k] * u1[0] - e1[k] * u2[0]) * inv; }
    normalize3(tv); normalize3(bv);
    for (uint32_t vi : {i0, i1, i2}) { for (int k = 0; k < 3; k++) { tsum[3 * vi + k] += tv[k]; bsum[3 * vi + k] += bv[k]; } deg[vi]++; }
  }
  for (size_t i = 0; i < n; i++) if (deg[i]) {
    const float inv_deg = 1.0f / (float)deg[i];
    float tv[3] = {tsum[3 * i] * inv_deg, tsum[3 * i + 1] * inv_deg, tsum[3 * i + 2] * inv_deg};
    float bv[3] = {bsum[3 * i] * inv_deg, bsum[3 * i + 1] * inv_deg, bsum[3 * i + 2] * inv_deg};
    normalize3(tv); normalize3(bv);
    for (int k = 0; k < 3; k++) { m.tangents[3 * i + k] = tv[k]; m.bitangents[3 * i + k] = bv[k]; }
  }
}

Mesh load_mesh(const std::string& path) {
  Mesh m = path.size() > 4 && path.compare(path.size() - 4, 4, ".npz") == 0 ? load_npz(path) : load_obj(path);
  calc_tangents(m);
  return m;
}

} // namespace spt_host
