// flatten.cpp — SptSceneDesc -> device-layout arrays.  Runs once per scene on the host and replaces the
// constructors the reference runs while loading: Instance::new (instance.rs:19-29), TriMesh::new + BvhAccel::new
// (triangle.rs:42-55, bvh.rs:23-141; tree rebuilt as a compressed wide BVH), generate_mipmap (image_tex.rs:63-98),
// EnvLight::new (environment.rs:18-50), AliasTable::new (alias_table.rs:8-56), PowerIsLightSampler::new
// (power_is.rs:19-42) with LightT::power (shape_light.rs:78-81, sphere.rs:122-135, triangle.rs:282-288).
#include "flatten.h"
#include <cmath>
#include <cstring>
#include <cfloat>
#include <chrono>
#include <thread>
#include <atomic>
#include <functional>
#include <string>
#include <cstdlib>
#include <algorithm>

namespace spt {
namespace {

struct V3 { float x, y, z; };
inline V3 operator+(V3 a, V3 b) { return {a.x + b.x, a.y + b.y, a.z + b.z}; }
inline V3 operator-(V3 a, V3 b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
inline V3 operator*(V3 a, float s) { return {a.x * s, a.y * s, a.z * s}; }
inline V3 neg(V3 a) { return {-a.x, -a.y, -a.z}; }
inline float dot(V3 a, V3 b) { return (a.x * b.x + a.y * b.y) + a.z * b.z; }
inline V3 cross(V3 a, V3 b) { return {a.y * b.z - b.y * a.z, a.z * b.x - b.z * a.x, a.x * b.y - b.x * a.y}; }
inline float length(V3 a) { return std::sqrt(dot(a, a)); }
struct M3 { V3 cx, cy, cz; };                        // column vectors (glam Mat3A)
inline V3 mul(const M3& m, V3 v) { return (m.cx * v.x + m.cy * v.y) + m.cz * v.z; }
inline M3 mul(const M3& a, const M3& b) { return {mul(a, b.cx), mul(a, b.cy), mul(a, b.cz)}; }
inline M3 transpose(const M3& m) { return {{m.cx.x, m.cy.x, m.cz.x}, {m.cx.y, m.cy.y, m.cz.y}, {m.cx.z, m.cy.z, m.cz.z}}; }
inline M3 inverse(const M3& m) { // glam Mat3A::inverse
  V3 t0 = cross(m.cy, m.cz), t1 = cross(m.cz, m.cx), t2 = cross(m.cx, m.cy);
  float inv = 1.0f / dot(m.cz, t2);
  return transpose(M3{t0 * inv, t1 * inv, t2 * inv});
}
struct A3 { M3 m; V3 t; };
inline A3 inverse(const A3& a) { M3 mi = inverse(a.m); return {mi, neg(mul(mi, a.t))}; }
inline V3 point(const A3& a, V3 p) { return mul(a.m, p) + a.t; }
inline float lum(const float* c) { return 0.299f * c[0] + 0.587f * c[1] + 0.114f * c[2]; }
inline float4 row(const M3& m, int i, float w) {
  const float* x = &m.cx.x; const float* y = &m.cy.x; const float* z = &m.cz.x;
  return make_float4(x[i], y[i], z[i], w);
}

// AliasTable::new, alias_table.rs:8-56 (the author's own pairing, not Vose); returns (u, k)
void alias_build(const std::vector<float>& props, std::vector<float2>& out) {
  const size_t n = props.size();
  std::vector<float> u(n); std::vector<size_t> k(n);
  for (size_t i = 0; i < n; i++) { u[i] = props[i] * (float)n; k[i] = i; }
  long poor = -1, rich = -1;
  for (size_t i = 0; i < n; i++) if (u[i] < 1.0f) { poor = (long)i; break; }
  long poor_max = poor;
  for (size_t i = 0; i < n; i++) if (u[i] > 1.0f) { rich = (long)i; break; }
  while (poor >= 0 && rich >= 0) {
    u[rich] -= 1.0f - u[poor];
    k[poor] = (size_t)rich;
    if (u[rich] < 1.0f && rich < poor_max) poor = rich;
    else {
      poor = -1;
      for (size_t i = (size_t)poor_max + 1; i < n; i++) if (u[i] < 1.0f) { poor = (long)i; poor_max = (long)i; break; }
    }
    const long r0 = rich; rich = -1;
    for (size_t i = (size_t)r0; i < n; i++) if (u[i] > 1.0f) { rich = (long)i; break; }
  }
  out.resize(n);
  for (size_t i = 0; i < n; i++) { int32_t ki = (int32_t)k[i]; float kf; memcpy(&kf, &ki, 4); out[i] = make_float2(u[i], kf); }
}

inline float srgb_lin(float s) { return s <= 0.04045f ? s / 12.92f : std::pow((s + 0.055f) / 1.055f, 2.4f); }

// TextureT::average_color (scalar.rs:35-37, image_tex.rs:41-44, binary_op.rs:59-61, srgb_tex.rs:30-32, input_modifier.rs:78-80)
void tex_average(const HostScene& hs, int id, float out[3]) {
  const DTexture& t = hs.textures[id];
  float a[3], b[3];
  switch (t.tag) {
    case SPT_TEX_SCALAR: out[0] = t.value[0]; out[1] = t.value[1]; out[2] = t.value[2]; return;
    case SPT_TEX_IMAGE: {
      const DMip& m = hs.mips[t.mip_base + t.n_mips - 1];
      const uchar4 p = hs.texels[m.texel_base];
      out[0] = p.x / 255.0f; out[1] = p.y / 255.0f; out[2] = p.z / 255.0f; return;
    }
    case SPT_TEX_ADD: case SPT_TEX_SUB: case SPT_TEX_MUL: case SPT_TEX_DIV:
      tex_average(hs, t.child0, a); tex_average(hs, t.child1, b);
      for (int c = 0; c < 3; c++) out[c] = t.tag == SPT_TEX_ADD ? a[c] + b[c] : t.tag == SPT_TEX_SUB ? a[c] - b[c] : t.tag == SPT_TEX_MUL ? a[c] * b[c] : a[c] / b[c];
      return;
    case SPT_TEX_SRGB: tex_average(hs, t.child0, a); for (int c = 0; c < 3; c++) out[c] = srgb_lin(a[c]); return;
    default: tex_average(hs, t.child0, out); return;
  }
}

} // namespace

// fn(begin, end) over [0, n) on several host threads: the per-vertex / per-triangle gathers of a big mesh are plain copies
// into disjoint ranges (10 M triangles: 2.4 GB written), so the result does not depend on the split.
static void parallel_ranges(size_t n, const std::function<void(size_t, size_t)>& fn) {
  const size_t min_per_thread = (size_t)1 << 16;
  size_t nt = std::min<size_t>(std::max(1u, std::thread::hardware_concurrency()), 32);
  nt = std::min(nt, n / min_per_thread);
  if (nt <= 1) { fn(0, n); return; }
  std::vector<std::thread> th;
  const size_t per = (n + nt - 1) / nt;
  for (size_t t = 0; t < nt; t++) { const size_t b = t * per, e = std::min(n, b + per); if (b < e) th.emplace_back(fn, b, e); }
  for (auto& x : th) x.join();
}


std::string flatten_scene(const SptSceneDesc& d, HostScene& hs) {
  auto bad = [](const char* m) { return std::string(m); };
  if (d.n_cameras <= 0) return bad("At least one camera is needed");
  // ---- textures (+ mip pyramids) -----------------------------------------------------------------------------
  hs.textures.resize(d.n_textures);
  std::vector<int> tex_depth(std::max(1, d.n_textures), 0); // nesting of binary operators (device limit: 3)
  for (int i = 0; i < d.n_textures; i++) {
    const SptTextureDesc& t = d.textures[i];
    DTexture o; memset(&o, 0, sizeof(o));
    o.tag = t.tag; o.child0 = t.child[0]; o.child1 = t.child[1]; o.mode = t.mode; o.wrap = t.wrap;
    for (int c = 0; c < 3; c++) { o.value[c] = t.value[c]; o.tiling[c] = t.tiling[c]; o.offset[c] = t.offset[c]; }
    if (t.tag < 0 || t.tag > SPT_TEX_INPUT_MODIFIER) return bad("texture: unknown tag");
    if (t.tag >= SPT_TEX_ADD && (t.child[0] < 0 || t.child[0] >= i)) return bad("texture: child must precede its parent");
    if (t.tag >= SPT_TEX_ADD && t.tag <= SPT_TEX_DIV && (t.child[1] < 0 || t.child[1] >= i)) return bad("texture: child must precede its parent");
    if (t.tag == SPT_TEX_IMAGE) {
      if (!t.rgba || t.width <= 0 || t.height <= 0) return bad("image texture without pixels");
      hs.has_image_tex = true;
      o.mip_base = (int32_t)hs.mips.size();
      int w = t.width, h = t.height;
      std::vector<uchar4> cur((size_t)w * h);
      memcpy(cur.data(), t.rgba, (size_t)4 * w * h);
      for (;;) {
        DMip m; m.w = w; m.h = h; m.texel_base = (uint32_t)hs.texels.size(); m.pad = 0;
        hs.mips.push_back(m);
        hs.texels.insert(hs.texels.end(), cur.begin(), cur.end());
        o.n_mips++;
        if (!(w > 1 || h > 1)) break;
        const int nw = (w + 1) >> 1, nh = (h + 1) >> 1; // generate_mipmap: box filter through u8 truncation
        std::vector<uchar4> nxt((size_t)nw * nh);
        for (int x = 0; x < nw; x++) for (int y = 0; y < nh; y++) {
          const int x0 = 2 * x, x1 = std::min(2 * x + 1, w - 1), y0 = 2 * y, y1 = std::min(2 * y + 1, h - 1);
          const uchar4 p0 = cur[(size_t)y0 * w + x0], p1 = cur[(size_t)y1 * w + x0], p2 = cur[(size_t)y0 * w + x1], p3 = cur[(size_t)y1 * w + x1];
          uchar4 q;
          q.x = (unsigned char)(((float)p0.x + (float)p1.x + (float)p2.x + (float)p3.x) * 0.25f);
          q.y = (unsigned char)(((float)p0.y + (float)p1.y + (float)p2.y + (float)p3.y) * 0.25f);
          q.z = (unsigned char)(((float)p0.z + (float)p1.z + (float)p2.z + (float)p3.z) * 0.25f);
          q.w = (unsigned char)(((float)p0.w + (float)p1.w + (float)p2.w + (float)p3.w) * 0.25f);
          nxt[(size_t)y * nw + x] = q;
        }
        cur.swap(nxt); w = nw; h = nh;
      }
    }
    if (t.tag >= SPT_TEX_ADD && t.tag <= SPT_TEX_DIV) tex_depth[i] = 1 + std::max(tex_depth[t.child[0]], tex_depth[t.child[1]]);
    else if (t.tag >= SPT_TEX_SRGB) tex_depth[i] = tex_depth[t.child[0]];
    if (tex_depth[i] > 3) return bad("texture: more than 3 nested binary operators are not supported");
    hs.textures[i] = o;
  }
  auto tex_ok = [&](int id) { return id >= 0 && id < d.n_textures; };
  // ---- materials / mediums / surfaces ---------------------------------------------------------------------------
  hs.materials.resize(d.n_materials);
  for (int i = 0; i < d.n_materials; i++) {
    const SptMaterialDesc& m = d.materials[i];
    if (m.tag < 0 || m.tag > SPT_MAT_SUBSURFACE) return bad("material: unknown tag");
    DMaterial o; o.tag = m.tag; o.ior = m.ior; o.rough_chan = m.rough_chan; o.metal_chan = m.metal_chan;
    for (int k = 0; k < 4; k++) o.tex[k] = m.tex[k];
    const bool need[8][4] = {{1, 0, 0, 0}, {0, 0, 0, 0}, {1, 1, 1, 1}, {0, 0, 1, 1}, {1, 0, 1, 1}, {1, 1, 1, 1}, {1, 1, 1, 1}, {1, 1, 1, 1}};
    for (int k = 0; k < 4; k++) if (need[m.tag][k] && !tex_ok(m.tex[k])) return bad("material: bad texture index");
    if (m.rough_chan < 0 || m.rough_chan > 3 || m.metal_chan < 0 || m.metal_chan > 3) return bad("material: rough_chan / metal_chan must be a texel channel 0..3");
    hs.materials[i] = o;
  }
  hs.mediums.resize(d.n_mediums);
  for (int i = 0; i < d.n_mediums; i++) { // Homogeneous::new, homogeneous.rs:12-19
    DMedium o; memset(&o, 0, sizeof(o));
    for (int c = 0; c < 3; c++) { o.sigma_t[c] = d.mediums[i].sigma_a[c] + d.mediums[i].sigma_s[c]; o.sigma_s[c] = d.mediums[i].sigma_s[c]; }
    o.g = d.mediums[i].asymmetric;
    hs.mediums[i] = o;
  }
  hs.surfaces.resize(d.n_surfaces);
  for (int i = 0; i < d.n_surfaces; i++) {
    const SptSurfaceDesc& s = d.surfaces[i];
    if (s.material < 0 || s.material >= d.n_materials) return bad("surface: bad material index");
    if ((s.normal_map >= 0 && !tex_ok(s.normal_map)) || (s.emissive_map >= 0 && !tex_ok(s.emissive_map))) return bad("surface: bad texture index");
    if (s.inside_medium >= d.n_mediums) return bad("surface: bad medium index");
    DSurface o; o.material = s.material; o.normal_map = s.normal_map; o.emissive_map = s.emissive_map;
    o.double_sided = s.double_sided; o.inside_medium = s.inside_medium;
    for (int c = 0; c < 3; c++) o.emissive[c] = s.emissive[c];
    hs.surfaces[i] = o;
  }
  hs.lean = true;
  for (int i = 0; i < d.n_textures; i++) hs.lean = hs.lean && d.textures[i].tag == SPT_TEX_SCALAR;
  for (int i = 0; i < d.n_surfaces; i++) hs.lean = hs.lean && d.surfaces[i].normal_map < 0 && d.surfaces[i].emissive_map < 0;
  // ---- spheres, meshes (+ BLAS) ------------------------------------------------------------------------------------
  hs.spheres.resize(d.n_spheres);
  for (int i = 0; i < d.n_spheres; i++) hs.spheres[i] = make_float4(d.spheres[i].center[0], d.spheres[i].center[1], d.spheres[i].center[2], d.spheres[i].radius);
  // reserve node 0.. for the TLAS: build BLASes into a temporary list first, then lay out TLAS | BLAS...
  std::vector<std::vector<CwNode>> blas_nodes(d.n_meshes);
  std::vector<Box3> mesh_box(d.n_meshes);
  hs.meshes.resize(d.n_meshes);
  std::vector<std::vector<int>> blas_order(d.n_meshes);
  // triangles per BLAS leaf (1..3).  Measured on C2 / C3 / C5: 2 beats 3 by 1.6-2.4 % whole-render (fewer triangle tests
  // per ray outweigh the extra nodes); 1 equals 2 with more node memory.  SPT_BLAS_LEAF overrides it for experiments.
  int blas_leaf = 2;
  if (const char* e = getenv("SPT_BLAS_LEAF")) blas_leaf = std::max(1, std::min(3, atoi(e)));
  // BLAS builder: meshes of >= SPT_DEVICE_BUILD_MIN_TRIS triangles are built on the GPU (LBVH, ~100x faster to build, a few
  // per cent slower to trace), smaller ones with the host SAH builder.  SPT_BVH_BUILD=host|device forces one of them.
  int bvh_mode = 1, bvh_algo = 1; // GPU builder: PLOC (near-SAH quality); SPT_BVH_BUILD=device_lbvh picks the plain radix tree
  if (const char* e = getenv("SPT_BVH_BUILD")) {
    bvh_mode = !strcmp(e, "host") ? 0 : (!strncmp(e, "device", 6) ? 2 : 1);
    if (!strcmp(e, "device_lbvh")) bvh_algo = 0;
  }
  if (const char* e = getenv("SPT_BVH_ALGO")) bvh_algo = !strcmp(e, "lbvh") ? 0 : 1;
  for (int mi = 0; mi < d.n_meshes; mi++) {
    const SptMeshDesc& m = d.meshes[mi];
    if (m.n_tris <= 0 || m.n_verts <= 0 || !m.position || !m.index) return bad("mesh: empty");
    DMesh dm; dm.vert_base = (int32_t)(hs.verts.size() / 4); dm.idx_base = (int32_t)(hs.indices.size() / 3); dm.n_tris = m.n_tris; dm.blas_root = 0;
    {
      const size_t v0 = hs.verts.size();
      hs.verts.resize(v0 + 4 * (size_t)m.n_verts);
      float4* out = hs.verts.data() + v0;
      parallel_ranges((size_t)m.n_verts, [&](size_t vb, size_t ve) {
        const float nz[3] = {0, 0, 1}, tx[3] = {1, 0, 0}, by[3] = {0, 1, 0}, uv0[2] = {0, 0};
        for (size_t v = vb; v < ve; v++) {
          const float* p = m.position + 3 * v;
          const float* n = m.normal ? m.normal + 3 * v : nz; const float* uv = m.texcoord ? m.texcoord + 2 * v : uv0;
          const float* t = m.tangent ? m.tangent + 3 * v : tx; const float* b = m.bitangent ? m.bitangent + 3 * v : by;
          out[4 * v + 0] = make_float4(p[0], p[1], p[2], uv[0]);
          out[4 * v + 1] = make_float4(n[0], n[1], n[2], uv[1]);
          out[4 * v + 2] = make_float4(t[0], t[1], t[2], 0.f);
          out[4 * v + 3] = make_float4(b[0], b[1], b[2], 0.f);
        }
      });
    }
    std::vector<Box3> boxes(m.n_tris);
    {
      const size_t i0 = hs.indices.size();
      hs.indices.resize(i0 + 3 * (size_t)m.n_tris);
      uint32_t* out = hs.indices.data() + i0;
      std::atomic<bool> bad_index(false);
      parallel_ranges((size_t)m.n_tris, [&](size_t tb, size_t te) {
        for (size_t t = tb; t < te; t++) {
          Box3 bx; for (int k = 0; k < 3; k++) { bx.lo[k] = FLT_MAX; bx.hi[k] = -FLT_MAX; }
          for (int c = 0; c < 3; c++) {
            const uint32_t vi = m.index[3 * t + c];
            if (vi >= (uint32_t)m.n_verts) { bad_index = true; return; }
            out[3 * t + c] = vi;
            for (int k = 0; k < 3; k++) { bx.lo[k] = std::min(bx.lo[k], m.position[3 * vi + k]); bx.hi[k] = std::max(bx.hi[k], m.position[3 * vi + k]); }
          }
          boxes[t] = bx;
        }
      });
      if (bad_index) return bad("mesh: vertex index out of range");
    }
    hs.meshes[mi] = dm;
    // node/prim bases are patched below once the TLAS size is known: build with base 0 and offset afterwards
    const bool on_device = bvh_mode == 2 || (bvh_mode == 1 && m.n_tris >= SPT_DEVICE_BUILD_MIN_TRIS);
    const auto tb0 = std::chrono::steady_clock::now();
    if (on_device && m.n_tris >= 2) {
      std::string berr;
      if (!build_cwbvh_device(boxes, blas_leaf, blas_nodes[mi], blas_order[mi], &mesh_box[mi], &berr, bvh_algo)) return bad(berr.c_str());
      hs.blas_on_device++;
    } else build_cwbvh(boxes, blas_leaf, 0, 0, blas_nodes[mi], blas_order[mi], &mesh_box[mi]);
    hs.seconds_bvh += std::chrono::duration<double>(std::chrono::steady_clock::now() - tb0).count();
  }
  // ---- instances ------------------------------------------------------------------------------------------------------
  hs.inst.resize(d.n_instances); hs.inst_shade.resize(d.n_instances);
  std::vector<Box3> inst_box(d.n_instances);
  std::vector<A3> inst_fwd(d.n_instances);
  bool tlas_tight = true;
  if (const char* e = getenv("SPT_TLAS_TIGHT")) tlas_tight = atoi(e) != 0;
  for (int i = 0; i < d.n_instances; i++) {
    const SptInstanceDesc& in = d.instances[i];
    if (in.surface < 0 || in.surface >= d.n_surfaces) return bad("instance: bad surface index");
    const float* x = in.xform;
    A3 a{{{x[0], x[1], x[2]}, {x[3], x[4], x[5]}, {x[6], x[7], x[8]}}, {x[9], x[10], x[11]}};
    const A3 ai = inverse(a);
    const M3 it = transpose(ai.m); // Transform::new, transform.rs:18-22
    DInstance o;
    o.inv0 = row(ai.m, 0, ai.t.x); o.inv1 = row(ai.m, 1, ai.t.y); o.inv2 = row(ai.m, 2, ai.t.z);
    o.kind = in.prim_kind; o.surface = in.surface; o.mesh = -1; o.root = 0;
    Box3 pb;
    if (in.prim_kind == SPT_PRIM_SPHERE) {
      if (in.prim_index < 0 || in.prim_index >= d.n_spheres) return bad("instance: bad sphere index");
      o.root = in.prim_index;
      const float4 s = hs.spheres[in.prim_index];
      pb.lo[0] = s.x - s.w; pb.lo[1] = s.y - s.w; pb.lo[2] = s.z - s.w; pb.hi[0] = s.x + s.w; pb.hi[1] = s.y + s.w; pb.hi[2] = s.z + s.w;
    } else if (in.prim_kind == SPT_PRIM_TRIMESH) {
      if (in.prim_index < 0 || in.prim_index >= d.n_meshes) return bad("instance: bad mesh index");
      o.mesh = in.prim_index; pb = mesh_box[in.prim_index];
    } else return bad("instance: unknown primitive kind");
    Box3 wb; for (int k = 0; k < 3; k++) { wb.lo[k] = FLT_MAX; wb.hi[k] = -FLT_MAX; }
    for (int c = 0; c < 8; c++) { // Bbox::transformed_by, bbox.rs:41-61
      const V3 p = point(a, V3{(c & 4) ? pb.hi[0] : pb.lo[0], (c & 2) ? pb.hi[1] : pb.lo[1], (c & 1) ? pb.hi[2] : pb.lo[2]});
      wb.lo[0] = std::min(wb.lo[0], p.x); wb.lo[1] = std::min(wb.lo[1], p.y); wb.lo[2] = std::min(wb.lo[2], p.z);
      wb.hi[0] = std::max(wb.hi[0], p.x); wb.hi[1] = std::max(wb.hi[1], p.y); wb.hi[2] = std::max(wb.hi[2], p.z);
    }
    // The TLAS box only culls (hits come from the leaf tests), so it may be tighter than the reference's box of the transformed
    // object box: for a rotated mesh instance the box of its transformed VERTICES (a hit point lies in their convex hull), padded
    // outward by 1e-5 of its extent against rounding.  C5's bunnies, turned by arbitrary angles about y: fewer false instance visits.
    if (tlas_tight && in.prim_kind == SPT_PRIM_TRIMESH && d.meshes[in.prim_index].n_verts <= (1 << 20) &&
        (a.m.cx.y != 0.f || a.m.cx.z != 0.f || a.m.cy.x != 0.f || a.m.cy.z != 0.f || a.m.cz.x != 0.f || a.m.cz.y != 0.f)) {
      const SptMeshDesc& m = d.meshes[in.prim_index];
      Box3 tb; for (int k = 0; k < 3; k++) { tb.lo[k] = FLT_MAX; tb.hi[k] = -FLT_MAX; }
      for (int v = 0; v < m.n_verts; v++) {
        const V3 p = point(a, V3{m.position[3 * v], m.position[3 * v + 1], m.position[3 * v + 2]});
        tb.lo[0] = std::min(tb.lo[0], p.x); tb.lo[1] = std::min(tb.lo[1], p.y); tb.lo[2] = std::min(tb.lo[2], p.z);
        tb.hi[0] = std::max(tb.hi[0], p.x); tb.hi[1] = std::max(tb.hi[1], p.y); tb.hi[2] = std::max(tb.hi[2], p.z);
      }
      for (int k = 0; k < 3; k++) {
        const float pad = 1e-5f * (tb.hi[k] - tb.lo[k]) + 1e-6f * std::max(std::fabs(tb.lo[k]), std::fabs(tb.hi[k]));
        wb.lo[k] = std::max(wb.lo[k], tb.lo[k] - pad); wb.hi[k] = std::min(wb.hi[k], tb.hi[k] + pad); // never looser than the reference's box
      }
    }
    inst_box[i] = wb; inst_fwd[i] = a;
    DInstanceShade sh;
    sh.fwd0 = row(a.m, 0, a.t.x); sh.fwd1 = row(a.m, 1, a.t.y); sh.fwd2 = row(a.m, 2, a.t.z);
    sh.it0 = row(it, 0, 0.f); sh.it1 = row(it, 1, 0.f); sh.it2 = row(it, 2, 0.f);
    hs.inst[i] = o; hs.inst_shade[i] = sh;
  }
  // ---- TLAS, then lay the BLASes out behind it ---------------------------------------------------------------------------
  std::vector<uint32_t> mesh_tri_base(std::max(1, d.n_meshes), 0u);
  std::vector<CwNode> tlas; std::vector<int> tlas_order;
  const auto tt0 = std::chrono::steady_clock::now();
  build_cwbvh(inst_box, 1, 0, 0, tlas, tlas_order, nullptr);
  hs.seconds_bvh += std::chrono::duration<double>(std::chrono::steady_clock::now() - tt0).count();
  hs.nodes = tlas; hs.tlas_nodes = (int64_t)tlas.size();
  hs.tlas_inst.assign(tlas_order.begin(), tlas_order.end());
  for (int mi = 0; mi < d.n_meshes; mi++) {
    const uint32_t node_base = (uint32_t)hs.nodes.size(), tri_base = (uint32_t)(hs.tris.size() / 3);
    hs.meshes[mi].blas_root = (int32_t)node_base;
    for (CwNode n : blas_nodes[mi]) { n.child_base += node_base; n.prim_base += tri_base; hs.nodes.push_back(n); }
    const SptMeshDesc& m = d.meshes[mi];
    mesh_tri_base[mi] = tri_base;
    const std::vector<int>& order = blas_order[mi];
    const size_t t0 = hs.tris.size(), s0 = hs.shade_tris.size();
    hs.tris.resize(t0 + 3 * order.size()); hs.shade_tris.resize(s0 + 9 * order.size());
    float4* out_t = hs.tris.data() + t0; float4* out_s = hs.shade_tris.data() + s0;
    parallel_ranges(order.size(), [&](size_t jb, size_t je) {
      const float nz[3] = {0, 0, 1}, tx[3] = {1, 0, 0}, by[3] = {0, 1, 0}, uv0[2] = {0, 0};
      for (size_t j = jb; j < je; j++) {
        const int t = order[j];
        float idf; int32_t id = t; memcpy(&idf, &id, 4);
        const float* P[3]; const float* N[3]; const float* T[3]; const float* B[3]; const float* UV[3];
        for (int c = 0; c < 3; c++) {
          const uint32_t vi = m.index[3 * t + c];
          P[c] = m.position + 3 * vi;
          N[c] = m.normal ? m.normal + 3 * vi : nz; UV[c] = m.texcoord ? m.texcoord + 2 * vi : uv0;
          T[c] = m.tangent ? m.tangent + 3 * vi : tx; B[c] = m.bitangent ? m.bitangent + 3 * vi : by;
        }
        // (p0, id), e1 = p1 - p0, e2 = p2 - p0: the two edge vectors Triangle::intersect_ray recomputes per test (triangle.rs:128-129)
        // are stored instead of the vertices — same IEEE subtraction, done once here (this file is built with -ffp-contract=off)
        out_t[3 * j] = make_float4(P[0][0], P[0][1], P[0][2], idf);
        for (int c = 1; c < 3; c++) out_t[3 * j + c] = make_float4(P[c][0] - P[0][0], P[c][1] - P[0][1], P[c][2] - P[0][2], 0.f);
        // Triangle::pdf / BvhAccel::pdf: 1 / max(area, 1e-3) / n_tris (triangle.rs:272-280, bvh.rs:300-302, Q6)
        const V3 e1{P[1][0] - P[0][0], P[1][1] - P[0][1], P[1][2] - P[0][2]}, e2{P[2][0] - P[0][0], P[2][1] - P[0][1], P[2][2] - P[0][2]};
        const float area = length(cross(e1, e2)) * 0.5f;
        const float prim_pdf = (1.0f / std::max(area, 0.001f)) / (float)m.n_tris;
        const float w[9] = {UV[0][0], UV[0][1], UV[1][0], UV[1][1], UV[2][0], UV[2][1], prim_pdf, 0.f, 0.f};
        for (int c = 0; c < 3; c++) out_s[9 * j + c] = make_float4(N[c][0], N[c][1], N[c][2], w[c]);
        for (int c = 0; c < 3; c++) out_s[9 * j + 3 + c] = make_float4(T[c][0], T[c][1], T[c][2], w[3 + c]);
        for (int c = 0; c < 3; c++) out_s[9 * j + 6 + c] = make_float4(B[c][0], B[c][1], B[c][2], w[6 + c]);
      }
    });
    hs.blas_nodes += (int64_t)blas_nodes[mi].size();
  }
  hs.inst_mat.assign(std::max(1, d.n_instances), 0);
  for (int i = 0; i < d.n_instances; i++) {
    DInstance& I = hs.inst[i];
    if (I.kind == SPT_PRIM_TRIMESH) {
      const DMesh& dm = hs.meshes[I.mesh];
      if (dm.n_tris <= SPT_SMALL_MESH) { I.kind = 2 | (dm.n_tris << 8); I.root = (int32_t)mesh_tri_base[I.mesh]; } // leaf triangles tested directly
      else I.root = dm.blas_root;
    }
    const DSurface& sf = hs.surfaces[I.surface];
    const int tag = hs.materials[sf.material].tag;
    hs.inst_mat[i] = (uint8_t)tag;
    hs.mat_mask |= 1u << tag;
    if (sf.inside_medium >= 0 && !sf.double_sided) hs.has_medium = true;
  }
  // ---- environment -----------------------------------------------------------------------------------------------------------
  float env_avg_power = 0.f;
  if (d.env) { // EnvLight::new, environment.rs:18-50 (Q7: theta lacks the factor pi)
    const int W = d.env->width, H = d.env->height;
    if (W <= 0 || H <= 0 || !d.env->rgb) return bad("environment: empty");
    hs.env_w = W; hs.env_h = H;
    for (int c = 0; c < 3; c++) hs.env_scale[c] = d.env->scale[c];
    std::vector<float> props((size_t)W * H);
    float sum = 0.0f; const float hinv = 1.0f / (float)H;
    for (int y = 0; y < H; y++) for (int x = 0; x < W; x++) {
      const float theta = ((float)y + 0.5f) * hinv;
      const float p = lum(d.env->rgb + 3 * ((size_t)y * W + x)) * std::sin(theta);
      sum += p; props[(size_t)y * W + x] = p;
    }
    const float sinv = 1.0f / sum;
    for (auto& p : props) p *= sinv;
    hs.env_tex.resize(props.size());
    for (size_t i = 0; i < props.size(); i++) hs.env_tex[i] = make_float4(d.env->rgb[3 * i], d.env->rgb[3 * i + 1], d.env->rgb[3 * i + 2], props[i]);
    alias_build(props, hs.env_alias);
    env_avg_power = sum / (float)props.size(); // environment.rs:39; power = avg_power * 4 pi (:139-141)
  }
  // ---- lights + light sampler ----------------------------------------------------------------------------------------------------
  if (d.light_sampler != SPT_LS_UNIFORM && d.light_sampler != SPT_LS_POWER_IS) return bad("light_sampler: unknown kind");
  // pdf_env_light() reads light_prob[env_light_index] (power_is.rs:91-92): the index must name an entry of the light table
  if (d.env_light_index >= std::max(1, d.n_lights)) return bad("env_light_index out of range");
  if (d.env_light_index >= 0 && !d.env) return bad("env_light_index without environment");
  hs.light_sampler = d.light_sampler; hs.env_light_index = d.env_light_index;
  hs.inst_light.assign(std::max(1, d.n_instances), -1);
  hs.lights.resize(d.n_lights);
  std::vector<float> power(d.n_lights, 0.f);
  for (int i = 0; i < d.n_lights; i++) {
    const SptLightDesc& l = d.lights[i];
    DLight o; memset(&o, 0, sizeof(o));
    o.tag = l.tag; o.instance = l.instance;
    for (int c = 0; c < 3; c++) { o.position[c] = l.position[c]; o.direction[c] = l.direction[c]; o.strength[c] = l.strength[c]; }
    o.cos_inner = std::cos(l.inner_angle); o.cos_outer = std::cos(l.outer_angle); // SpotLight::new, spot.rs:22-23
    switch (l.tag) {
      case SPT_LIGHT_DIRECTIONAL: { // DirLight::new normalises, directional.rs:11-16
        V3 dn{l.direction[0], l.direction[1], l.direction[2]}; const float len = length(dn);
        o.direction[0] = dn.x / len; o.direction[1] = dn.y / len; o.direction[2] = dn.z / len;
        power[i] = lum(l.strength); break;
      }
      case SPT_LIGHT_POINT: case SPT_LIGHT_SPOT: power[i] = lum(l.strength); break;
      case SPT_LIGHT_SHAPE: {
        if (l.instance < 0 || l.instance >= d.n_instances) return bad("shape light: bad instance index");
        hs.inst_light[l.instance] = i;
        const SptInstanceDesc& in = d.instances[l.instance];
        const A3& a = inst_fwd[l.instance];
        float area = 0.f;
        if (in.prim_kind == SPT_PRIM_SPHERE) { // Sphere::surface_area, sphere.rs:122-135
          const float r = d.spheres[in.prim_index].radius * 0.5f;
          const V3 v0 = mul(a.m, V3{-r, -r, -r}), v1 = mul(a.m, V3{-r, -r, r}), v2 = mul(a.m, V3{-r, r, -r}), v3 = mul(a.m, V3{r, -r, -r});
          const float a2 = dot(v0 - v1, v0 - v1), b2 = dot(v0 - v2, v0 - v2), c2 = dot(v0 - v3, v0 - v3);
          area = 4.0f * 3.14159265358979323846f * std::sqrt((a2 * b2 + b2 * c2 + c2 * a2) / 3.0f);
        } else { // sum of Triangle::surface_area, triangle.rs:282-288
          const SptMeshDesc& m = d.meshes[in.prim_index];
          for (int t = 0; t < m.n_tris; t++) {
            V3 p[3];
            for (int c = 0; c < 3; c++) { const float* q = m.position + 3 * m.index[3 * t + c]; p[c] = point(a, V3{q[0], q[1], q[2]}); }
            area += length(cross(p[1] - p[0], p[2] - p[0])) * 0.5f;
          }
        }
        const SptSurfaceDesc& sf = d.surfaces[in.surface];
        float e[3] = {sf.emissive[0], sf.emissive[1], sf.emissive[2]};
        if (sf.emissive_map >= 0) { float av[3]; tex_average(hs, sf.emissive_map, av); for (int c = 0; c < 3; c++) e[c] *= av[c]; }
        power[i] = area * lum(e);
        break;
      }
      case SPT_LIGHT_ENV:
        if (!d.env) return bad("env light without environment");
        power[i] = env_avg_power * 4.0f * 3.14159265358979323846f; break;
      default: return bad("light: unknown tag");
    }
    hs.lights[i] = o;
  }
  hs.light_prob.assign(std::max(1, d.n_lights), 0.f);
  hs.light_alias.assign(std::max(1, d.n_lights), make_float2(1.f, 0.f));
  if (d.light_sampler == SPT_LS_POWER_IS && d.n_lights > 0) { // power_is.rs:24-34
    float sum = 0.f;
    for (float p : power) sum += p;
    const float sinv = 1.0f / sum;
    std::vector<float> props(d.n_lights);
    for (int i = 0; i < d.n_lights; i++) props[i] = power[i] * sinv;
    hs.light_prob = props;
    alias_build(props, hs.light_alias);
  }
  // ---- SS_CDF_TABLE, substrate.rs:187-196 ---------------------------------------------------------------------------------------------
  hs.ss_cdf.resize(1024);
  for (int i = 0; i < 512; i++) {
    float x = (float)i / 512.0f;
    x = -2.0f * std::log(1.0f - x);
    hs.ss_cdf[2 * i] = x;
    hs.ss_cdf[2 * i + 1] = 1.0f - std::exp(-x) * 0.25f - std::exp(-x / 3.0f) * 0.75f;
  }
  // ---- cameras, PerspectiveCamera::new (perspective.rs:15-27) ---------------------------------------------------------------------------
  hs.cameras.resize(d.n_cameras);
  for (int i = 0; i < d.n_cameras; i++) {
    const SptCameraDesc& c = d.cameras[i];
    V3 f{c.forward[0], c.forward[1], c.forward[2]}, u{c.up[0], c.up[1], c.up[2]};
    const float fl = length(f); f = {f.x / fl, f.y / fl, f.z / fl};
    V3 r = cross(f, u); const float rl = length(r); r = {r.x / rl, r.y / rl, r.z / rl};
    u = cross(r, f);
    DCamera o;
    for (int k = 0; k < 3; k++) { o.eye[k] = c.eye[k]; o.forward[k] = (&f.x)[k]; o.right[k] = (&r.x)[k]; o.up[k] = (&u.x)[k]; }
    o.half_cot_half_fov = 0.5f / std::tan(c.fov * 0.5f);
    hs.cameras[i] = o;
  }
  return std::string();
}

} // namespace spt
