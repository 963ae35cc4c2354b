// oracle/geom.h — TEST INFRASTRUCTURE ONLY.  Restates the reference's geometry layer:
//   core/ray.rs, core/bbox.rs, core/intersection.rs, core/transform.rs,
//   primitive/{bvh,triangle,sphere,instance,group,mod}.rs
// keeping its binary BVH (incl. quirks Q1/Q2) and traversal order (Q3) so that tie-breaking of
// equal-t hits is the reference's.  PARITY UNPINNED (see vec.h).
#pragma once
#include <vector>
#include <unordered_set>
#include "vec.h"
#include "rng.h"

namespace orc {

struct Counters { uint64_t nodes = 0, prims = 0, insts = 0; };
inline thread_local Counters g_cnt;
// Oracle-only switches (both default to the reference's behaviour):
//  g_fix_q1          build the BVH with the evidently intended bucket rule (centroid relative to the NODE box).
//                    Hits are topology-independent (except exact-t ties), so this only makes the checker fast
//                    enough for high-spp image comparisons; never used for the timed CPU baseline.
//  g_file_order_pick BvhAccel::sample picks primitives in file order instead of BVH order (a permutation of a
//                    uniform pick).  Set in keyed/paired RNG mode so the CUDA core, whose BVH differs, draws the
//                    same triangle from the same random number.
inline bool g_fix_q1 = false;
inline thread_local bool g_file_order_pick = false;

struct AuxRay { Vec3 x_origin, x_direction, y_origin, y_direction; }; // ray.rs:9-15
struct Ray {                                                          // ray.rs:1-7
  Vec3 origin, direction;
  float t_min = 0.0001f; // Ray::T_MIN_EPS, ray.rs:18
  bool has_aux = false;
  AuxRay aux;
  Ray() = default;
  Ray(Vec3 o, Vec3 d) : origin(o), direction(d) {}
  Vec3 point_at(float t) const { return origin + direction * t; } // ray.rs:29-31
  Ray transformed_by(const Affine3& a) const {                     // ray.rs:33-41
    Ray r = *this;
    r.origin = transform_point(a, origin);
    r.direction = transform_vector(a, direction);
    return r;
  }
};
constexpr float T_MIN_EPS = 0.0001f;

struct Bbox { // bbox.rs
  Vec3 p_min{FLT_MAX, FLT_MAX, FLT_MAX}, p_max{-FLT_MAX, -FLT_MAX, -FLT_MAX}; // empty(), :25-30
  static Bbox from_points(const Vec3* p, int n) {
    Bbox b; b.p_min = b.p_max = p[0];
    for (int i = 1; i < n; i++) { b.p_min = vmin(b.p_min, p[i]); b.p_max = vmax(b.p_max, p[i]); }
    return b;
  }
  bool is_empty() const { return p_min.x > p_max.x || p_min.y > p_max.y || p_min.z > p_max.z; }
  Bbox merge(const Bbox& o) const { Bbox b; b.p_min = vmin(p_min, o.p_min); b.p_max = vmax(p_max, o.p_max); return b; }
  Bbox transformed_by(const Affine3& t) const { // :41-61
    Vec3 p[8];
    int k = 0;
    for (int ix = 0; ix < 2; ix++) for (int iy = 0; iy < 2; iy++) for (int iz = 0; iz < 2; iz++)
      p[k++] = transform_point(t, Vec3(ix ? p_max.x : p_min.x, iy ? p_max.y : p_min.y, iz ? p_max.z : p_min.z));
    return from_points(p, 8);
  }
  bool intersect_ray(const Ray& ray, float& t0, float& t1) const { // :63-85
    if (is_empty()) return false;
    float x0 = (p_min.x - ray.origin.x) / ray.direction.x, x1 = (p_max.x - ray.origin.x) / ray.direction.x;
    float xa = fmin_(x0, x1), xb = fmax_(x0, x1);
    float y0 = (p_min.y - ray.origin.y) / ray.direction.y, y1 = (p_max.y - ray.origin.y) / ray.direction.y;
    float ya = fmin_(y0, y1), yb = fmax_(y0, y1);
    float z0 = (p_min.z - ray.origin.z) / ray.direction.z, z1 = (p_max.z - ray.origin.z) / ray.direction.z;
    float za = fmin_(z0, z1), zb = fmax_(z0, z1);
    t0 = fmax_(xa, fmax_(ya, za));
    t1 = fmin_(xb, fmin_(yb, zb));
    return t0 <= t1;
  }
  bool intersect_test(const Ray& ray, float t_max) const { // :87-93
    float t0, t1;
    if (intersect_ray(ray, t0, t1)) return t1 > ray.t_min && t0 < t_max;
    return false;
  }
  float surface_area() const { // :95-102 — really the volume (Q2)
    if (is_empty()) return 0.0f;
    Vec3 d = p_max - p_min;
    return d.x * d.y * d.z;
  }
  Vec3 centroid() const { return (p_min + p_max) * 0.5f; }
};

struct Transform { // transform.rs:1-48
  Affine3 trans;
  Mat3 trans_it;
  Transform() = default;
  explicit Transform(const Affine3& t) : trans(t) { trans_it = transpose(inverse(t).m); }
  Vec3 point(Vec3 p) const { return transform_point(trans, p); }
  Vec3 vector(Vec3 v) const { return transform_vector(trans, v); }
  Vec3 normal(Vec3 n) const { return normalize(trans_it * n); }
  Transform inverse_() const { Transform r; r.trans = inverse(trans); r.trans_it = inverse(trans_it); return r; }
  Transform mul(const Transform& rhs) const { Transform r; r.trans = trans * rhs.trans; r.trans_it = trans_it * rhs.trans_it; return r; }
};

struct Instance; struct Surface; struct Triangle; struct Sphere;

struct PrimRef { // BasicPrimitiveRef, primitive/mod.rs:79-84
  enum Kind { kNone, kSphere, kTriangle } kind = kNone;
  const Sphere* sphere = nullptr;
  const Triangle* tri = nullptr;
};

struct Intersection { // intersection.rs:6-18, Default :86-102
  float t = FLT_MAX;
  Vec3 position{0, 0, 0}, tangent{1, 0, 0}, bitangent{0, 1, 0}, normal{0, 0, 1};
  Vec2 texcoords{0, 0};
  const Instance* instance = nullptr;
  PrimRef primitive;
  const Surface* surface = nullptr;
  Vec2 duvdx{0, 0}, duvdy{0, 0};
  float bu = 0, bv = 0; // oracle-only: barycentric (v,w) of the accepted triangle hit, for SptHit
  static Intersection with_t_max(float t) { Intersection i; i.t = t; return i; }
  void calc_differential(const Ray& ray);
};

inline bool solve2x2(float a00, float a01, float a10, float a11, Vec2 b, float& x1, float& x2) { // :104-118
  // a.col(0) = (a00, a01), a.col(1) = (a10, a11)
  float det = a00 * a11 - a10 * a01; // glam Mat2::determinant
  if (det != 0.0f) {
    float temp = b.y * a00 - b.x * a01;
    x2 = temp / det;
    x1 = std::fabs(a00) > std::fabs(a01) ? (b.x - a10 * x2) / a00 : (b.y - a11 * x2) / a01;
    return true;
  }
  return false;
}

inline void Intersection::calc_differential(const Ray& ray) { // intersection.rs:28-83
  if (!ray.has_aux) return;
  const AuxRay& ar = ray.aux;
  Vec3 p = ray.point_at(t);
  float d = dot(p, normal);
  float tx = (d - dot(ar.x_origin, normal)) / dot(ar.x_direction, normal);
  Vec3 px = ar.x_origin + ar.x_direction * tx;
  float ty = (d - dot(ar.y_origin, normal)) / dot(ar.y_direction, normal);
  Vec3 py = ar.y_origin + ar.y_direction * ty;
  Vec3 dpdx = px - p, dpdy = py - p;
  Vec2 bx, by;
  float a00, a01, a10, a11;
  float ax = std::fabs(normal.x), ay = std::fabs(normal.y), az = std::fabs(normal.z);
  if (ax >= ay && ax >= az) {
    bx = {dpdx.y, dpdx.z}; by = {dpdy.y, dpdy.z};
    a00 = tangent.y; a01 = tangent.z; a10 = bitangent.y; a11 = bitangent.z;
  } else if (ay >= az) {
    bx = {dpdx.z, dpdx.x}; by = {dpdy.z, dpdy.x};
    a00 = tangent.z; a01 = tangent.x; a10 = bitangent.z; a11 = bitangent.x;
  } else {
    bx = {dpdx.x, dpdx.y}; by = {dpdy.x, dpdy.y};
    a00 = tangent.x; a01 = tangent.y; a10 = bitangent.x; a11 = bitangent.y;
  }
  float x1, x2;
  if (solve2x2(a00, a01, a10, a11, bx, x1, x2)) duvdx = {x1, x2};
  if (solve2x2(a00, a01, a10, a11, by, x1, x2)) duvdy = {x1, x2};
}

// ---- primitive/bvh.rs ---------------------------------------------------------------------------
template <class P>
struct BvhAccel {
  struct Node { int lc = -1, rc = -1; Bbox bbox; size_t start = 0, end = 0; size_t size() const { return end - start; } bool is_leaf() const { return lc < 0; } };
  std::vector<Node> nodes; // nodes[0] = root (empty vector = no root)
  std::vector<const P*> primitives;
  std::vector<const P*> file_order; // oracle-only, see g_file_order_pick

  static void find_best_split(const std::vector<Bbox>& boxes, const std::vector<std::vector<size_t>>& idx,
                              size_t total, size_t nb, float& best_cost, size_t& best_split) { // :143-174
    best_cost = FLT_MAX; best_split = 0;
    std::vector<Bbox> bl = boxes, br = boxes;
    std::vector<size_t> cnt(nb, 0);
    cnt[0] = idx[0].size();
    for (size_t i = 1; i < nb; i++) { bl[i] = bl[i].merge(bl[i - 1]); cnt[i] = cnt[i - 1] + idx[i].size(); }
    for (size_t i = nb - 1; i-- > 0;) br[i] = br[i].merge(br[i + 1]);
    for (size_t i = 1; i < nb; i++) {
      float la = bl[i - 1].surface_area(), lc = (float)cnt[i - 1];
      float ra = br[i].surface_area(), rc = (float)total - lc;
      float cost = la * lc + ra * rc;
      if (cost < best_cost) { best_cost = cost; best_split = i; }
    }
  }

  void build(std::vector<const P*> prims, size_t max_leaf, size_t nb) { // BvhAccel::new, :23-141
    primitives = std::move(prims);
    file_order = primitives;
    nodes.clear();
    if (primitives.empty()) return;
    Bbox bbox = primitives[0]->bbox();
    for (size_t i = 1; i < primitives.size(); i++) bbox = bbox.merge(primitives[i]->bbox());
    nodes.push_back(Node{-1, -1, bbox, 0, primitives.size()});
    std::vector<int> stack{0};
    while (!stack.empty()) {
      int ui = stack.back(); stack.pop_back();
      if (nodes[ui].size() <= max_leaf) continue;
      Bbox ub = nodes[ui].bbox;
      size_t ustart = nodes[ui].start, uend = nodes[ui].end;
      Vec3 len = (ub.p_max - ub.p_min) / (float)nb;
      std::vector<Bbox> bx(nb), by(nb), bz(nb);
      std::vector<std::vector<size_t>> ix(nb), iy(nb), iz(nb);
      for (size_t i = ustart; i < uend; i++) {
        Bbox pb = primitives[i]->bbox(); // shadows the node bbox: Q1 (bvh.rs:55-58)
        Vec3 c = pb.centroid();
        Vec3 ref_min = g_fix_q1 ? ub.p_min : pb.p_min;
        float x = (c.x - ref_min.x) / len.x;
        if (g_fix_q1 && x >= (float)nb) x = (float)(nb - 1); // centroid on the node's max face
        if (x >= 0.0f && x < (float)nb) { size_t k = (size_t)x; bx[k] = bx[k].merge(pb); ix[k].push_back(i); }
        float y = (c.y - ref_min.y) / len.y;
        if (g_fix_q1 && y >= (float)nb) y = (float)(nb - 1); // centroid on the node's max face
        if (y >= 0.0f && y < (float)nb) { size_t k = (size_t)y; by[k] = by[k].merge(pb); iy[k].push_back(i); }
        float z = (c.z - ref_min.z) / len.z;
        if (g_fix_q1 && z >= (float)nb) z = (float)(nb - 1); // centroid on the node's max face
        if (z >= 0.0f && z < (float)nb) { size_t k = (size_t)z; bz[k] = bz[k].merge(pb); iz[k].push_back(i); }
      }
      float cx = FLT_MAX, cy = FLT_MAX, cz = FLT_MAX;
      size_t sx = nb / 2, sy = nb / 2, sz = nb / 2;
      if (len.x > 0.0001f) find_best_split(bx, ix, uend - ustart, nb, cx, sx);
      if (len.y > 0.0001f) find_best_split(by, iy, uend - ustart, nb, cy, sy);
      if (len.z > 0.0001f) find_best_split(bz, iz, uend - ustart, nb, cz, sz);
      Node lc, rc;
      if (cx <= cy && cx <= cz) split_at(sx, nb, bx, ix, ustart, uend, lc, rc);
      else if (cy <= cx && cy <= cz) split_at(sy, nb, by, iy, ustart, uend, lc, rc);
      else split_at(sz, nb, bz, iz, ustart, uend, lc, rc);
      if (lc.size() == 0 || rc.size() == 0) continue;
      int li = (int)nodes.size(); nodes.push_back(lc);
      int ri = (int)nodes.size(); nodes.push_back(rc);
      nodes[ui].lc = li; nodes[ui].rc = ri;
      stack.push_back(li); stack.push_back(ri);
    }
  }

  void split_at(size_t split, size_t nb, const std::vector<Bbox>& boxes, std::vector<std::vector<size_t>>& idx,
                size_t start, size_t end, Node& lc, Node& rc) { // :176-215
    Bbox lb, rb;
    std::unordered_set<size_t> left, right;
    for (size_t i = 0; i < split; i++) { lb = lb.merge(boxes[i]); left.insert(idx[i].begin(), idx[i].end()); }
    for (size_t i = split; i < nb; i++) { rb = rb.merge(boxes[i]); right.insert(idx[i].begin(), idx[i].end()); }
    size_t rp = end - 1, mid = start + left.size();
    for (size_t lp = start; lp < mid; lp++) {
      if (!left.count(lp)) {
        while (right.count(rp)) rp--;
        std::swap(primitives[lp], primitives[rp]);
        rp--;
      }
    }
    lc = Node{-1, -1, lb, start, mid};
    rc = Node{-1, -1, rb, mid, end};
  }

  bool intersect_test(const Ray& ray, float t_max) const { // :238-260
    if (nodes.empty()) return false;
    int stack[128]; int sp = 0; stack[sp++] = 0;
    while (sp) {
      const Node& u = nodes[stack[--sp]];
      g_cnt.nodes++;
      if (!u.bbox.intersect_test(ray, t_max)) continue;
      if (u.is_leaf()) {
        for (size_t i = u.start; i < u.end; i++) if (primitives[i]->intersect_test(ray, t_max)) return true;
      } else { stack[sp++] = u.lc; stack[sp++] = u.rc; }
    }
    return false;
  }
  bool intersect(const Ray& ray, Intersection& inter) const { // :262-283
    if (nodes.empty()) return false;
    int stack[128]; int sp = 0; stack[sp++] = 0;
    bool result = false;
    while (sp) {
      const Node& u = nodes[stack[--sp]];
      g_cnt.nodes++;
      if (!u.bbox.intersect_test(ray, inter.t)) continue;
      if (u.is_leaf()) {
        for (size_t i = u.start; i < u.end; i++) result |= primitives[i]->intersect(ray, inter);
      } else { stack[sp++] = u.lc; stack[sp++] = u.rc; }
    }
    return result;
  }
  Bbox bbox() const { return nodes.empty() ? Bbox() : nodes[0].bbox; }
  Intersection sample(Rng& rng, float& pdf) const { // :293-298
    float fi = rng.uniform_1d() * (float)primitives.size();
    size_t index = std::min((size_t)fi, primitives.size() - 1);
    Intersection inter = (g_file_order_pick ? file_order[index] : primitives[index])->sample(rng, pdf);
    pdf = pdf / (float)primitives.size();
    return inter;
  }
  float surface_area(const Transform& t) const { // :304-309
    float s = 0; for (auto* p : primitives) s += p->surface_area(t); return s;
  }
};

} // namespace orc
