// oracle/prims.h — TEST INFRASTRUCTURE ONLY.  Restates primitive/{triangle,sphere,instance,group}.rs.
// PARITY UNPINNED (see vec.h).
#pragma once
#include <memory>
#include "geom.h"

namespace orc {

struct MeshVertex { Vec3 position, normal; Vec2 texcoords; Vec3 tangent, bitangent; }; // triangle.rs:10-17

struct Triangle { // triangle.rs:23-27
  const std::vector<MeshVertex>* vertices = nullptr;
  size_t indices[3] = {0, 0, 0};
  Bbox bbox_;
  int tri_id = 0; // oracle-only: index in the mesh's file order (the "primitive ID")

  void init(const std::vector<MeshVertex>* v, size_t i0, size_t i1, size_t i2, int id) { // Triangle::new :111-121
    vertices = v; indices[0] = i0; indices[1] = i1; indices[2] = i2; tri_id = id;
    Vec3 p[3] = {(*v)[i0].position, (*v)[i1].position, (*v)[i2].position};
    bbox_ = Bbox::from_points(p, 3);
  }
  const MeshVertex& V(int k) const { return (*vertices)[indices[k]]; }

  bool intersect_ray(const Ray& ray, float& t, float& u, float& v, float& w) const { // :124-147 (Q4)
    g_cnt.prims++;
    Vec3 p0 = V(0).position, p1 = V(1).position, p2 = V(2).position;
    Vec3 e1 = p1 - p0, e2 = p2 - p0;
    Vec3 q = cross(ray.direction, e2);
    float det = dot(e1, q);
    if (det != 0.0f) {
      det = 1.0f / det;
      Vec3 s = ray.origin - p0;
      v = dot(s, q) * det;
      if (v >= 0.0f) {
        Vec3 r = cross(s, e1);
        w = dot(ray.direction, r) * det;
        u = 1.0f - v - w;
        if (w >= 0.0f && u >= 0.0f) { t = dot(e2, r) * det; return true; }
      }
    }
    return false;
  }
  bool intersect_test(const Ray& ray, float t_max) const { // :177-183
    float t, u, v, w;
    if (intersect_ray(ray, t, u, v, w)) return t > ray.t_min && t < t_max;
    return false;
  }
  bool intersect(const Ray& ray, Intersection& inter) const { // :185-218
    float t, u, v, w;
    if (intersect_ray(ray, t, u, v, w)) {
      if (t > ray.t_min && t < inter.t) {
        inter.t = t;
        inter.normal = normalize(V(0).normal * u + V(1).normal * v + V(2).normal * w);
        Vec2 a = V(0).texcoords, b = V(1).texcoords, c = V(2).texcoords;
        inter.texcoords = {a.x * u + b.x * v + c.x * w, a.y * u + b.y * v + c.y * w}; // lerp_point2 :290-302
        inter.tangent = V(0).tangent * u + V(1).tangent * v + V(2).tangent * w;
        inter.bitangent = V(0).bitangent * u + V(1).bitangent * v + V(2).bitangent * w;
        inter.primitive = PrimRef{PrimRef::kTriangle, nullptr, this};
        inter.bu = v; inter.bv = w;
        return true;
      }
    }
    return false;
  }
  Bbox bbox() const { return bbox_; }
  float area() const { return length(cross(V(1).position - V(0).position, V(2).position - V(0).position)) * 0.5f; }
  Intersection sample(Rng& rng, float& pdf) const { // :224-270
    float r0, r1; rng.uniform_2d(r0, r1);
    float r0s = std::sqrt(r0);
    float u = 1.0f - r0s, v = r0s * (1.0f - r1), w = 1.0f - u - v;
    Intersection inter;
    inter.position = V(0).position * u + V(1).position * v + V(2).position * w;
    inter.normal = V(0).normal * u + V(1).normal * v + V(2).normal * w; // not normalised here (:259)
    inter.tangent = V(0).tangent * u + V(1).tangent * v + V(2).tangent * w;
    inter.bitangent = V(0).bitangent * u + V(1).bitangent * v + V(2).bitangent * w;
    Vec2 a = V(0).texcoords, b = V(1).texcoords, c = V(2).texcoords;
    inter.texcoords = (a * u + b * v) + c * w; // glam Vec2 ops, :262
    inter.primitive = PrimRef{PrimRef::kTriangle, nullptr, this};
    pdf = 1.0f / fmax_(area(), 0.001f); // Q6
    return inter;
  }
  float pdf(const Intersection&) const { return 1.0f / fmax_(area(), 0.001f); } // :272-280
  float surface_area(const Transform& t) const {                                // :282-288
    Vec3 p0 = t.point(V(0).position), p1 = t.point(V(1).position), p2 = t.point(V(2).position);
    return length(cross(p1 - p0, p2 - p0)) * 0.5f;
  }
};

inline Vec2 sphere_normal_to_texcoords(Vec3 p) { // sphere.rs:138-145
  float theta = std::acos(p.y);
  float phi = std::atan2(p.x, p.z) + PI;
  return {phi * 0.5f * FRAC_1_PI, theta * FRAC_1_PI};
}

struct Sphere { // sphere.rs:8-12
  Vec3 center; float radius = 1; Bbox bbox_;
  void init(Vec3 c, float r) { center = c; radius = r; Vec3 d{r, r, r}; bbox_.p_min = c - d; bbox_.p_max = c + d; }
  bool intersect_ray(const Ray& ray, float& mn, float& mx) const { // :25-39
    g_cnt.prims++;
    Vec3 oc = ray.origin - center;
    float a = length_squared(ray.direction);
    float b = dot(ray.direction, oc);
    float c = length_squared(oc) - radius * radius;
    float delta = b * b - a * c;
    if (delta >= 0.0f) { delta = std::sqrt(delta); mn = (-b - delta) / a; mx = (-b + delta) / a; return true; }
    return false;
  }
  bool intersect_test(const Ray& ray, float t_max) const { // :51-57
    float mn, mx;
    if (intersect_ray(ray, mn, mx)) return mn < t_max && mx > ray.t_min;
    return false;
  }
  static void frame(Vec3 norm, Vec3& tangent, Vec3& bitangent) { // :66-77, :104-114
    float sin_theta = std::sqrt(1.0f - norm.y * norm.y);
    if (sin_theta != 0.0f) {
      bitangent = norm * (-norm.y / sin_theta);
      bitangent.y = sin_theta;
      tangent = cross(bitangent, norm);
    } else if (norm.y > 0.0f) { bitangent = {1, 0, 0}; tangent = {0, 0, 1}; }
    else { bitangent = {-1, 0, 0}; tangent = {0, 0, -1}; }
  }
  bool intersect(const Ray& ray, Intersection& inter) const { // :59-84
    float mn, mx;
    if (intersect_ray(ray, mn, mx)) {
      float t = mn < ray.t_min ? mx : mn;
      if (ray.t_min < t && t < inter.t) {
        inter.t = t;
        Vec3 norm = (ray.point_at(t) - center) / radius;
        inter.normal = norm;
        frame(norm, inter.tangent, inter.bitangent);
        inter.texcoords = sphere_normal_to_texcoords(norm);
        inter.primitive = PrimRef{PrimRef::kSphere, this, nullptr};
        inter.bu = inter.texcoords.x; inter.bv = inter.texcoords.y;
        return true;
      }
    }
    return false;
  }
  Bbox bbox() const { return bbox_; }
  Intersection sample(Rng& rng, float& pdf) const { // :90-116 (Q5)
    Vec3 norm = rng.uniform_on_sphere();
    Intersection inter;
    inter.position = center + norm * radius;
    inter.normal = norm;
    inter.texcoords = sphere_normal_to_texcoords(norm);
    inter.primitive = PrimRef{PrimRef::kSphere, this, nullptr};
    frame(norm, inter.tangent, inter.bitangent);
    pdf = 0.25f * FRAC_1_PI;
    return inter;
  }
  float pdf(const Intersection&) const { return 0.25f * FRAC_1_PI; } // :118-120
  float surface_area(const Transform& t) const {                     // :122-135
    float r = radius * 0.5f;
    Vec3 v0 = t.vector({-r, -r, -r}), v1 = t.vector({-r, -r, r}), v2 = t.vector({-r, r, -r}), v3 = t.vector({r, -r, -r});
    float a2 = distance_squared(v0, v1), b2 = distance_squared(v0, v2), c2 = distance_squared(v0, v3);
    return 4.0f * PI * std::sqrt((a2 * b2 + b2 * c2 + c2 * a2) / 3.0f);
  }
};

struct TriMesh { // triangle.rs:19-21, 42-55
  std::vector<MeshVertex> vertices;
  std::vector<Triangle> tris;
  BvhAccel<Triangle> bvh;
  int mesh_id = 0;
  void build(const std::vector<uint32_t>& idx) {
    size_t n = idx.size() / 3;
    tris.resize(n);
    std::vector<const Triangle*> ptrs(n);
    for (size_t i = 0; i < n; i++) { tris[i].init(&vertices, idx[3 * i], idx[3 * i + 1], idx[3 * i + 2], (int)i); ptrs[i] = &tris[i]; }
    bvh.build(ptrs, 4, 16);
  }
};

// The reference's `Primitive` enum restricted to the in-scope leaves an Instance can hold.
struct Shape {
  const Sphere* sphere = nullptr;
  const TriMesh* mesh = nullptr;
  bool intersect_test(const Ray& r, float t_max) const { return sphere ? sphere->intersect_test(r, t_max) : mesh->bvh.intersect_test(r, t_max); }
  bool intersect(const Ray& r, Intersection& i) const { return sphere ? sphere->intersect(r, i) : mesh->bvh.intersect(r, i); }
  Bbox bbox() const { return sphere ? sphere->bbox() : mesh->bvh.bbox(); }
  Intersection sample(Rng& rng, float& pdf) const { return sphere ? sphere->sample(rng, pdf) : mesh->bvh.sample(rng, pdf); }
  float pdf(const Intersection& inter) const { // BvhAccel::pdf bvh.rs:300-302 / Sphere::pdf
    if (sphere) return sphere->pdf(inter);
    float p = inter.primitive.kind == PrimRef::kTriangle ? inter.primitive.tri->pdf(inter) : inter.primitive.sphere->pdf(inter);
    return p / (float)mesh->bvh.primitives.size();
  }
  float surface_area(const Transform& t) const { return sphere ? sphere->surface_area(t) : mesh->bvh.surface_area(t); }
};

struct Instance { // instance.rs:10-16
  Shape primitive;
  Transform trans;
  Affine3 trans_inv;
  Bbox bbox_;
  const Surface* surface = nullptr;
  int inst_id = 0;
  void init(Shape p, const Affine3& t, const Surface* s, int id) { // Instance::new :19-29
    primitive = p; trans_inv = inverse(t); bbox_ = p.bbox().transformed_by(t); trans = Transform(t); surface = s; inst_id = id;
  }
  bool intersect_test(const Ray& ray, float t_max) const { // :89-92
    g_cnt.insts++;
    return primitive.intersect_test(ray.transformed_by(trans_inv), t_max);
  }
  bool intersect(const Ray& ray, Intersection& inter) const { // :94-109
    g_cnt.insts++;
    Ray tr = ray.transformed_by(trans_inv);
    if (primitive.intersect(tr, inter)) {
      inter.instance = this;
      inter.surface = surface;
      inter.position = ray.point_at(inter.t);
      inter.normal = trans.normal(inter.normal);
      inter.tangent = trans.vector(inter.tangent);
      inter.bitangent = trans.vector(inter.bitangent);
      return true;
    }
    return false;
  }
  Bbox bbox() const { return bbox_; }
  Intersection sample(Rng& rng, float& pdf) const { // :115-129
    float p;
    Intersection inter = primitive.sample(rng, p);
    inter.surface = surface;
    float original_area = length(cross(inter.tangent, inter.bitangent));
    inter.position = trans.point(inter.position);
    inter.normal = trans.normal(inter.normal);
    inter.bitangent = trans.vector(inter.bitangent);
    inter.tangent = trans.vector(inter.tangent);
    float transformed_area = length(cross(inter.tangent, inter.bitangent));
    pdf = p * original_area / transformed_area;
    return inter;
  }
  float pdf(const Intersection& inter) const { // :131-141
    Transform ti = trans.inverse_();
    Vec3 tangent = ti.vector(inter.tangent), bitangent = ti.vector(inter.bitangent);
    float original_area = length(cross(tangent, bitangent));
    float transformed_area = length(cross(inter.tangent, inter.bitangent));
    return primitive.pdf(inter) * original_area / transformed_area;
  }
  float surface_area(const Transform& t) const { return primitive.surface_area(t.mul(trans)); } // :143-146
};

struct Aggregate { // Scene::aggregate: Group<Instance> (group.rs) or BvhAccel<Instance>
  bool use_bvh = true;
  std::vector<const Instance*> group;
  BvhAccel<Instance> bvh;
  bool intersect_test(const Ray& ray, float t_max) const {
    if (use_bvh) return bvh.intersect_test(ray, t_max);
    for (auto* p : group) if (p->intersect_test(ray, t_max)) return true; // group.rs:25-32
    return false;
  }
  bool intersect(const Ray& ray, Intersection& inter) const {
    if (use_bvh) return bvh.intersect(ray, inter);
    bool r = false;
    for (auto* p : group) r |= p->intersect(ray, inter); // group.rs:34-40
    return r;
  }
};

inline bool primref_intersect(const PrimRef& p, const Ray& ray, Intersection& inter) { // primitive/mod.rs:95-101
  if (p.kind == PrimRef::kTriangle) return p.tri->intersect(ray, inter);
  if (p.kind == PrimRef::kSphere) return p.sphere->intersect(ray, inter);
  return false;
}

} // namespace orc
