// traverse.cuh — closest-hit and any-hit ray queries against the two-level compressed wide BVH.
//
// Replaces, for a whole ray at a time, the reference's
//   BvhAccel::<Instance>::intersect / intersect_test        src/primitive/bvh.rs:262-283 / :238-260
//   Instance::intersect / intersect_test                    src/primitive/instance.rs:94-109 / :89-92
//   BvhAccel::<Triangle>::intersect(_test) + Triangle::*    src/primitive/triangle.rs:124-147, :177-218
//   Sphere::intersect / intersect_test                      src/primitive/sphere.rs:25-84
//   Group::intersect(_test)                                 src/primitive/group.rs:25-40 (same results, any order)
// The leaf tests reproduce the reference's f32 formulas operation by operation without FMA contraction, on the
// same un-normalised object-space ray (core/ray.rs:33-41), so `t` is comparable bit for bit; only the order in
// which equal-t candidates are met differs.
//
// Traversal: one thread per ray, node groups (base index + hit bitmask) on a short per-thread stack, children
// visited in octant order, 5 x 16-byte loads per node, 3 x 16-byte loads per triangle, 4 per instance.
#pragma once
#include "scene.h"
#include "vecmath.cuh"

struct HitRec {
  float t, u, v;
  int prim, inst;
};
struct TravCounters { uint32_t nodes, tris, insts; };

#define SPT_STACK 32

SPT_D uint32_t sext_s8x4(uint32_t x) { uint32_t v; asm("prmt.b32 %0, %1, 0x0, 0x0000BA98;" : "=r"(v) : "r"(x)); return v; }
SPT_D uint32_t bfind_u32(uint32_t x) { uint32_t v; asm("bfind.u32 %0, %1;" : "=r"(v) : "r"(x)); return v; }
SPT_D float4 ldg4(const float4* p) { return __ldg(p); }

struct RaySetup { float3 o, d, idir; uint32_t octinv4; };
SPT_D void ray_setup(RaySetup& r, float3 o, float3 d) {
  r.o = o; r.d = d;
  const float eps = 1e-18f;
  float dx = fabsf(d.x) < eps ? copysignf(eps, d.x) : d.x;
  float dy = fabsf(d.y) < eps ? copysignf(eps, d.y) : d.y;
  float dz = fabsf(d.z) < eps ? copysignf(eps, d.z) : d.z;
  r.idir = mk3(1.0f / dx, 1.0f / dy, 1.0f / dz);
  uint32_t oct = (d.x < 0.f ? 4u : 0u) | (d.y < 0.f ? 2u : 0u) | (d.z < 0.f ? 1u : 0u);
  r.octinv4 = (7u - oct) * 0x01010101u;
}

// Slab test of the 8 quantised child boxes of one node; returns the node-group / primitive-group pair.
SPT_D void node_children(const float4* __restrict__ nodes, uint32_t node_index, const RaySetup& r, float tmin, float tmax,
                         uint2& ng, uint2& tg) {
  const float4* np = nodes + (size_t)node_index * 5;
  const float4 n0 = ldg4(np), n1 = ldg4(np + 1), n2 = ldg4(np + 2), n3 = ldg4(np + 3), n4 = ldg4(np + 4);
  const uint32_t eim = __float_as_uint(n0.w);
  const float sx = __uint_as_float((uint32_t)((int)(int8_t)(eim & 0xff) + 127) << 23);
  const float sy = __uint_as_float((uint32_t)((int)(int8_t)((eim >> 8) & 0xff) + 127) << 23);
  const float sz = __uint_as_float((uint32_t)((int)(int8_t)((eim >> 16) & 0xff) + 127) << 23);
  const float ax = sx * r.idir.x, ay = sy * r.idir.y, az = sz * r.idir.z;
  const float bx = (n0.x - r.o.x) * r.idir.x, by = (n0.y - r.o.y) * r.idir.y, bz = (n0.z - r.o.z) * r.idir.z;
  // near/far planes per axis according to the ray direction sign
  const bool nx = r.idir.x < 0.f, ny = r.idir.y < 0.f, nz = r.idir.z < 0.f;
  const uint32_t lox[2] = {__float_as_uint(n2.x), __float_as_uint(n2.y)}, loy[2] = {__float_as_uint(n2.z), __float_as_uint(n2.w)};
  const uint32_t loz[2] = {__float_as_uint(n3.x), __float_as_uint(n3.y)}, hix[2] = {__float_as_uint(n3.z), __float_as_uint(n3.w)};
  const uint32_t hiy[2] = {__float_as_uint(n4.x), __float_as_uint(n4.y)}, hiz[2] = {__float_as_uint(n4.z), __float_as_uint(n4.w)};
  const uint32_t meta[2] = {__float_as_uint(n1.z), __float_as_uint(n1.w)};
  uint32_t hitmask = 0;
#pragma unroll
  for (int h = 0; h < 2; h++) {
    const uint32_t m4 = meta[h];
    const uint32_t inner4 = (m4 & (m4 << 1)) & 0x10101010u;
    const uint32_t imask4 = sext_s8x4(inner4 << 3);
    const uint32_t bit4 = (m4 ^ (r.octinv4 & imask4)) & 0x1f1f1f1fu;
    const uint32_t cb4 = (m4 >> 5) & 0x07070707u;
    const uint32_t nearx = nx ? hix[h] : lox[h], farx = nx ? lox[h] : hix[h];
    const uint32_t neary = ny ? hiy[h] : loy[h], fary = ny ? loy[h] : hiy[h];
    const uint32_t nearz = nz ? hiz[h] : loz[h], farz = nz ? loz[h] : hiz[h];
#pragma unroll
    for (int j = 0; j < 4; j++) {
      const int sh = 8 * j;
      const float t0x = fmaf((float)((nearx >> sh) & 0xff), ax, bx), t1x = fmaf((float)((farx >> sh) & 0xff), ax, bx);
      const float t0y = fmaf((float)((neary >> sh) & 0xff), ay, by), t1y = fmaf((float)((fary >> sh) & 0xff), ay, by);
      const float t0z = fmaf((float)((nearz >> sh) & 0xff), az, bz), t1z = fmaf((float)((farz >> sh) & 0xff), az, bz);
      const float cmin = fmaxf(fmaxf(t0x, t0y), fmaxf(t0z, tmin));
      const float cmax = fminf(fminf(t1x, t1y), fminf(t1z, tmax)) * 1.0000005f; // conservative against rounding
      if (cmin <= cmax) hitmask |= ((cb4 >> sh) & 0xffu) << ((bit4 >> sh) & 0xffu);
    }
  }
  ng.x = __float_as_uint(n1.x);
  tg.x = __float_as_uint(n1.y);
  ng.y = (hitmask & 0xff000000u) | (eim >> 24);
  tg.y = hitmask & 0x00ffffffu;
}

// Triangle::intersect_ray (triangle.rs:124-147, Q4: all three edges inclusive, two-sided, det != 0 only).
SPT_D bool tri_test(const float4* __restrict__ tris, uint32_t tri_index, float3 o, float3 d, float& t, float& v, float& w, int& tri_id) {
  const float4* tp = tris + (size_t)tri_index * 3;
  const float4 a = ldg4(tp), b = ldg4(tp + 1), c = ldg4(tp + 2);
  const float3 p0 = mk3(a), e1 = xsub3(mk3(b), p0), e2 = xsub3(mk3(c), p0);
  const float3 q = xcross3(d, e2);
  float det = xdot3(e1, q);
  if (det != 0.0f) {
    det = __fdiv_rn(1.0f, det);
    const float3 s = xsub3(o, p0);
    v = xmul(xdot3(s, q), det);
    if (v >= 0.0f) {
      const float3 r = xcross3(s, e1);
      w = xmul(xdot3(d, r), det);
      const float u = xsub(xsub(1.0f, v), w);
      if (w >= 0.0f && u >= 0.0f) {
        t = xmul(xdot3(e2, r), det);
        tri_id = __float_as_int(a.w);
        return true;
      }
    }
  }
  return false;
}

// Sphere::intersect_ray (sphere.rs:25-39)
SPT_D bool sphere_roots(float4 cr, float3 o, float3 d, float& mn, float& mx) {
  const float3 oc = xsub3(o, mk3(cr));
  const float a = xdot3(d, d), b = xdot3(d, oc);
  const float c = xsub(xdot3(oc, oc), xmul(cr.w, cr.w));
  float delta = xsub(xmul(b, b), xmul(a, c));
  if (delta >= 0.0f) {
    delta = __fsqrt_rn(delta);
    mn = __fdiv_rn(xsub(-b, delta), a);
    mx = __fdiv_rn(xadd(-b, delta), a);
    return true;
  }
  return false;
}

// ANY = false: closest hit in (tmin, hit.t) -> fills hit, returns whether something was hit.
// ANY = true : returns true as soon as any primitive is hit in (tmin, tmax = hit.t).
template <bool ANY, bool COUNT>
__device__ bool traverse(const DScene& sc, float3 ow, float3 dw, float tmin, HitRec& hit, TravCounters* cnt) {
  if (sc.tlas_empty) return false;
  uint2 stack[SPT_STACK];
  int sp = 0, blas_base = 0, cur_inst = -1;
  bool found = false;
  RaySetup r;
  ray_setup(r, ow, dw);
  uint2 ng = make_uint2(0u, 0x80000000u), tg = make_uint2(0u, 0u);
  for (;;) {
    if (ng.y > 0x00ffffffu) {
      const uint32_t bit = bfind_u32(ng.y);
      const uint32_t base = ng.x, imask = ng.y;
      ng.y &= ~(1u << bit);
      if (ng.y > 0x00ffffffu) stack[sp++] = ng;
      const uint32_t slot = (bit - 24u) ^ (r.octinv4 & 0xffu);
      const uint32_t rel = __popc(imask & ~(0xffffffffu << slot));
      if (COUNT) cnt->nodes++;
      node_children(sc.nodes, base + rel, r, tmin, hit.t, ng, tg);
    } else {
      tg = ng;
      ng = make_uint2(0u, 0u);
    }
    while (tg.y != 0u) {
      const uint32_t k = bfind_u32(tg.y);
      tg.y &= ~(1u << k);
      const uint32_t pidx = tg.x + k;
      if (cur_inst < 0) { // TLAS leaf: an instance (instance.rs:89-109)
        const int iid = sc.tlas_inst[pidx];
        const float4* ip = reinterpret_cast<const float4*>(sc.inst + iid);
        const float4 r0 = ldg4(ip), r1 = ldg4(ip + 1), r2 = ldg4(ip + 2), mt = ldg4(ip + 3);
        if (COUNT) cnt->insts++;
        const float3 o2 = xmat_point(r0, r1, r2, ow), d2 = xmat_vec(r0, r1, r2, dw);
        const int kind = __float_as_int(mt.x), root = __float_as_int(mt.y);
        if (kind == 0) { // sphere
          if (COUNT) cnt->tris++;
          float mn, mx;
          if (sphere_roots(ldg4(sc.spheres + root), o2, d2, mn, mx)) {
            if (ANY) { if (mn < hit.t && mx > tmin) return true; } // sphere.rs:51-57
            else {
              const float t = mn < tmin ? mx : mn;                 // sphere.rs:59-63
              if (tmin < t && t < hit.t) { hit.t = t; hit.u = 0.f; hit.v = 0.f; hit.prim = 0; hit.inst = iid; found = true; }
            }
          }
        } else { // enter the BLAS with the object-space ray
          if (tg.y != 0u) stack[sp++] = tg;
          if (ng.y > 0x00ffffffu) stack[sp++] = ng;
          blas_base = sp; cur_inst = iid;
          ray_setup(r, o2, d2);
          ng = make_uint2((uint32_t)root, 0x80000000u);
          tg = make_uint2(0u, 0u);
          break;
        }
      } else {
        if (COUNT) cnt->tris++;
        float t, v, w; int tid;
        if (tri_test(sc.tris, pidx, r.o, r.d, t, v, w, tid)) {
          if (t > tmin && t < hit.t) { // triangle.rs:177-187
            if (ANY) return true;
            hit.t = t; hit.u = v; hit.v = w; hit.prim = tid; hit.inst = cur_inst; found = true;
          }
        }
      }
    }
    if (ng.y <= 0x00ffffffu) {
      if (cur_inst >= 0 && sp == blas_base) { cur_inst = -1; ray_setup(r, ow, dw); }
      if (sp == 0) break;
      ng = stack[--sp];
    }
  }
  return found;
}

// sphere_normal_to_texcoords (sphere.rs:138-145)
SPT_D float2 sphere_uv(float3 n) {
  const float theta = acosf(n.y), phi = atan2f(n.x, n.z) + SPT_PI;
  return make_float2(phi * 0.5f * SPT_INV_PI, theta * SPT_INV_PI);
}
