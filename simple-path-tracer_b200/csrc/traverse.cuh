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
// Traversal: one thread per ray, node groups (base index + hit bitmask) on a short per-thread stack kept in SHARED
// memory (overflow in local memory), children visited in octant order, 5 x 16-byte loads per node, 3 x 16-byte
// loads per triangle, 4 per instance.  Warps are PERSISTENT: a lane whose ray is finished pulls the next ray from a
// warp-private chunk of the launch's ray range (one global atomic per 256 rays), so short and long rays do not
// leave lanes idle; all lanes of a warp advance in rounds of exactly one node step followed by the leaf work that
// node produced (triangles / instances), which keeps the 200-instruction node step converged across the warp.
#pragma once
#include "launch.h"
#include "vecmath.cuh"

struct HitRec {
  float t, u, v;
  int prim, inst; // prim: triangle id inside its mesh (file order) or 0 for a sphere; inst: instance id, -1 = miss
  int tri;        // global index of the triangle in `tris` / `shade_tris` (leaf order), -1 for a sphere
};
struct TravCounters { // only maintained by the COUNT instantiations
  uint32_t nodes, tris, insts;
  // warp-round profile of the persistent driver (per lane): rounds spent with a ray, and lanes active in each stage;
  // lane 0 also counts the rounds of its warp and how many rounds executed each stage at all
  uint32_t busy, pop, node, inst, tri, w_rounds, w_pop, w_node, w_inst, w_tri;
};

#ifndef SPT_SSTACK
#define SPT_SSTACK 16       // stack entries per thread in shared memory
#endif
#ifndef SPT_WORLD_STASH
#define SPT_WORLD_STASH 1   // 1: while a lane is inside a BLAS its world-space ray lives in shared memory, not in 10 registers
#endif
#define SPT_WSTASH (SPT_WORLD_STASH ? 5 : 0) // extra uint2 rows of the shared stack that hold the stashed world ray
#define SPT_LSTACK 32       // overflow entries in local memory (untouched unless a ray goes deeper than SPT_SSTACK)
#ifndef SPT_TRAV_PREFETCH
#define SPT_TRAV_PREFETCH 0 // prefetching the next child node after a node step: measured -7 % (C2), kept as a tuning knob
#endif
#ifndef SPT_NODE_PER_ROUND
#define SPT_NODE_PER_ROUND 1
#endif
#ifndef SPT_TRI_FRCP
#define SPT_TRI_FRCP 1
#endif
#ifndef SPT_INST_UNIFORM_SETUP
#define SPT_INST_UNIFORM_SETUP 1 // 1: the object-space ray setup of an instance visit runs for small meshes too (no divergent branch)
#endif
#ifndef SPT_FETCH_MIN_IDLE
#define SPT_FETCH_MIN_IDLE 6 // refill the warp once this many lanes are without a ray (1 / 3 re-measured with the pop stage at the end of the round: -0.5 % / 0 %)
#endif
#ifndef SPT_CHUNK_PREFETCH
#define SPT_CHUNK_PREFETCH 1 // 1: a warp that reserves a chunk prefetches the chunk's rays into L2 (extend -2.7 %, frame +1 % on C2; 0 = off)
#endif
#ifndef SPT_POP_END
#define SPT_POP_END 1        // 1: the pop stage closes the round instead of opening it (see trace_persistent)
#endif
#ifndef SPT_TRI_HELP
#define SPT_TRI_HELP 0       // 1: ONE triangle pass per round in which lanes without triangle work test the second pending triangle of lanes that have
                             // two.  Measured on B200 (profiles/README.md, round 2): -10 % on C2 against two plain triangle stages — the pairing
                             // (3 ballots, a shared-memory map, 11 shuffles, 72 instead of 64 registers) costs more than the lanes it fills return.
#endif

SPT_D uint32_t sext_s8x4(uint32_t x) { uint32_t v; asm("prmt.b32 %0, %1, 0x0, 0x0000BA98;" : "=r"(v) : "r"(x)); return v; }
SPT_D uint32_t bfind_u32(uint32_t x) { uint32_t v; asm("bfind.u32 %0, %1;" : "=r"(v) : "r"(x)); return v; }
SPT_D float4 ldg4(const float4* p) { return __ldg(p); }

struct RaySetup { float3 o, d, idir; uint32_t octinv4; };
SPT_D float rcp_approx(float x) { float y; asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x)); return y; } // MUFU.RCP, 1 ulp
// |d| clamped away from zero (keeps its sign) so that 1/d stays finite; 1-ulp reciprocal: only the conservative box
// tests use idir.  Octant from the sign bits.
SPT_D float clamp_dir(float d) { return __uint_as_float((__float_as_uint(fmaxf(fabsf(d), 1e-18f)) & 0x7fffffffu) | (__float_as_uint(d) & 0x80000000u)); }
SPT_D void ray_setup(RaySetup& r, float3 o, float3 d) {
  r.o = o; r.d = d;
  r.idir = mk3(rcp_approx(clamp_dir(d.x)), rcp_approx(clamp_dir(d.y)), rcp_approx(clamp_dir(d.z)));
  const uint32_t oct = ((__float_as_uint(d.x) >> 31) << 2) | ((__float_as_uint(d.y) >> 31) << 1) | (__float_as_uint(d.z) >> 31);
  r.octinv4 = (7u - oct) * 0x01010101u;
}

// float(byte j of x).  I2F.U8 runs on the XU pipe (4 lanes/clk per scheduler: ncu showed XU at 50 % of its peak with 48
// conversions per node step); SPT_BYTE2F_MAGIC=1 builds 2^23 + q with one PRMT and subtracts 2^23 (ALU + FMA pipes, exact).
// SPT_BYTE2F_BIAS=1 (default) avoids the conversion altogether: one PRMT drops byte j into mantissa bits 8..15 of 1.0f, giving
// 1 + q * 2^-15 exactly; node_children() folds the 2^15 into the plane scale and the "1 +" into the plane offset, so the
// slab FFMA absorbs both (48 XU instructions per node step become 48 ALU ones; see node_children for the error bound).
#ifndef SPT_BYTE2F_MAGIC
#define SPT_BYTE2F_MAGIC 0
#endif
#ifndef SPT_BYTE2F_BIAS
#define SPT_BYTE2F_BIAS 0
#endif
SPT_D float byte_to_float(uint32_t x, int j) { // j is a compile-time constant after unrolling
#if SPT_BYTE2F_BIAS
  return __uint_as_float(__byte_perm(x, 0x3F800000u, 0x7604u | ((uint32_t)j << 4)));
#elif SPT_BYTE2F_MAGIC
  return __uint_as_float(__byte_perm(x, 0x4B000000u, 0x7650u | (uint32_t)j)) - 8388608.0f;
#else
  return (float)((x >> (8 * j)) & 0xff);
#endif
}

// Blackwell packed single precision: one FFMA2 does the near- and the far-plane FMA of an axis (fma.rn.f32x2, sm_100+; each half is
// rounded like the scalar FFMA, so the result bits are the same).  SPT_FFMA2=1 uses it in the slab test: 48 FFMA per node step become 24.
#ifndef SPT_FFMA2
#define SPT_FFMA2 0
#endif
SPT_D unsigned long long pack2f(float lo, float hi) { unsigned long long r; asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi)); return r; }
SPT_D void ffma2(float& d0, float& d1, float a0, float a1, unsigned long long b, unsigned long long c) {
  unsigned long long d;
  asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(pack2f(a0, a1)), "l"(b), "l"(c));
  asm("mov.b64 {%0, %1}, %2;" : "=f"(d0), "=f"(d1) : "l"(d));
}

// Slab test of the 8 quantised child boxes of one node; returns the node-group / primitive-group pair.
SPT_D void node_children(const float4* __restrict__ nodes, uint32_t node_index, const RaySetup& r, float tmin, float tmax,
                         uint2& ng, uint2& tg) {
  const float4* np = nodes + (size_t)node_index * 5;
  const float4 n0 = ldg4(np), n1 = ldg4(np + 1), n2 = ldg4(np + 2), n3 = ldg4(np + 3), n4 = ldg4(np + 4);
  const uint32_t eim = __float_as_uint(n0.w);
#if SPT_BYTE2F_BIAS
  // byte_to_float() yields v = 1 + q * 2^-15, so with A = 2^15 * s / d and B = (lo - o) / d - A the plane distance q * s / d +
  // (lo - o) / d is v * A + B.  The rounding of B and of the FFMA costs at most ~2^-8 of one quantisation step (|A| * 2^-23)
  // when the origin is within a few node widths, and the usual relative 2^-24 otherwise; the near planes are pulled in and
  // the far planes pushed out by |A| * 2^-21 (2^-6 of a step) per axis, which keeps the test conservative.
  const float sx = __uint_as_float((uint32_t)((int)(int8_t)(eim & 0xff) + 142) << 23);
  const float sy = __uint_as_float((uint32_t)((int)(int8_t)((eim >> 8) & 0xff) + 142) << 23);
  const float sz = __uint_as_float((uint32_t)((int)(int8_t)((eim >> 16) & 0xff) + 142) << 23);
  const float ax = sx * r.idir.x, ay = sy * r.idir.y, az = sz * r.idir.z;
  const float cx = fmaf(n0.x - r.o.x, r.idir.x, -ax), cy = fmaf(n0.y - r.o.y, r.idir.y, -ay), cz = fmaf(n0.z - r.o.z, r.idir.z, -az);
  const float bx = fmaf(fabsf(ax), -4.76837158e-7f, cx), by = fmaf(fabsf(ay), -4.76837158e-7f, cy), bz = fmaf(fabsf(az), -4.76837158e-7f, cz);
  const float fx = fmaf(fabsf(ax), 4.76837158e-7f, cx), fy = fmaf(fabsf(ay), 4.76837158e-7f, cy), fz = fmaf(fabsf(az), 4.76837158e-7f, cz);
#else
  const float sx = __uint_as_float((uint32_t)((int)(int8_t)(eim & 0xff) + 127) << 23);
  const float sy = __uint_as_float((uint32_t)((int)(int8_t)((eim >> 8) & 0xff) + 127) << 23);
  const float sz = __uint_as_float((uint32_t)((int)(int8_t)((eim >> 16) & 0xff) + 127) << 23);
  const float ax = sx * r.idir.x, ay = sy * r.idir.y, az = sz * r.idir.z;
  const float bx = (n0.x - r.o.x) * r.idir.x, by = (n0.y - r.o.y) * r.idir.y, bz = (n0.z - r.o.z) * r.idir.z;
  const float fx = bx, fy = by, fz = bz;
#endif
  // near/far planes per axis according to the ray direction sign
  const bool nx = r.idir.x < 0.f, ny = r.idir.y < 0.f, nz = r.idir.z < 0.f;
  const uint32_t lox[2] = {__float_as_uint(n2.x), __float_as_uint(n2.y)}, loy[2] = {__float_as_uint(n2.z), __float_as_uint(n2.w)};
  const uint32_t loz[2] = {__float_as_uint(n3.x), __float_as_uint(n3.y)}, hix[2] = {__float_as_uint(n3.z), __float_as_uint(n3.w)};
  const uint32_t hiy[2] = {__float_as_uint(n4.x), __float_as_uint(n4.y)}, hiz[2] = {__float_as_uint(n4.z), __float_as_uint(n4.w)};
  const uint32_t meta[2] = {__float_as_uint(n1.z), __float_as_uint(n1.w)};
  uint32_t hitmask = 0;
#if SPT_FFMA2
  const unsigned long long ax2 = pack2f(ax, ax), ay2 = pack2f(ay, ay), az2 = pack2f(az, az);
  const unsigned long long bx2 = pack2f(bx, fx), by2 = pack2f(by, fy), bz2 = pack2f(bz, fz);
#endif
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
#if SPT_FFMA2
      float t0x, t1x, t0y, t1y, t0z, t1z;
      ffma2(t0x, t1x, byte_to_float(nearx, j), byte_to_float(farx, j), ax2, bx2);
      ffma2(t0y, t1y, byte_to_float(neary, j), byte_to_float(fary, j), ay2, by2);
      ffma2(t0z, t1z, byte_to_float(nearz, j), byte_to_float(farz, j), az2, bz2);
#else
      const float t0x = fmaf(byte_to_float(nearx, j), ax, bx), t1x = fmaf(byte_to_float(farx, j), ax, fx);
      const float t0y = fmaf(byte_to_float(neary, j), ay, by), t1y = fmaf(byte_to_float(fary, j), ay, fy);
      const float t0z = fmaf(byte_to_float(nearz, j), az, bz), t1z = fmaf(byte_to_float(farz, j), az, fz);
#endif
      const float cmin = fmaxf(fmaxf(t0x, t0y), fmaxf(t0z, tmin));
      const float cmax = fminf(fminf(t1x, t1y), fminf(t1z, tmax)) * 1.000002f; // conservative against rounding (incl. the approximate 1/d)
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
  const float3 p0 = mk3(a), e1 = mk3(b), e2 = mk3(c); // edges precomputed by flatten() with the same subtraction
  const float3 q = xcross3(d, e2);
  float det = xdot3(e1, q);
  if (det != 0.0f) {
#if SPT_TRI_FRCP
    det = __frcp_rn(det); // the correctly rounded reciprocal IS the IEEE quotient 1.0 / det (triangle.rs:131), two instructions shorter
#else
    det = __fdiv_rn(1.0f, det);
#endif
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

// ---- per-ray traversal state (registers) + the two-level stack ------------------------------------------------------
struct TravStack {
  uint2* sh;     // &s_stack[0][threadIdx.x]; entry i at sh[i * SPT_TRAV_BLOCK]
  uint2* lo;     // local-memory overflow
  SPT_D void push(int& sp, uint2 e) const { if (sp < SPT_SSTACK) sh[sp * SPT_TRAV_BLOCK] = e; else lo[sp - SPT_SSTACK] = e; sp++; }
  SPT_D uint2 pop(int& sp) const { --sp; return sp < SPT_SSTACK ? sh[sp * SPT_TRAV_BLOCK] : lo[sp - SPT_SSTACK]; }
#if SPT_WORLD_STASH
  SPT_D void stash(const RaySetup& r) const {
    sh[(SPT_SSTACK + 0) * SPT_TRAV_BLOCK] = make_uint2(__float_as_uint(r.o.x), __float_as_uint(r.o.y));
    sh[(SPT_SSTACK + 1) * SPT_TRAV_BLOCK] = make_uint2(__float_as_uint(r.o.z), __float_as_uint(r.d.x));
    sh[(SPT_SSTACK + 2) * SPT_TRAV_BLOCK] = make_uint2(__float_as_uint(r.d.y), __float_as_uint(r.d.z));
    sh[(SPT_SSTACK + 3) * SPT_TRAV_BLOCK] = make_uint2(__float_as_uint(r.idir.x), __float_as_uint(r.idir.y));
    sh[(SPT_SSTACK + 4) * SPT_TRAV_BLOCK] = make_uint2(__float_as_uint(r.idir.z), r.octinv4);
  }
  SPT_D void unstash(RaySetup& r) const {
    const uint2 a = sh[(SPT_SSTACK + 0) * SPT_TRAV_BLOCK], b = sh[(SPT_SSTACK + 1) * SPT_TRAV_BLOCK], c = sh[(SPT_SSTACK + 2) * SPT_TRAV_BLOCK],
                d = sh[(SPT_SSTACK + 3) * SPT_TRAV_BLOCK], e = sh[(SPT_SSTACK + 4) * SPT_TRAV_BLOCK];
    r.o = mk3(__uint_as_float(a.x), __uint_as_float(a.y), __uint_as_float(b.x));
    r.d = mk3(__uint_as_float(b.y), __uint_as_float(c.x), __uint_as_float(c.y));
    r.idir = mk3(__uint_as_float(d.x), __uint_as_float(d.y), __uint_as_float(e.x));
    r.octinv4 = e.y;
  }
#endif
};
struct Trav {
  float tmin;
  RaySetup r;    // the ray of the current space: world (core/ray.rs) in the TLAS, object space inside a BLAS
#if !SPT_WORLD_STASH
  float3 ow, dw; // the world-space ray
  float3 widir; uint32_t woct; // idir / octant of the world-space ray, restored when a BLAS is left
#endif
  HitRec hit;    // hit.t doubles as the running t_max (Intersection::t, intersection.rs:21-26)
  uint2 ng, tg;  // pending node group / primitive group of the current space (either may be empty)
  int sp, blas_base, cur_inst;
};
SPT_D void trav_begin(Trav& T, float3 o, float3 d, float tmin, float tmax) {
  T.tmin = tmin;
  ray_setup(T.r, o, d);
#if !SPT_WORLD_STASH
  T.ow = o; T.dw = d; T.widir = T.r.idir; T.woct = T.r.octinv4;
#endif
  T.hit.t = tmax; T.hit.u = T.hit.v = 0.f; T.hit.prim = -1; T.hit.inst = -1; T.hit.tri = -1; // Intersection::default
  T.ng = make_uint2(0u, 0x80000000u); T.tg = make_uint2(0u, 0u);
  T.sp = 0; T.blas_base = 0; T.cur_inst = -1;
}
SPT_D bool trav_has_node(const Trav& T) { return T.ng.y > 0x00ffffffu; }
SPT_D bool trav_has_leaf(const Trav& T) { return T.tg.y != 0u; }

// Node step: take the nearest pending child of T.ng, intersect its 8 children -> new T.ng / T.tg.
template <bool COUNT>
SPT_D void trav_node(const DScene& sc, Trav& T, const TravStack& st, TravCounters& cnt) {
  const uint32_t bit = bfind_u32(T.ng.y);
  const uint32_t base = T.ng.x, imask = T.ng.y;
  T.ng.y &= ~(1u << bit);
  if (T.ng.y > 0x00ffffffu) st.push(T.sp, T.ng);
  const uint32_t slot = (bit - 24u) ^ (T.r.octinv4 & 0xffu);
  const uint32_t rel = __popc(imask & ~(0xffffffffu << slot));
  if (COUNT) cnt.nodes++;
  node_children(sc.nodes, base + rel, T.r, T.tmin, T.hit.t, T.ng, T.tg);
#if SPT_TRAV_PREFETCH
  // the child this lane will step into next round (or the first triangle it will test this round): start the fetch now,
  // the instance / triangle stages of the other lanes run in between
  if (T.ng.y > 0x00ffffffu) {
    const uint32_t nb = bfind_u32(T.ng.y), ns = (nb - 24u) ^ (T.r.octinv4 & 0xffu);
    const float4* np = sc.nodes + (size_t)(T.ng.x + __popc(T.ng.y & ~(0xffffffffu << ns))) * 5;
    asm volatile("prefetch.global.L1 [%0];" ::"l"(np));
    asm volatile("prefetch.global.L1 [%0];" ::"l"(np + 4));
  }
#endif
}
// One triangle of T.tg against the object-space ray (BLAS leaf or small mesh).  Returns true when ANY found an occluder.
template <bool ANY, bool COUNT>
SPT_D bool trav_tri(const DScene& sc, Trav& T, TravCounters& cnt) {
  const uint32_t k = bfind_u32(T.tg.y);
  T.tg.y &= ~(1u << k);
  const uint32_t pidx = T.tg.x + k;
  if (COUNT) cnt.tris++;
  float t, v, w; int tid;
  if (tri_test(sc.tris, pidx, T.r.o, T.r.d, t, v, w, tid) && t > T.tmin && t < T.hit.t) { // triangle.rs:177-187
    if (ANY) { T.hit.inst = 0; return true; }
    T.hit.t = t; T.hit.u = v; T.hit.v = w; T.hit.prim = tid; T.hit.tri = (int)pidx; T.hit.inst = T.cur_inst;
  }
  return false;
}
// One instance of T.tg (a TLAS leaf, instance.rs:89-109): sphere test, or switch to the object-space ray and queue the
// instance's BLAS root (mesh) or its few triangles (small mesh).  Returns true when ANY found an occluder.
template <bool ANY, bool COUNT>
SPT_D bool trav_inst(const DScene& sc, Trav& T, const TravStack& st, TravCounters& cnt) {
  const uint32_t k = bfind_u32(T.tg.y);
  T.tg.y &= ~(1u << k);
  const int iid = sc.tlas_inst[T.tg.x + k];
  const float4* ip = reinterpret_cast<const float4*>(sc.inst + iid);
  const float4 r0 = ldg4(ip), r1 = ldg4(ip + 1), r2 = ldg4(ip + 2), mt = ldg4(ip + 3);
  if (COUNT) cnt.insts++;
  const float3 o2 = xmat_point(r0, r1, r2, T.r.o), d2 = xmat_vec(r0, r1, r2, T.r.d); // instances are only visited from the TLAS: r is the world ray
  const int kind = __float_as_int(mt.x), root = __float_as_int(mt.y);
  if (kind == 0) { // sphere
    if (COUNT) cnt.tris++;
    float mn, mx;
    if (sphere_roots(ldg4(sc.spheres + root), o2, d2, mn, mx)) {
      if (ANY) { if (mn < T.hit.t && mx > T.tmin) { T.hit.inst = 0; return true; } } // sphere.rs:51-57
      else {
        const float t = mn < T.tmin ? mx : mn;                                      // sphere.rs:59-63
        if (T.tmin < t && t < T.hit.t) { T.hit.t = t; T.hit.u = 0.f; T.hit.v = 0.f; T.hit.prim = 0; T.hit.tri = -1; T.hit.inst = iid; }
      }
    }
    return false;
  }
#if SPT_SMALL_INLINE
  if (kind & 2) { // small mesh: test its few triangles right here with the object-space ray; the TLAS state is untouched
    const int n = kind >> 8;
    for (int j = 0; j < n; j++) {
      if (COUNT) cnt.tris++;
      float t, v, w; int tid;
      if (tri_test(sc.tris, (uint32_t)(root + j), o2, d2, t, v, w, tid) && t > T.tmin && t < T.hit.t) {
        if (ANY) { T.hit.inst = 0; return true; }
        T.hit.t = t; T.hit.u = v; T.hit.v = w; T.hit.prim = tid; T.hit.tri = root + j; T.hit.inst = iid;
      }
    }
    return false;
  }
#endif
  if (T.tg.y != 0u) st.push(T.sp, T.tg);
  if (T.ng.y > 0x00ffffffu) st.push(T.sp, T.ng);
  T.blas_base = T.sp; T.cur_inst = iid;
#if SPT_WORLD_STASH
  st.stash(T.r);
#endif
#if SPT_INST_UNIFORM_SETUP
  // small mesh (kind & 2): its triangles become the pending primitive group, no BLAS node step.  The reciprocal direction is
  // computed for it all the same: three MUFUs cost less than the divergent branch around them.
  const bool small = (kind & 2) != 0;
  ray_setup(T.r, o2, d2);
  T.ng = make_uint2(small ? 0u : (uint32_t)root, small ? 0u : 0x80000000u);
  T.tg = make_uint2(small ? (uint32_t)root : 0u, small ? (1u << (kind >> 8)) - 1u : 0u);
  return false;
#endif
  if (kind & 2) { // small mesh: its triangles become the pending primitive group, no BLAS node step
    T.r.o = o2; T.r.d = d2;
    T.ng = make_uint2(0u, 0u);
    T.tg = make_uint2((uint32_t)root, (1u << (kind >> 8)) - 1u);
  } else {
    ray_setup(T.r, o2, d2);
    T.ng = make_uint2((uint32_t)root, 0x80000000u);
    T.tg = make_uint2(0u, 0u);
  }
  return false;
}
// Called when neither a node nor a primitive group is pending: leave the BLAS if it is exhausted, pop the next group.
// Returns true when the ray is finished.
SPT_D bool trav_pop(Trav& T, const TravStack& st) {
#if SPT_WORLD_STASH
  if (T.cur_inst >= 0 && T.sp == T.blas_base) { T.cur_inst = -1; st.unstash(T.r); }
#else
  if (T.cur_inst >= 0 && T.sp == T.blas_base) { T.cur_inst = -1; T.r.o = T.ow; T.r.d = T.dw; T.r.idir = T.widir; T.r.octinv4 = T.woct; }
#endif
  if (T.sp == 0) return true;
  const uint2 e = st.pop(T.sp);
  if (e.y > 0x00ffffffu) { T.ng = e; T.tg = make_uint2(0u, 0u); } else { T.tg = e; T.ng = make_uint2(0u, 0u); }
  return false;
}

// Single-ray query (used inside shading for the BSSRDF probe).
// ANY = false: closest hit in (tmin, hit.t) -> hit.  ANY = true: any primitive in (tmin, tmax = hit.t).
template <bool ANY, bool COUNT>
SPT_D bool traverse(const DScene& sc, float3 ow, float3 dw, float tmin, HitRec& hit, TravCounters& cnt, const TravStack& st) {
  if (sc.tlas_empty) return false;
  Trav T;
  trav_begin(T, ow, dw, tmin, hit.t);
  for (;;) {
    if (trav_has_node(T)) trav_node<COUNT>(sc, T, st, cnt);
    bool stop = false;
    while (T.tg.y != 0u && !stop) stop = T.cur_inst >= 0 ? trav_tri<ANY, COUNT>(sc, T, cnt) : trav_inst<ANY, COUNT>(sc, T, st, cnt);
    if (stop) break;
    if (!trav_has_node(T) && T.tg.y == 0u && trav_pop(T, st)) break;
  }
  hit = T.hit;
  return hit.inst >= 0;
}

// The same single-ray query run WARP-SYNCHRONOUSLY by the lanes `m` that reach it together (the BSSRDF probe inside k_shade<7>): the
// lanes advance in the fixed-stage rounds of the persistent driver below — node step, instance visit, triangle test x2, pop — instead of
// drifting through the loops of traverse() one by one under independent thread scheduling (ncu round 1: 3.8 of 32 lanes in such loops).
// Every lane of `m` must call it; a lane whose ray is finished idles until the last one is done.  Same hits as traverse().
template <bool ANY, bool COUNT>
SPT_D bool traverse_staged(unsigned m, const DScene& sc, float3 ow, float3 dw, float tmin, HitRec& hit, TravCounters& cnt, const TravStack& st) {
  Trav T;
  bool has = !sc.tlas_empty;
  trav_begin(T, ow, dw, tmin, hit.t);
  while (__any_sync(m, has)) {
    bool fin = false;
    if (has && trav_has_node(T) && !trav_has_leaf(T)) trav_node<COUNT>(sc, T, st, cnt);
    if (has && trav_has_leaf(T) && T.cur_inst < 0) fin = trav_inst<ANY, COUNT>(sc, T, st, cnt);
#pragma unroll
    for (int k = 0; k < 2; k++)
      if (has && !fin && trav_has_leaf(T) && T.cur_inst >= 0) fin = trav_tri<ANY, COUNT>(sc, T, cnt);
    if (has && !fin && !trav_has_node(T) && !trav_has_leaf(T)) fin = trav_pop(T, st);
    if (fin) has = false;
  }
  hit = T.hit;
  return hit.inst >= 0;
}

// Persistent-warp driver.  Src provides
//   uint32_t chunk() const;                                            candidates a warp reserves at a time (one atomic); warp-uniform
//   uint32_t total() const;                                            number of candidate rays of this launch
//   bool load(uint32_t i, float3& o, float3& d, float& tmin, float& tmax);   false: nothing to trace at i
//   void store(uint32_t i, const HitRec& h);                           closest hit (inst = -1: miss) / ANY: inst = 0 occluded
// `next` is a zero-initialised device counter of this launch.  Must be called by all threads of a block of
// SPT_TRAV_BLOCK threads (uses a static shared-memory stack).
//
// Every round the whole warp passes through the same fixed chain of stages, each executed once by all the lanes that
// have that kind of work pending, in the order a single ray needs them:
//   refill -> node step -> instance visit -> triangle pass -> pop (stack / leave BLAS / finish)
// so a lane can advance through several stages in one round, no lane ever waits for another lane's leaf work, and
// lanes doing the same kind of work execute it together instead of drifting apart under independent thread scheduling.
// The pop stage CLOSES the round (SPT_POP_END): a ray whose last test ran in this round finds its stack empty now and gives its
// lane back for the next round's refill; opening the round with it (round 1) cost every ray one whole round at the end of its life
// in which its lane did nothing else (tools/trav_sim.py: 4.6 -> 3.6 lane-rounds per ray on C2).
// Triangle pass variant (SPT_TRI_HELP, off: measured slower): every lane with pending triangles tests its next one; lanes that have
// NO triangle work in this round each take the second pending triangle of the i-th lane that has two or more — the ray goes over by
// shuffles, the result comes back the same way and is applied after the owner's own, i.e. in the order a sequential loop would.
// A lane drains its pending primitive group before it takes its next node step (same order as a depth-first loop).
#ifndef SPT_TRI_PER_ROUND
#define SPT_TRI_PER_ROUND (SPT_TRI_HELP ? 1 : 2)
#endif
#ifndef SPT_INST_PER_ROUND
#define SPT_INST_PER_ROUND 1
#endif
#ifndef SPT_STAGE_ORDER
#define SPT_STAGE_ORDER 0   // 1: an extra instance stage before the node step (a BLAS entered there gets its root step in the same round)
#endif
#ifndef SPT_SMALL_INLINE
#define SPT_SMALL_INLINE 0  // 1: small meshes are tested inside the instance stage instead of becoming a pending triangle group
#endif
template <bool ANY, bool COUNT, class Src>
SPT_D void trace_persistent(const DScene& sc, Src& src, uint32_t* next, TravCounters& cnt) {
  __shared__ uint2 s_stack[SPT_SSTACK + SPT_WSTASH][SPT_TRAV_BLOCK];
#if SPT_TRI_HELP
  __shared__ uint32_t s_pair[SPT_TRAV_BLOCK / 32][64]; // per warp: the lane pairing of the triangle pass
#endif
  uint2 l_stack[SPT_LSTACK];
  TravStack st; st.sh = &s_stack[0][threadIdx.x]; st.lo = l_stack;
  const uint32_t total = src.total(), lane = threadIdx.x & 31u, FULL = 0xffffffffu;
  // The FIRST chunk of every warp is dealt statically by its index in the grid, further chunks come from the atomic cursor (which
  // therefore starts behind the static ones).  A launch with little work — a draining pool — then needs no atomic at all, and warps
  // beyond the work leave at once: 4736 warps hitting one address with atomicAdd was a ~50 us floor under every such launch.
  const uint32_t n_warps = gridDim.x * (SPT_TRAV_BLOCK / 32), gwarp = blockIdx.x * (SPT_TRAV_BLOCK / 32) + (threadIdx.x >> 5);
  const uint32_t static_end = n_warps * src.chunk();
  uint32_t w_next = min(gwarp * src.chunk(), total), w_end = min(w_next + src.chunk(), total), idx = 0;
  bool exhausted = false, has = false;
  Trav T;
  T.ng = T.tg = make_uint2(0u, 0u); T.cur_inst = -1; T.sp = 0;
  for (;;) {
    // ---- refill
    const uint32_t idle = __ballot_sync(FULL, !has);
    if (idle == FULL || __popc(idle) >= SPT_FETCH_MIN_IDLE) {
      if (w_next >= w_end && !exhausted) { // next chunk of this warp
        uint32_t base = 0;
        if (static_end >= total) base = total; // all the work was dealt statically
        else {
          if (lane == 0) base = atomicAdd(next, src.chunk()) + static_end;
          base = __shfl_sync(FULL, base, 0);
        }
        if (base >= total) exhausted = true;
        else {
          w_next = base; w_end = min(base + src.chunk(), total);
#if SPT_CHUNK_PREFETCH
          for (uint32_t j = base + lane; j < w_end; j += 32u) src.prefetch(j); // the chunk's rays on their way to L2 before the refills ask for them
#endif
        }
      }
      const uint32_t avail = w_end - w_next; // w_end >= w_next always
      const uint32_t take = min((uint32_t)__popc(idle), avail);
      if (take == 0 && idle == FULL) break;  // nothing in flight and nothing left (exhausted)
      const uint32_t rank = __popc(idle & ((1u << lane) - 1u));
      if (!has && rank < take) {
        idx = w_next + rank;
        float3 o, d; float tmin, tmax;
        has = src.load(idx, o, d, tmin, tmax);
        if (has) {
          if (sc.tlas_empty) { HitRec h; h.t = tmax; h.u = h.v = 0.f; h.prim = h.inst = h.tri = -1; src.store(idx, h); has = false; }
          else trav_begin(T, o, d, tmin, tmax);
        }
      }
      w_next += take;
    }
    bool fin = false;
    if (COUNT) { cnt.busy += has ? 1u : 0u; if (lane == 0) cnt.w_rounds++; }
#if !SPT_POP_END
    // ---- pop: nothing pending -> leave the BLAS / next stack entry / ray finished
    {
      const bool p = has && !trav_has_node(T) && !trav_has_leaf(T);
      if (COUNT) { const uint32_t m = __ballot_sync(FULL, p); cnt.pop += p ? 1u : 0u; if (lane == 0 && m) cnt.w_pop++; }
      if (p) fin = trav_pop(T, st);
    }
#endif
#if SPT_STAGE_ORDER == 1
    if (has && !fin && trav_has_leaf(T) && T.cur_inst < 0) fin = trav_inst<ANY, COUNT>(sc, T, st, cnt);
#endif
    // ---- node step
#pragma unroll
    for (int k = 0; k < SPT_NODE_PER_ROUND; k++) {
      const bool p = has && !fin && trav_has_node(T) && !trav_has_leaf(T);
      if (COUNT) { const uint32_t m = __ballot_sync(FULL, p); cnt.node += p ? 1u : 0u; if (lane == 0 && m) cnt.w_node++; }
      if (p) trav_node<COUNT>(sc, T, st, cnt);
    }
    // ---- instance visit (TLAS leaf)
#pragma unroll
    for (int k = 0; k < SPT_INST_PER_ROUND; k++) {
      const bool p = has && !fin && trav_has_leaf(T) && T.cur_inst < 0;
      if (COUNT) { const uint32_t m = __ballot_sync(FULL, p); cnt.inst += p ? 1u : 0u; if (lane == 0 && m) cnt.w_inst++; }
      if (p) fin = trav_inst<ANY, COUNT>(sc, T, st, cnt);
    }
    // ---- triangle tests (BLAS leaf / small mesh)
#if SPT_TRI_HELP
#pragma unroll
    for (int k = 0; k < SPT_TRI_PER_ROUND; k++) {
      const bool own = has && !fin && trav_has_leaf(T) && T.cur_inst >= 0;
      const uint32_t m_own = __ballot_sync(FULL, own);
      if (m_own != 0u) { // warp-uniform
        uint32_t k1 = 0, k2 = 0;
        if (own) { k1 = bfind_u32(T.tg.y); T.tg.y &= ~(1u << k1); }
        const bool rich = own && T.tg.y != 0u;                 // a second triangle is pending
        if (rich) k2 = bfind_u32(T.tg.y);
        const uint32_t m_rich = __ballot_sync(FULL, rich), m_poor = ~m_own, lt = (1u << lane) - 1u;
        const uint32_t n_pairs = min(__popc(m_rich), __popc(m_poor)); // the i-th lane without triangle work helps the i-th rich lane
        const uint32_t r_rich = __popc(m_rich & lt), r_poor = __popc(m_poor & lt);
        const bool helped = rich && r_rich < n_pairs, helper = !own && r_poor < n_pairs;
        uint32_t* map = s_pair[threadIdx.x >> 5];              // [0..31]: rank -> rich lane, [32..63]: rank -> helper lane
        if (n_pairs) {
          if (helped) map[r_rich] = lane;
          if (helper) map[32 + r_poor] = lane;
          __syncwarp();
        }
        const uint32_t from = helper ? map[r_poor] : lane;     // whose ray this lane tests
        const uint32_t back = helped ? map[32 + r_rich] : lane; // who tests this lane's second triangle
        if (helped) T.tg.y &= ~(1u << k2);
        float3 o = T.r.o, d = T.r.d;
        uint32_t pidx = T.tg.x + k1;
        if (n_pairs) { // hand the rays (and the second triangle's index) over; own lanes read themselves
          const uint32_t second = T.tg.x + k2;
          o.x = __shfl_sync(FULL, T.r.o.x, from); o.y = __shfl_sync(FULL, T.r.o.y, from); o.z = __shfl_sync(FULL, T.r.o.z, from);
          d.x = __shfl_sync(FULL, T.r.d.x, from); d.y = __shfl_sync(FULL, T.r.d.y, from); d.z = __shfl_sync(FULL, T.r.d.z, from);
          const uint32_t p2 = __shfl_sync(FULL, second, from);
          if (helper) pidx = p2;
          __syncwarp(); // map is rewritten by the next pass / round
        }
        float t = __int_as_float(0x7fc00000), v = 0.f, w = 0.f; int tid = -1; // NaN = no hit: fails t > tmin for every tmin
        if (own || helper) {
          if (COUNT) cnt.tris++;
          float th;
          if (tri_test(sc.tris, pidx, o, d, th, v, w, tid)) t = th;
        }
        if (COUNT) { const uint32_t m = __ballot_sync(FULL, own || helper); cnt.tri += (own || helper) ? 1u : 0u; if (lane == 0 && m) cnt.w_tri++; }
        float t2 = __int_as_float(0x7fc00000), v2 = 0.f, w2 = 0.f; int tid2 = -1;
        if (n_pairs) { // results of the helpers back to the owners
          t2 = __shfl_sync(FULL, t, back); v2 = __shfl_sync(FULL, v, back); w2 = __shfl_sync(FULL, w, back); tid2 = __shfl_sync(FULL, tid, back);
        }
        if (own) { // the owner's own triangle first, then the helped one: the order of the sequential loop (triangle.rs:177-187)
          if (t > T.tmin && t < T.hit.t) {
            if (ANY) { T.hit.inst = 0; fin = true; }
            else { T.hit.t = t; T.hit.u = v; T.hit.v = w; T.hit.prim = tid; T.hit.tri = (int)(T.tg.x + k1); T.hit.inst = T.cur_inst; }
          }
          if (helped && t2 > T.tmin && t2 < T.hit.t) {
            if (ANY) { T.hit.inst = 0; fin = true; }
            else { T.hit.t = t2; T.hit.u = v2; T.hit.v = w2; T.hit.prim = tid2; T.hit.tri = (int)(T.tg.x + k2); T.hit.inst = T.cur_inst; }
          }
        }
      }
    }
#else
#pragma unroll
    for (int k = 0; k < SPT_TRI_PER_ROUND; k++) {
      const bool p = has && !fin && trav_has_leaf(T) && T.cur_inst >= 0;
      if (COUNT) { const uint32_t m = __ballot_sync(FULL, p); cnt.tri += p ? 1u : 0u; if (lane == 0 && m) cnt.w_tri++; }
      if (p) fin = trav_tri<ANY, COUNT>(sc, T, cnt);
    }
#endif
#if SPT_POP_END
    // ---- pop: nothing pending -> leave the BLAS / next stack entry / ray finished
    {
      const bool p = has && !fin && !trav_has_node(T) && !trav_has_leaf(T);
      if (COUNT) { const uint32_t m = __ballot_sync(FULL, p); cnt.pop += p ? 1u : 0u; if (lane == 0 && m) cnt.w_pop++; }
      if (p) fin = trav_pop(T, st);
    }
#endif
    if (fin) { src.store(idx, T.hit); has = false; }
  }
}

// sphere_normal_to_texcoords (sphere.rs:138-145)
SPT_D float2 sphere_uv(float3 n) {
  const float theta = acosf(n.y), phi = atan2f(n.x, n.z) + SPT_PI;
  return make_float2(phi * 0.5f * SPT_INV_PI, theta * SPT_INV_PI);
}
