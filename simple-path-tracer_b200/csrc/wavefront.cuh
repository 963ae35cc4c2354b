// wavefront.cuh — shared definitions of the wavefront path-tracing kernels (kernels.cuh).
//
// Replaces PathTracer::render + trace_ray (src/renderer/pt.rs:39-295): instead of one recursive loop per pixel
// sample, a pool of resident paths (SoA in HBM/L2) is advanced one bounce per iteration by a chain of kernels:
//     logic    : classify every live path by what its hit needs next -> per-class queues of slot ids
//                (8 material classes binned by material tag, + "special" = inside a medium / environment miss)
//     special  : medium events (Homogeneous::sample_pi/sample_wi, NEE from inside the medium) and environment hits
//     shade<M> : one kernel per material class present in the scene: rebuild the hit, emission + MIS, BxDF sample,
//                NEE shadow ray into a compacted queue, Russian roulette.  Binning removes the 8-way BxDF divergence
//                and lets each class compile to its own register budget.
//     generate : finished slots are REGENERATED in place from a global camera-sample cursor (pixel sampler + camera),
//                so the pool stays full until the image is done
//     connect  : any-hit test of the queued shadow rays, unoccluded contributions go to the film
//     extend   : closest-hit traversal of every live path's ray
// Film: 64-bit fixed-point sums (2^-24) updated with integer atomics, so the image is independent of the order
// in which paths finish — bit-identical across pool sizes, tile ownership and GPU counts.
#pragma once
#include "shade.cuh"

#include "wavefront_types.h"

#define FILM_SCALE 16777216.0f
#define FILM_CLAMP 67108864.0f

SPT_D void film_add(const RenderParams& P, uint32_t pix, float3 c, uint32_t& nonfinite) {
  if (!finite3(c)) { nonfinite++; return; }
  unsigned long long* f = reinterpret_cast<unsigned long long*>(P.film) + 3 * (size_t)pix;
  const long long r = __float2ll_rn(clampf_(c.x, -FILM_CLAMP, FILM_CLAMP) * FILM_SCALE);
  const long long g = __float2ll_rn(clampf_(c.y, -FILM_CLAMP, FILM_CLAMP) * FILM_SCALE);
  const long long b = __float2ll_rn(clampf_(c.z, -FILM_CLAMP, FILM_CLAMP) * FILM_SCALE);
  if (r) atomicAdd(f, (unsigned long long)r);
  if (g) atomicAdd(f + 1, (unsigned long long)g);
  if (b) atomicAdd(f + 2, (unsigned long long)b);
}

SPT_D uint32_t lane_id() { return threadIdx.x & 31u; }
// Warp-aggregated fetch-and-add: one atomic per warp for all lanes with `want` set.
template <typename T>
SPT_D T warp_reserve(T* counter, bool want) {
  const unsigned active = __activemask();
  const unsigned m = __ballot_sync(active, want);
  T base = 0;
  if (want) {
    const int leader = __ffs(m) - 1;
    if ((int)lane_id() == leader) base = atomicAdd(counter, (T)__popc(m));
    base = __shfl_sync(m, base, leader);
    base += (T)__popc(m & ((1u << lane_id()) - 1u));
  }
  return base;
}
SPT_D void warp_sum_to(unsigned long long* dst, uint32_t v) {
  const unsigned active = __activemask();
  const uint32_t s = __reduce_add_sync(active, v);
  if (s && lane_id() == (uint32_t)(__ffs(active) - 1)) atomicAdd(dst, (unsigned long long)s);
}

SPT_D void warp_sum_to32(uint32_t* dst, uint32_t v) {
  const unsigned active = __activemask();
  const uint32_t s = __reduce_add_sync(active, v);
  if (s && lane_id() == (uint32_t)(__ffs(active) - 1)) atomicAdd(dst, s);
}

// Camera sample ordinal -> (pixel, k).  Ordinals run tile-major, then sample-major, then 8x4 pixel blocks inside
// the tile so that a warp holds an 8x4 block of primary rays.  Returns false for the padding of edge tiles.
SPT_D bool ordinal_to_sample(const RenderParams& P, unsigned long long s, uint32_t& pixel, uint32_t& k) {
  const uint32_t tp = (uint32_t)(P.tile_w * P.tile_h);
  const unsigned long long per_tile = (unsigned long long)tp * (unsigned long long)P.spp;
  const uint32_t tl = (uint32_t)(s / per_tile);
  const uint32_t rem = (uint32_t)(s % per_tile);
  k = (uint32_t)P.spp_first + rem / tp;
  const uint32_t q = rem % tp, blk = q >> 5, l = q & 31u, bpr = (uint32_t)P.tile_w >> 3;
  const uint32_t tx = (blk % bpr) * 8u + (l & 7u), ty = (blk / bpr) * 4u + (l >> 3);
  const uint32_t tile = P.tiles[tl];
  const uint32_t x = (tile % (uint32_t)P.tiles_x) * (uint32_t)P.tile_w + tx, y = (tile / (uint32_t)P.tiles_x) * (uint32_t)P.tile_h + ty;
  pixel = y * (uint32_t)P.width + x;
  return x < (uint32_t)P.width && y < (uint32_t)P.height;
}

// PixelSamplerT::next_sample in closed form (recurrence.rs:41-56, random.rs:31-38, jittered.rs:44-57 with Q16 fixed)
template <class RNG>
SPT_D void pixel_offset(const RenderParams& P, uint32_t pixel, uint32_t k, RNG& rng, float& ox, float& oy) {
  if (P.sampler == 2) r2_offset((uint32_t)((unsigned long long)pixel * (unsigned long long)P.spp_frame + k) + 1u, ox, oy);
  else if (P.sampler == 0) { ox = rng.next(); oy = rng.next(); }
  else {
    const uint32_t ix = k % (uint32_t)P.div_x, iy = k / (uint32_t)P.div_x;
    ox = ((float)ix + rng.next()) * (1.0f / (float)P.div_x);
    oy = ((float)iy + rng.next()) * (1.0f / (float)P.div_y);
  }
}
// pt.rs:269-275 + PerspectiveCamera::generate_ray (camera/perspective.rs:41-47)
SPT_D float3 camera_dir(const DCamera& c, float x, float y) {
  const float3 f = mk3(c.forward[0], c.forward[1], c.forward[2]), r = mk3(c.right[0], c.right[1], c.right[2]), u = mk3(c.up[0], c.up[1], c.up[2]);
  return norm3(f * c.half_cot_half_fov + r * x + u * y);
}
SPT_D void film_xy(const RenderParams& P, uint32_t pixel, float ox, float oy, float& x, float& y) {
  const uint32_t i = pixel % (uint32_t)P.width, j = pixel / (uint32_t)P.width;
  const float aspect = (float)P.width / (float)P.height;
  x = (((float)i + ox) * (1.0f / (float)P.width) - 0.5f) * aspect;
  y = ((float)((uint32_t)P.height - j - 1u) + oy) * (1.0f / (float)P.height) - 0.5f;
}

// Intersect ONE leaf primitive with a ray — BasicPrimitiveRef::intersect as used by shadow_ray_from_medium
// (pt.rs:212-233).  The reference passes the WORLD-space ray to the object-space primitive; reproduced.
SPT_D bool leaf_intersect(const DScene& sc, int inst, int tri, float3 o, float3 d, float tmin, float tmax, float& t_out) {
  const DInstance I = sc.inst[inst];
  if (tri < 0) {
    float mn, mx;
    if (!sphere_roots(sc.spheres[I.root], o, d, mn, mx)) return false;
    const float t = mn < tmin ? mx : mn;
    if (tmin < t && t < tmax) { t_out = t; return true; }
    return false;
  }
  const float4* tp = sc.tris + 3 * (size_t)tri;
  const float3 p0 = mk3(ldg4(tp)), e1 = mk3(ldg4(tp + 1)), e2 = mk3(ldg4(tp + 2)), q = xcross3(d, e2);
  float det = xdot3(e1, q);
  if (det == 0.0f) return false;
  det = __fdiv_rn(1.0f, det);
  const float3 s = xsub3(o, p0);
  const float v = xmul(xdot3(s, q), det);
  if (!(v >= 0.f)) return false;
  const float3 r = xcross3(s, e1);
  const float w = xmul(xdot3(d, r), det), u = xsub(xsub(1.0f, v), w);
  if (!(w >= 0.f && u >= 0.f)) return false;
  const float t = xmul(xdot3(e2, r), det);
  if (t > tmin && t < tmax) { t_out = t; return true; }
  return false;
}

SPT_D int ss_cdf_lower_bound(const float* __restrict__ tab, float rand) { // first i >= 1 with y[i] >= rand, or 512
  int lo = 1, hi = 512;
  while (lo < hi) { const int mid = (lo + hi) >> 1; if (tab[2 * mid + 1] >= rand) hi = mid; else lo = mid + 1; }
  return lo;
}
