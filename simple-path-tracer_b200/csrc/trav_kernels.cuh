// trav_kernels.cuh — the traversal kernels: extend (closest hit), connect (any hit + film), caller ray batches.
#pragma once
#include "wavefront.cuh"

#define SPT_BLOCK SPT_TRAV_BLOCK
#ifndef SPT_PF_L1
#define SPT_PF_L1 0
#endif
#if SPT_PF_L1
#define SPT_PF_INSTR "prefetch.global.L1 [%0];"
#else
#define SPT_PF_INSTR "prefetch.global.L2 [%0];"
#endif

SPT_D void flush_trav_counters(DevCounters* ctr, const TravCounters& tc) {
  warp_sum_to(&ctr->nodes, tc.nodes); warp_sum_to(&ctr->tris, tc.tris); warp_sum_to(&ctr->insts, tc.insts);
  const uint32_t v[10] = {tc.busy, tc.pop, tc.node, tc.inst, tc.tri, tc.w_rounds, tc.w_pop, tc.w_node, tc.w_inst, tc.w_tri};
#pragma unroll
  for (int i = 0; i < 10; i++) warp_sum_to(&ctr->trav_profile[i], v[i]);
}

// ---- ray sources of the persistent traversal driver -------------------------------------------------------------
struct ExtendSrc { // every live path of the pool: closest hit (pt.rs:49-50)
  const RenderParams& P;
  uint32_t n_rays;
  uint32_t n_list; // 0: scan the pool slots; > 0 (draining pool): the live slots are listed in P.ext_queue[P.parity][0 .. n_list)
  uint32_t chunk_;
  SPT_D uint32_t chunk() const { return chunk_; }
  SPT_D uint32_t total() const { return n_list ? n_list : (uint32_t)P.pool.n; }
  SPT_D uint32_t slot_of(uint32_t i) const { return n_list ? P.ext_queue[P.parity][i] : i; }
  SPT_D void prefetch(uint32_t i) const { if (!n_list) { asm volatile(SPT_PF_INSTR ::"l"(P.pool.rec + i)); if ((i & 7u) == 0u) asm volatile("prefetch.global.L2 [%0];" ::"l"(P.pool.flags + i)); } }
  SPT_D bool load(uint32_t i, float3& o, float3& d, float& tmin, float& tmax) {
    const uint32_t s = slot_of(i);
    if (!(P.pool.flags[s] & 1u)) return false;
    const float4 o4 = P.pool.rec[s].ray_o, d4 = P.pool.rec[s].ray_d;
    o = mk3(o4); d = mk3(d4); tmin = o4.w; tmax = FLT_MAX; // Intersection::default, intersection.rs:86-102
    n_rays++;
    return true;
  }
  SPT_D void store(uint32_t i, const HitRec& h) {
    const uint32_t s = slot_of(i);
    P.pool.hit[s] = make_float4(h.t, h.u, h.v, __int_as_float(h.tri));
    P.pool.hit_inst[s] = h.inst;
  }
};
struct ConnectSrc { // the queued shadow rays: any hit (pt.rs:81, :169); unoccluded contributions go to the film
  const RenderParams& P;
  uint32_t n, nonfinite;
  uint32_t chunk_;
  SPT_D uint32_t chunk() const { return chunk_; }
  SPT_D uint32_t total() const { return n; }
  SPT_D void prefetch(uint32_t i) const { if ((i & 1u) == 0u) { asm volatile("prefetch.global.L2 [%0];" ::"l"(P.sq.o + i)); asm volatile("prefetch.global.L2 [%0];" ::"l"(P.sq.d + i)); } }
  SPT_D bool load(uint32_t i, float3& o, float3& d, float& tmin, float& tmax) {
    const float4 o4 = P.sq.o[i], d4 = P.sq.d[i];
    o = mk3(o4); d = mk3(d4); tmin = o4.w; tmax = d4.w;
    return true;
  }
  SPT_D void store(uint32_t i, const HitRec& h) {
    if (h.inst < 0) { const float4 c = P.sq.c[i]; film_add(P, __float_as_uint(c.w), mk3(c), nonfinite); }
  }
};

// Rays a warp reserves per atomic.  With plenty of work the measured optimum (SPT_EXTEND_CHUNK / SPT_CONNECT_CHUNK = 64); when a launch
// holds fewer rays than that gives every warp of the grid (a draining pool, the last iterations of a render) the launch is
// latency-bound — its duration is the time of the warp with the most rays — so the rays are spread over all warps, down to
// SPT_MIN_CHUNK per warp.
#ifndef SPT_MIN_CHUNK
#define SPT_MIN_CHUNK 8
#endif
SPT_D uint32_t spread_chunk(uint32_t n_rays, uint32_t full) {
  const uint32_t warps = gridDim.x * (SPT_BLOCK / 32);
  return max((uint32_t)SPT_MIN_CHUNK, min(full, (n_rays + warps - 1) / warps));
}

// Tuning builds may force a register budget with -DSPT_TRAV_MINB=n (resident blocks per SM).  Measured on C2/C3/C5: the
// unconstrained allocation (72 registers, 7 blocks/SM) beats both n=8 (64 registers, -2 %) and n=1 (-18 %).
#ifdef SPT_TRAV_MINB
#define SPT_TRAV_BOUNDS __launch_bounds__(SPT_BLOCK, SPT_TRAV_MINB)
#else
#define SPT_TRAV_BOUNDS __launch_bounds__(SPT_BLOCK)
#endif

template <bool COUNT>
__global__ void SPT_TRAV_BOUNDS k_extend(const DScene sc, const RenderParams P) {
  TravCounters tc = {};
  ExtendSrc src{P, 0, P.ctr->it.drain ? P.ctr->it.n_ext[P.parity] : 0u, SPT_EXTEND_CHUNK};
  if (P.ctr->it.drain && src.n_list == 0u) return; // a draining pool without a live path
  if (src.n_list) src.chunk_ = spread_chunk(src.n_list, SPT_EXTEND_CHUNK);
  trace_persistent<false, COUNT>(sc, src, &P.ctr->it.next_ext, tc);
  warp_sum_to(&P.ctr->ext_rays, src.n_rays);
  if (COUNT) flush_trav_counters(P.ctr, tc);
}

template <bool COUNT>
__global__ void SPT_TRAV_BOUNDS k_connect(const DScene sc, const RenderParams P) {
  TravCounters tc = {};
  ConnectSrc src{P, P.ctr->it.sq, 0, SPT_CONNECT_CHUNK};
  src.chunk_ = spread_chunk(src.n, SPT_CONNECT_CHUNK);
  trace_persistent<true, COUNT>(sc, src, &P.ctr->it.next_con, tc);
  if (blockIdx.x == 0 && threadIdx.x == 0) atomicAdd(&P.ctr->shadow_rays, (unsigned long long)src.n);
  warp_sum_to(&P.ctr->nonfinite, src.nonfinite);
  if (COUNT) flush_trav_counters(P.ctr, tc);
}

// ---- seam B2 on caller-supplied ray batches -------------------------------------------------------------------

struct BatchSrc {
  const DScene& sc;
  const float4* __restrict__ rays; // 2 per ray: (o, tmin) (d, tmax)
  uint32_t n;
  HitIO* __restrict__ hits;       // closest-hit output or null
  uint8_t* __restrict__ occ;      // any-hit output or null
  SPT_D uint32_t chunk() const { return SPT_FETCH_CHUNK; }
  SPT_D void prefetch(uint32_t i) const { asm volatile("prefetch.global.L2 [%0];" ::"l"(rays + 2 * (size_t)i)); }
  SPT_D uint32_t total() const { return n; }
  SPT_D bool load(uint32_t i, float3& o, float3& d, float& tmin, float& tmax) {
    const float4 a = ldg4(rays + 2 * (size_t)i), b = ldg4(rays + 2 * (size_t)i + 1);
    o = mk3(a); d = mk3(b); tmin = a.w; tmax = b.w;
    return true;
  }
  SPT_D void store(uint32_t i, const HitRec& h) {
    if (occ) { occ[i] = h.inst >= 0 ? 1 : 0; return; }
    HitIO o; o.t = h.t; o.u = h.u; o.v = h.v; o.prim = h.prim; o.inst = h.inst;
    if (h.inst >= 0 && h.tri < 0) { // sphere: report texcoords like Intersection::texcoords
      const DInstance I = sc.inst[h.inst];
      const float4 a = ldg4(rays + 2 * (size_t)i), b = ldg4(rays + 2 * (size_t)i + 1);
      const float3 o2 = xmat_point(I.inv0, I.inv1, I.inv2, mk3(a)), d2 = xmat_vec(I.inv0, I.inv1, I.inv2, mk3(b));
      const float4 cr = sc.spheres[I.root];
      const float2 uv = sphere_uv((xpoint_at(o2, d2, h.t) - mk3(cr)) / cr.w);
      o.u = uv.x; o.v = uv.y;
    }
    hits[i] = o;
  }
};

template <bool ANY, bool COUNT>
__global__ void SPT_TRAV_BOUNDS k_batch(const DScene sc, const float4* __restrict__ rays, uint32_t n, HitIO* __restrict__ hits,
                                                     uint8_t* __restrict__ occ, uint32_t* next, DevCounters* ctr) {
  TravCounters tc = {};
  BatchSrc src{sc, rays, n, hits, occ};
  trace_persistent<ANY, COUNT>(sc, src, next, tc);
  if (COUNT && ctr) flush_trav_counters(ctr, tc);
}
