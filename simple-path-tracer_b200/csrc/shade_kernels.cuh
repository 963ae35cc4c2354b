// shade_kernels.cuh — shade<MAT> kernels and the per-path register state shared with k_special (see wavefront.cuh for the stage chain).
#pragma once
#include "wavefront.cuh"

#define SPT_BLOCK SPT_TRAV_BLOCK // the shade kernels share the shared-memory stack layout of the traversal (BSSRDF probe)

// ---- per-path state in registers during shading ---------------------------------------------------------------------------
struct PathRegs {
  float3 ro, rd, thr;
  float rtmin, last_pdf;
  PathRng rng;
  uint32_t pixel, sample;
  int depth, medium;
};
struct ShadowRay { float3 o, d, c; float tmin, tmax; bool emit; };

#ifndef SPT_SHADE_DEFER_SAMPLE
#define SPT_SHADE_DEFER_SAMPLE 1
#endif
#ifndef SPT_PROBE_STAGED
#define SPT_PROBE_STAGED 1 // 1: the BSSRDF probe is traced warp-synchronously in fixed-stage rounds (traverse_staged), 0: per-lane loop
#endif
#ifndef SPT_SHADE_PREFETCH
#define SPT_SHADE_PREFETCH 0 // 0 off (default: measured neutral on C2/C3/C4), 1 prefetch.global.L2, 2 prefetch.global.L1
#endif
SPT_D void prefetch_line(const void* p) {
#if SPT_SHADE_PREFETCH == 2
  asm volatile("prefetch.global.L1 [%0];" ::"l"(p));
#else
  asm volatile("prefetch.global.L2 [%0];" ::"l"(p));
#endif
}
SPT_D void prefetch_path(const RenderParams& P, int slot) {
  prefetch_line(P.pool.rec + slot);
  prefetch_line((const char*)(P.pool.rec + slot) + 32);
  prefetch_line(P.pool.hit + slot);
  prefetch_line(P.pool.hit_inst + slot);
  prefetch_line(P.pool.flags + slot);
}

SPT_D void load_path(const RenderParams& P, int slot, uint32_t flags, PathRegs& s, HitRec& hit) {
  const PathRec* rec = P.pool.rec + slot;
  const float4 o4 = rec->ray_o, d4 = rec->ray_d, t4 = rec->thr, h4 = P.pool.hit[slot];
  const uint4 aux = rec->aux;
  s.ro = mk3(o4); s.rd = mk3(d4); s.rtmin = o4.w;
  hit.t = h4.x; hit.u = h4.y; hit.v = h4.z; hit.tri = __float_as_int(h4.w); hit.prim = 0; hit.inst = P.pool.hit_inst[slot];
  s.thr = mk3(t4); s.last_pdf = t4.w;
  s.rng.state = (uint64_t)aux.x | ((uint64_t)aux.y << 32);
  s.pixel = aux.z; s.sample = aux.w;
  s.depth = (int)((flags >> 8) & 0xffu);
  s.medium = (int)(flags >> 16) - 1;
}
// Russian roulette + depth count (pt.rs:195-206), then write the path back or mark its slot free.
SPT_D void finish_bounce(const RenderParams& P, int slot, PathRegs& s, bool terminate) {
  if (!terminate) {
    if (!finite3(s.thr)) terminate = true;
    else {
      const float rr = s.rng.next();
      const float p = clampf_(lum3(s.thr), 0.001f, 0.95f);
      if (rr > p) terminate = true;
      else { s.thr = s.thr * (1.0f / p); s.depth++; if (s.depth >= P.max_depth) terminate = true; }
    }
  }
  if (terminate) { P.pool.flags[slot] = 0u; return; }
  PathRec* rec = P.pool.rec + slot; // the whole record: two full sectors, no partial-sector fills
  rec->ray_o = make_float4(s.ro.x, s.ro.y, s.ro.z, s.rtmin);
  rec->ray_d = make_float4(s.rd.x, s.rd.y, s.rd.z, 0.f);
  rec->thr = make_float4(s.thr.x, s.thr.y, s.thr.z, s.last_pdf);
  rec->aux = make_uint4((uint32_t)s.rng.state, (uint32_t)(s.rng.state >> 32), s.pixel, s.sample);
  P.pool.flags[slot] = 1u | ((uint32_t)s.depth << 8) | ((uint32_t)(s.medium + 1) << 16);
}
// Shadow-ray emission into the compacted queue, one atomic per warp.
// A sample whose contribution is exactly zero (light behind the surface, BxDF zero across the horizon, black emitter)
// is not queued: its any-hit result would be multiplied by zero.  NaN contributions are kept (they must be counted).
SPT_D void push_shadow(const RenderParams& P, const ShadowRay& sh, uint32_t pixel, uint32_t& n_culled) {
  const bool zero = sh.c.x == 0.0f && sh.c.y == 0.0f && sh.c.z == 0.0f;
  const bool emit = sh.emit && !zero;
  if (sh.emit && zero) n_culled++;
  const uint32_t qi = warp_reserve<uint32_t>(&P.ctr->it.sq, emit);
  if (emit) {
    P.sq.o[qi] = make_float4(sh.o.x, sh.o.y, sh.o.z, sh.tmin);
    P.sq.d[qi] = make_float4(sh.d.x, sh.d.y, sh.d.z, sh.tmax);
    P.sq.c[qi] = make_float4(sh.c.x, sh.c.y, sh.c.z, __uint_as_float(pixel));
  }
}

// ---- shade<MAT>: surface hits of one material class (pt.rs:112-193) --------------------------------------------------------------
// Resident blocks per SM the shade kernels are compiled for (register budget = 65536 / (128 * min_blocks)).
#if SPT_LEAN
// the lean kernels (texture.cuh: SPT_LEAN) need neither the TextureInput record nor the texture evaluator's stack frame: at 128
// registers they do not spill at all, which leaves room for more resident blocks (measured: profiles/README.md, round 2)
template <int MAT> struct ShadeCfg { static constexpr int min_blocks = SPT_SHADE_LEAN_MINB; };
#elif defined(SPT_SHADE_MINB)
template <int MAT> struct ShadeCfg { static constexpr int min_blocks = SPT_SHADE_MINB; };
#else
template <int MAT> struct ShadeCfg { static constexpr int min_blocks = 4; }; // measured best of 2..6 on every config
#endif

// LEAN only distinguishes the symbols of the two compilations of this kernel (shade_*.cu and shade_*_lean.cu); the body is selected
// by the SPT_LEAN macro of the translation unit.
template <int MAT, bool COUNT, int LEAN>
__global__ void __launch_bounds__(SPT_BLOCK, ShadeCfg<MAT>::min_blocks) k_shade(const DScene sc, const RenderParams P) {
  static_assert(LEAN == SPT_LEAN, "instantiate k_shade with LEAN = SPT_LEAN");
  constexpr bool PROBE = MAT < 0 || MAT == 7; // the BSSRDF probe ray needs the traversal stack
  __shared__ uint2 s_stack[PROBE ? SPT_SSTACK + SPT_WSTASH : 1][PROBE ? SPT_TRAV_BLOCK : 1];
  uint32_t nonfinite = 0, n_probe = 0, n_culled = 0;
  TravCounters tc = {};
  constexpr int QI = MAT < 0 ? SPT_Q_GENERIC : MAT;
  const uint32_t n = P.ctr->it.q[QI];
  const uint32_t* queue = P.queue + (size_t)QI * P.pool.n;
  const uint32_t stride = gridDim.x * blockDim.x;
  uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  int slot_next = i < n ? (int)queue[i] : 0;
  for (; i < n; i += stride) {
    const int slot = slot_next;
#if SPT_SHADE_PREFETCH
    // The per-path chain queue -> path record / hit -> instance -> triangle -> surface -> material is a string of dependent
    // loads (ncu: 40 % of the stall samples sit on its first three links).  Fetch the NEXT path's queue entry now and
    // prefetch its record, hit and flags, so that only the L2-resident scene links remain exposed in the next trip.
    if (i + stride < n) {
      slot_next = (int)queue[i + stride];
      prefetch_path(P, slot_next);
    }
#else
    if (i + stride < n) slot_next = (int)queue[i + stride];
#endif
    const uint32_t flags = P.pool.flags[slot];
    PathRegs s; HitRec hit;
    load_path(P, slot, flags, s, hit);
    ShadowRay sh; sh.emit = false; sh.o = sh.d = sh.c = mk3(0.f); sh.tmin = sh.tmax = 0.f;
    bool terminate = false;

    HitGeom g;
    load_hit(sc, s.ro, s.rd, hit, g);
    if (!SPT_LEAN && s.depth == 0 && sc.has_image_tex) { // ray differentials exist only for camera rays (Q18)
      const uint32_t k = s.sample;
      PathRng r2; r2.state = rng_init(P.seed, s.pixel, k);
      float ox, oy, x, y; pixel_offset(P, s.pixel, k, r2, ox, oy); film_xy(P, s.pixel, ox, oy, x, y);
      const float aspect = (float)P.width / (float)P.height, ssi = 1.0f / sqrtf((float)P.spp_frame);
      const float3 eye = mk3(sc.camera.eye[0], sc.camera.eye[1], sc.camera.eye[2]);
      calc_differential(g, eye, camera_dir(sc.camera, x + aspect / (float)P.width * ssi, y), eye, camera_dir(sc.camera, x, y + ssi / (float)P.height));
    }
    const DSurface sf = sc.surfaces[g.surface];
    const TexIn tin = texin_from(g);
    const Bsdf bsdf = make_bsdf<MAT>(sc, sc.materials[sf.material], tin);
    Frame fr = surface_frame(sc, sf, s.rd, g, tin);
    const float3 em = surface_emissive(sc, sf, tin);
    if (lum3(em) > 0.0f) {
      const float w = s.depth == 0 ? 1.0f : power_heuristic(s.last_pdf, pdf_shape_light(sc, s.ro, g, sf.double_sided != 0));
      film_add(P, s.pixel, s.thr * em * w, nonfinite);
    }
    float3 po = g.p;
    const float3 wo = to_local(fr, -s.rd);
    BsdfSample bs;
    // Lambert draws exactly two numbers for its sample: take a copy of the stream for them, step the path stream past
    // them, and compute the sample AFTER next-event estimation, so that its 10 registers are not live across the light
    // sampling code (same draws in the same order as sampling first; pt.rs:120-179).
    PathRng rng_bs; rng_bs.state = s.rng.state;
    if constexpr (MAT == 0 && SPT_SHADE_DEFER_SAMPLE) s.rng.skip2();
    else bsdf_sample<MAT>(bsdf, wo, s.rng, bs);
    bool dead_after_nee = false;
    if constexpr (PROBE) {
      if (bs.need_probe) { // Subsurface::sample, substrate.rs:231-333
        float rand_u = s.rng.next();
        const float rand_x = s.rng.next(), rand_y = s.rng.next();
        float3 st, sb, sn;
        if (rand_u < 0.5f) { rand_u = rand_u * 2.0f; st = fr.x; sb = fr.y; sn = fr.z; }
        else if (rand_u < 0.75f) { rand_u = rand_u * 4.0f - 2.0f; st = fr.y; sb = fr.z; sn = fr.x; }
        else { rand_u = rand_u * 4.0f - 3.0f; st = fr.z; sb = fr.x; sn = fr.y; }
        const float sp_d = rand_u < 1.0f / 3.0f ? bsdf.s_d.x : (rand_u < 2.0f / 3.0f ? bsdf.s_d.y : bsdf.s_d.z);
        const int ci = ss_cdf_lower_bound(sc.ss_cdf, rand_x);
        bool ok = ci < 512;
        float sample_r = -1.0f;
        if (ok) {
          const float y0 = sc.ss_cdf[2 * ci - 1], y1 = sc.ss_cdf[2 * ci + 1], x0 = sc.ss_cdf[2 * ci - 2], x1 = sc.ss_cdf[2 * ci];
          const float t = (rand_x - y0) / (y1 - y0);
          sample_r = (x1 * t + x0 * (1.0f - t)) * sp_d;
        }
        const float r_max = sc.ss_cdf[2 * 511] * sp_d;
        ok = ok && !(sample_r < 0.0f);
        if (ok) {
          float sphi, cphi; sincos_2pi(rand_y, sphi, cphi);
          const float sample_l = sqrtf(r_max * r_max + sample_r * sample_r);
          const float3 start = po + st * cphi * sample_r + sb * sphi * sample_r + sn * sample_l;
          HitRec ph; ph.t = 2.0f * sample_l; ph.u = ph.v = 0.f; ph.prim = -1; ph.inst = -1; ph.tri = -1;
          n_probe++;
          uint2 l_stack[SPT_LSTACK];
          TravStack tst; tst.sh = &s_stack[0][threadIdx.x]; tst.lo = l_stack;
#if SPT_PROBE_STAGED
          ok = traverse_staged<false, COUNT>(__activemask(), sc, start, -sn, 1e-4f, ph, tc, tst); // the lanes that probe in this trip, together
#else
          ok = traverse<false, COUNT>(sc, start, -sn, 1e-4f, ph, tc, tst);
#endif
          if (ok) {
            HitGeom pg; load_hit(sc, start, -sn, ph, pg);
            const DSurface psf = sc.surfaces[pg.surface];
            const Frame pfr = surface_frame(sc, psf, -sn, pg, texin_from(pg));
            const float3 pi = pg.p;
            const float3 spv = ss_sp(bsdf, len3(pi - po));
            const float3 off = to_local(fr, pi - po), nl = to_local(fr, pg.n);
            const float r_xy = sqrtf(off.x * off.x + off.y * off.y), r_yz = sqrtf(off.y * off.y + off.z * off.z), r_zx = sqrtf(off.z * off.z + off.x * off.x);
            const float pdf_pi = 0.5f * fabsf(nl.z) * avg3(ss_sp(bsdf, r_xy)) + 0.25f * fabsf(nl.x) * avg3(ss_sp(bsdf, r_yz)) + 0.25f * fabsf(nl.y) * avg3(ss_sp(bsdf, r_zx));
            float3 wi, fs; float ps;
            sub_sample_diffuse<MAT>(bsdf, wo, s.rng, wi, fs, ps);
            plastic_combine<MAT>(bsdf, wo, wi, fs, ps, LOBE_DIFFUSE, bs);
            po = pi; fr = pfr;                 // pt.rs:147-151
            s.thr *= spv / pdf_pi;
          }
        }
        if (!ok) dead_after_nee = true; // sample has wi = 0, bxdf = 0: the path contributes nothing further
      }
    }
    if (!bsdf_is_delta<MAT>(bsdf) && sc.n_lights > 0) { // pt.rs:158-179
      LightSample ls;
      sample_one_light(sc, po, s.rng, ls);
      const float3 wi = to_local(fr, ls.dir);
      const float3 f = bsdf_eval<MAT>(bsdf, wo, wi);
      const float mat_pdf = bsdf_pdf<MAT>(bsdf, wo, wi);
      if (ls.pdf != 0.0f && isfinite(ls.pdf)) {
        const float w = ls.delta ? 1.0f : power_heuristic(ls.pdf, mat_pdf);
        const float3 li = ls.L * f * fabsf(wi.z) * w / fmaxf(ls.pdf, 0.00001f);
        sh.emit = true; sh.o = po; sh.d = ls.dir; sh.tmin = 1e-4f / fmaxf(fabsf(wi.z), 0.00001f); sh.tmax = ls.dist - 0.001f; sh.c = s.thr * li;
      }
    }
    if constexpr (MAT == 0 && SPT_SHADE_DEFER_SAMPLE) {
      push_shadow(P, sh, s.pixel, n_culled);
      bsdf_sample<MAT>(bsdf, wo, rng_bs, bs);
    }
    if (dead_after_nee) terminate = true;
    else {
      s.last_pdf = bs.pdf;
      const float3 wi_world = to_world(fr, bs.wi);
      s.ro = po; s.rd = wi_world; s.rtmin = 1e-4f / fmaxf(fabsf(bs.wi.z), 0.00001f);
      s.thr *= bs.f * fabsf(bs.wi.z) / fmaxf(bs.pdf, 0.00001f);
      const float hd = dot3(wi_world, fr.hemi); // Coordinate::in_expected_hemisphere, coord.rs:53-59
      if (!(bs.dir == DIR_REFLECT ? hd >= 0.f : hd <= 0.f)) terminate = true;
      else if (dot3(wi_world, g.n) < 0.f) s.medium = sf.double_sided ? -1 : sf.inside_medium; // pt.rs:190-192, surface.rs:105-111
    }
    if constexpr (!(MAT == 0 && SPT_SHADE_DEFER_SAMPLE)) push_shadow(P, sh, s.pixel, n_culled);
    finish_bounce(P, slot, s, terminate);
  }
  warp_sum_to(&P.ctr->nonfinite, nonfinite);
  warp_sum_to(&P.ctr->shadow_culled, n_culled);
  if (PROBE) warp_sum_to(&P.ctr->probe_rays, n_probe);
  if (COUNT && PROBE) { warp_sum_to(&P.ctr->nodes, tc.nodes); warp_sum_to(&P.ctr->tris, tc.tris); warp_sum_to(&P.ctr->insts, tc.insts); }
}

