// kernels.cuh — __global__ entry points of the wavefront core (see wavefront.cuh for the stage chain).
#pragma once
#include "wavefront.cuh"

#define SPT_BLOCK 128 // == SPT_TRAV_BLOCK: the traversal and shade kernels share the shared-memory stack layout
#define SPT_LOGIC_BLOCK 256

// ---- ray sources of the persistent traversal driver -------------------------------------------------------------
struct ExtendSrc { // every live path of the pool: closest hit (pt.rs:49-50)
  const RenderParams& P;
  uint32_t n_rays;
  SPT_D uint32_t total() const { return (uint32_t)P.pool.n; }
  SPT_D bool load(uint32_t i, float3& o, float3& d, float& tmin, float& tmax) {
    if (!(P.pool.flags[i] & 1u)) return false;
    const float4 o4 = P.pool.ray_o[i], d4 = P.pool.ray_d[i];
    o = mk3(o4); d = mk3(d4); tmin = o4.w; tmax = FLT_MAX; // Intersection::default, intersection.rs:86-102
    n_rays++;
    return true;
  }
  SPT_D void store(uint32_t i, const HitRec& h) {
    P.pool.hit[i] = make_float4(h.t, h.u, h.v, __int_as_float(h.tri));
    P.pool.hit_inst[i] = h.inst;
  }
};
struct ConnectSrc { // the queued shadow rays: any hit (pt.rs:81, :169); unoccluded contributions go to the film
  const RenderParams& P;
  uint32_t n, nonfinite;
  SPT_D uint32_t total() const { return n; }
  SPT_D bool load(uint32_t i, float3& o, float3& d, float& tmin, float& tmax) {
    const float4 o4 = P.sq.o[i], d4 = P.sq.d[i];
    o = mk3(o4); d = mk3(d4); tmin = o4.w; tmax = d4.w;
    return true;
  }
  SPT_D void store(uint32_t i, const HitRec& h) {
    if (h.inst < 0) { const float4 c = P.sq.c[i]; film_add(P, __float_as_uint(c.w), mk3(c), nonfinite); }
  }
};

template <bool COUNT>
__global__ void __launch_bounds__(SPT_BLOCK) k_extend(const DScene sc, const RenderParams P) {
  TravCounters tc = {0, 0, 0};
  if (blockIdx.x == 0 && threadIdx.x == 0) { P.ctr->it.sq = 0; P.ctr->it.next_con = 0; } // connect of this iteration is done
  ExtendSrc src{P, 0};
  trace_persistent<false, COUNT>(sc, src, &P.ctr->it.next_ext, tc);
  warp_sum_to(&P.ctr->ext_rays, src.n_rays);
  if (COUNT) { warp_sum_to(&P.ctr->nodes, tc.nodes); warp_sum_to(&P.ctr->tris, tc.tris); warp_sum_to(&P.ctr->insts, tc.insts); }
}

template <bool COUNT>
__global__ void __launch_bounds__(SPT_BLOCK) k_connect(const DScene sc, const RenderParams P) {
  TravCounters tc = {0, 0, 0};
  ConnectSrc src{P, P.ctr->it.sq, 0};
  trace_persistent<true, COUNT>(sc, src, &P.ctr->it.next_con, tc);
  if (blockIdx.x == 0 && threadIdx.x == 0) atomicAdd(&P.ctr->shadow_rays, (unsigned long long)src.n);
  warp_sum_to(&P.ctr->nonfinite, src.nonfinite);
  if (COUNT) { warp_sum_to(&P.ctr->nodes, tc.nodes); warp_sum_to(&P.ctr->tris, tc.tris); warp_sum_to(&P.ctr->insts, tc.insts); }
}

// ---- logic: bin the live paths by the kind of work their hit needs ---------------------------------------------------
// Dense pass in slot order, one slot per thread; queue space is reserved with ONE global atomic per class per block
// (shared-memory counters inside the block), so the queues stay nearly slot-ordered and the shade kernels' gathers
// stay mostly coalesced.
__global__ void __launch_bounds__(SPT_LOGIC_BLOCK) k_logic(const DScene sc, const RenderParams P) {
  __shared__ uint32_t s_cnt[SPT_NQ], s_base[SPT_NQ];
  if (threadIdx.x < SPT_NQ) s_cnt[threadIdx.x] = 0;
  if (blockIdx.x == 0 && threadIdx.x == 0) { P.ctr->it.live = 0; P.ctr->it.next_ext = 0; }
  __syncthreads();
  const int slot = blockIdx.x * blockDim.x + threadIdx.x;
  int cls = -1; uint32_t rank = 0;
  if (slot < P.pool.n) {
    const uint32_t flags = P.pool.flags[slot];
    if (flags & 1u) {
      const int inst = P.pool.hit_inst[slot];
      if ((flags >> 16) != 0u) cls = SPT_Q_SPECIAL;                       // inside a medium (pt.rs:56-96)
      else if (inst < 0) { if (sc.env_tex) cls = SPT_Q_SPECIAL; else P.pool.flags[slot] = 0u; } // miss (pt.rs:97-111)
      else cls = P.sort_materials ? (int)sc.inst_mat[inst] : SPT_Q_GENERIC; // surface hit (pt.rs:112-193)
    }
    if (cls >= 0) rank = atomicAdd(&s_cnt[cls], 1u);
  }
  __syncthreads();
  if (threadIdx.x < SPT_NQ && s_cnt[threadIdx.x]) s_base[threadIdx.x] = atomicAdd(&P.ctr->it.q[threadIdx.x], s_cnt[threadIdx.x]);
  __syncthreads();
  if (cls >= 0) P.queue[(size_t)cls * P.pool.n + s_base[cls] + rank] = (uint32_t)slot;
}

// ---- per-path state in registers during shading ---------------------------------------------------------------------------
struct PathRegs {
  float3 ro, rd, thr;
  float rtmin, last_pdf;
  PathRng rng;
  uint32_t pixel;
  int depth, medium;
};
struct ShadowRay { float3 o, d, c; float tmin, tmax; bool emit; };

SPT_D void load_path(const RenderParams& P, int slot, uint32_t flags, PathRegs& s, HitRec& hit) {
  const float4 o4 = P.pool.ray_o[slot], d4 = P.pool.ray_d[slot], h4 = P.pool.hit[slot], t4 = P.pool.thr[slot];
  s.ro = mk3(o4); s.rd = mk3(d4); s.rtmin = o4.w;
  hit.t = h4.x; hit.u = h4.y; hit.v = h4.z; hit.tri = __float_as_int(h4.w); hit.prim = 0; hit.inst = P.pool.hit_inst[slot];
  s.thr = mk3(t4); s.last_pdf = t4.w;
  s.rng.state = P.pool.rng[slot];
  s.pixel = P.pool.pixel[slot];
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
  P.pool.ray_o[slot] = make_float4(s.ro.x, s.ro.y, s.ro.z, s.rtmin);
  P.pool.ray_d[slot] = make_float4(s.rd.x, s.rd.y, s.rd.z, 0.f);
  P.pool.thr[slot] = make_float4(s.thr.x, s.thr.y, s.thr.z, s.last_pdf);
  P.pool.rng[slot] = s.rng.state;
  P.pool.flags[slot] = 1u | ((uint32_t)s.depth << 8) | ((uint32_t)(s.medium + 1) << 16);
}
// Shadow-ray emission into the compacted queue, one atomic per warp.
SPT_D void push_shadow(const RenderParams& P, const ShadowRay& sh, uint32_t pixel) {
  const uint32_t qi = warp_reserve<uint32_t>(&P.ctr->it.sq, sh.emit);
  if (sh.emit) {
    P.sq.o[qi] = make_float4(sh.o.x, sh.o.y, sh.o.z, sh.tmin);
    P.sq.d[qi] = make_float4(sh.d.x, sh.d.y, sh.d.z, sh.tmax);
    P.sq.c[qi] = make_float4(sh.c.x, sh.c.y, sh.c.z, __uint_as_float(pixel));
  }
}

// ---- special: paths inside a medium (pt.rs:56-96) and environment hits (pt.rs:97-111) -----------------------------------------
// A path that leaves its medium is handed on to the material queue of the surface it left through, to be shaded in
// this same iteration with the cached hit (Q21: the reference re-traces the same ray and finds the same hit).
__global__ void __launch_bounds__(SPT_BLOCK) k_special(const DScene sc, const RenderParams P) {
  uint32_t nonfinite = 0;
  const uint32_t n = P.ctr->it.q[SPT_Q_SPECIAL];
  const uint32_t* queue = P.queue + (size_t)SPT_Q_SPECIAL * P.pool.n;
  for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    const int slot = (int)queue[i];
    const uint32_t flags = P.pool.flags[slot];
    PathRegs s; HitRec hit;
    load_path(P, slot, flags, s, hit);
    const bool does_hit = hit.inst >= 0;
    ShadowRay sh; sh.emit = false; sh.o = sh.d = sh.c = mk3(0.f); sh.tmin = sh.tmax = 0.f;
    bool terminate = false, scattered = false;
    if (s.medium >= 0) {
      const DMedium md = sc.mediums[s.medium];
      const float3 sig_t = mk3(md.sigma_t[0], md.sigma_t[1], md.sigma_t[2]), sig_s = mk3(md.sigma_s[0], md.sigma_s[1], md.sigma_s[2]);
      const float3 wo = -s.rd;
      // Homogeneous::sample_pi, homogeneous.rs:31-60
      const float rx = s.rng.next(), ry = s.rng.next();
      const float sst = rx < 1.0f / 3.0f ? sig_t.x : (rx < 2.0f / 3.0f ? sig_t.y : sig_t.z);
      const float sample_t = -logf(1.0f - ry) / sst;
      const float tm = fminf(sample_t, hit.t);
      const float3 att = exp3(-sig_t * tm);
      const float3 pi = s.ro - wo * tm;
      if (sample_t < hit.t) {
        s.thr *= att * sig_s / avg3(sig_t * att);
        scattered = true;
        if (!does_hit) terminate = true; // the reference panics here (pt.rs:77 unwrap on None)
        else {
          if (sc.n_lights > 0) {
            LightSample ls;
            sample_one_light(sc, pi, s.rng, ls);
            const float phase = hg_phase(md.g, dot3(wo, ls.dir));
            // shadow_ray_from_medium, pt.rs:212-233
            float tl, transported, tmin_s = 1e-4f;
            if (leaf_intersect(sc, hit.inst, hit.tri, pi, ls.dir, 1e-4f, ls.dist - 0.001f, tl)) { transported = tl; tmin_s += tl; }
            else { transported = ls.dist; tmin_s += ls.dist - 0.001f; }
            const float3 atten = exp3(-sig_t * transported);
            if (ls.pdf != 0.0f && isfinite(ls.pdf)) {
              const float3 li = ls.delta ? atten * phase * ls.L / ls.pdf : atten * phase * ls.L * power_heuristic(ls.pdf, phase) / ls.pdf;
              sh.emit = true; sh.o = pi; sh.d = ls.dir; sh.tmin = tmin_s; sh.tmax = ls.dist - 0.001f; sh.c = s.thr * li;
            }
          }
          // Homogeneous::sample_wi, homogeneous.rs:62-71
          const float u0 = s.rng.next(), u1 = s.rng.next();
          const float ct = hg_cdf_inverse(md.g, u0), st = sqrtf(1.0f - ct * ct);
          float sp, cp; sincos_2pi(u1, sp, cp);
          s.rd = medium_local_to_world(wo, mk3(st * cp, st * sp, ct));
          s.last_pdf = hg_phase(md.g, ct);
          s.ro = pi; s.rtmin = 1e-4f;
        }
      } else { // left the medium
        s.thr *= att / avg3(att);
        s.medium = -1;
      }
    }
    bool to_surface = false;
    if (!scattered) {
      if (!does_hit) { // pt.rs:97-111
        if (sc.env_tex) {
          float3 env; float env_pdf;
          env_eval_dir(sc, s.rd, env, env_pdf);
          const float w = s.depth == 0 ? 1.0f : power_heuristic(s.last_pdf, pdf_env_light(sc) * env_pdf);
          film_add(P, s.pixel, s.thr * env * w, nonfinite);
        }
        terminate = true;
      } else to_surface = true;
    }
    push_shadow(P, sh, s.pixel);
    if (to_surface) { // write back what the medium changed, then queue for surface shading
      P.pool.thr[slot] = make_float4(s.thr.x, s.thr.y, s.thr.z, s.last_pdf);
      P.pool.rng[slot] = s.rng.state;
      P.pool.flags[slot] = 1u | ((uint32_t)s.depth << 8);
      const int cls = P.sort_materials ? (int)sc.inst_mat[hit.inst] : SPT_Q_GENERIC;
      const uint32_t qi = atomicAdd(&P.ctr->it.q[cls], 1u);
      P.queue[(size_t)cls * P.pool.n + qi] = (uint32_t)slot;
    } else finish_bounce(P, slot, s, terminate);
  }
  warp_sum_to(&P.ctr->nonfinite, nonfinite);
}

// ---- shade<MAT>: surface hits of one material class (pt.rs:112-193) --------------------------------------------------------------
template <int MAT> struct ShadeCfg { static constexpr int min_blocks = 3; };
template <> struct ShadeCfg<0> { static constexpr int min_blocks = 4; };
template <> struct ShadeCfg<1> { static constexpr int min_blocks = 4; };

template <int MAT, bool COUNT>
__global__ void __launch_bounds__(SPT_BLOCK, ShadeCfg<MAT>::min_blocks) k_shade(const DScene sc, const RenderParams P) {
  constexpr bool PROBE = MAT < 0 || MAT == 7; // the BSSRDF probe ray needs the traversal stack
  __shared__ uint2 s_stack[PROBE ? SPT_SSTACK : 1][PROBE ? SPT_TRAV_BLOCK : 1];
  uint32_t nonfinite = 0, n_probe = 0;
  TravCounters tc = {0, 0, 0};
  constexpr int QI = MAT < 0 ? SPT_Q_GENERIC : MAT;
  const uint32_t n = P.ctr->it.q[QI];
  const uint32_t* queue = P.queue + (size_t)QI * P.pool.n;
  for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    const int slot = (int)queue[i];
    const uint32_t flags = P.pool.flags[slot];
    PathRegs s; HitRec hit;
    load_path(P, slot, flags, s, hit);
    ShadowRay sh; sh.emit = false; sh.o = sh.d = sh.c = mk3(0.f); sh.tmin = sh.tmax = 0.f;
    bool terminate = false;

    HitGeom g;
    load_hit(sc, s.ro, s.rd, hit, g);
    if (s.depth == 0 && sc.has_image_tex) { // ray differentials exist only for camera rays (Q18)
      const uint32_t k = P.pool.sample[slot];
      PathRng r2; r2.state = rng_init(P.seed, s.pixel, k);
      float ox, oy, x, y; pixel_offset(P, s.pixel, k, r2, ox, oy); film_xy(P, s.pixel, ox, oy, x, y);
      const float aspect = (float)P.width / (float)P.height, ssi = 1.0f / sqrtf((float)P.spp);
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
    bsdf_sample<MAT>(bsdf, wo, s.rng, bs);
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
          ok = traverse<false, COUNT>(sc, start, -sn, 1e-4f, ph, tc, tst);
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
    push_shadow(P, sh, s.pixel);
    finish_bounce(P, slot, s, terminate);
  }
  warp_sum_to(&P.ctr->nonfinite, nonfinite);
  if (PROBE) warp_sum_to(&P.ctr->probe_rays, n_probe);
  if (COUNT && PROBE) { warp_sum_to(&P.ctr->nodes, tc.nodes); warp_sum_to(&P.ctr->tris, tc.tris); warp_sum_to(&P.ctr->insts, tc.insts); }
}

// ---- generate: refill finished / empty slots from the sample cursor (render loop pt.rs:263-280 + pixel sampler + camera) ----
__global__ void __launch_bounds__(SPT_LOGIC_BLOCK) k_generate(const DScene sc, const RenderParams P) {
  if (blockIdx.x == 0 && threadIdx.x < SPT_NQ) P.ctr->it.q[threadIdx.x] = 0; // the shade kernels of this iteration are done
  uint32_t n_new = 0, n_live = 0;
  for (int slot = blockIdx.x * blockDim.x + threadIdx.x; slot < P.pool.n; slot += gridDim.x * blockDim.x) {
    if (P.pool.flags[slot] & 1u) { n_live++; continue; }
    bool got = false;
    uint32_t k = 0, pixel = 0;
    unsigned long long s = warp_reserve<unsigned long long>(&P.ctr->cursor, true);
    for (;;) {
      if (s >= P.total_samples) break;
      if (ordinal_to_sample(P, s, pixel, k)) { got = true; break; }
      s = atomicAdd(&P.ctr->cursor, 1ull);
    }
    if (!got) continue;
    PathRng rng; rng.state = rng_init(P.seed, pixel, k);
    float ox, oy, x, y;
    pixel_offset(P, pixel, k, rng, ox, oy);
    film_xy(P, pixel, ox, oy, x, y);
    const float3 rd = camera_dir(sc.camera, x, y);
    P.pool.ray_o[slot] = make_float4(sc.camera.eye[0], sc.camera.eye[1], sc.camera.eye[2], 1e-4f);
    P.pool.ray_d[slot] = make_float4(rd.x, rd.y, rd.z, 0.f);
    P.pool.thr[slot] = make_float4(1.f, 1.f, 1.f, 0.f);
    P.pool.rng[slot] = rng.state;
    P.pool.pixel[slot] = pixel;
    P.pool.sample[slot] = k;
    P.pool.flags[slot] = 1u;
    n_new++; n_live++;
  }
  warp_sum_to(&P.ctr->paths, n_new);
  {
    const unsigned active = __activemask();
    const uint32_t s = __reduce_add_sync(active, n_live);
    if (s && lane_id() == (uint32_t)(__ffs(active) - 1)) atomicAdd(&P.ctr->it.live, s);
  }
}

// ---- resolve: Film::filter_pixel for box r<=0.5 (own-pixel mean, film.rs:68-91) + color_to_rgb (film.rs:94-99) -----
__global__ void k_resolve(const long long* __restrict__ film, int n_pixels, float inv_spp, float* __restrict__ out_mean, uint8_t* __restrict__ out_rgb8) {
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < 3 * n_pixels; i += gridDim.x * blockDim.x) {
    const float sum = (float)((double)film[i] * (1.0 / 16777216.0));
    const float m = sum * inv_spp;
    out_mean[i] = m;
    if (out_rgb8) out_rgb8[i] = (uint8_t)clampf_(m * 255.0f, 0.0f, 255.0f);
  }
}

// ---- seam B2 on caller-supplied ray batches -------------------------------------------------------------------
struct HitIO { float t, u, v; int prim, inst; };

struct BatchSrc {
  const DScene& sc;
  const float4* __restrict__ rays; // 2 per ray: (o, tmin) (d, tmax)
  uint32_t n;
  HitIO* __restrict__ hits;       // closest-hit output or null
  uint8_t* __restrict__ occ;      // any-hit output or null
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
__global__ void __launch_bounds__(SPT_BLOCK) k_batch(const DScene sc, const float4* __restrict__ rays, uint32_t n, HitIO* __restrict__ hits,
                                                     uint8_t* __restrict__ occ, uint32_t* next, DevCounters* ctr) {
  TravCounters tc = {0, 0, 0};
  BatchSrc src{sc, rays, n, hits, occ};
  trace_persistent<ANY, COUNT>(sc, src, next, tc);
  if (COUNT && ctr) { warp_sum_to(&ctr->nodes, tc.nodes); warp_sum_to(&ctr->tris, tc.tris); warp_sum_to(&ctr->insts, tc.insts); }
}
