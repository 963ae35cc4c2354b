// kernels.cuh — __global__ entry points of the wavefront core.
#pragma once
#include "wavefront.cuh"

#define SPT_BLOCK 128
#ifndef SPT_SHADE_MINB
#define SPT_SHADE_MINB 1
#endif

// ---- extend: closest hit for every live path (pt.rs:49-50) ---------------------------------------------------
template <bool COUNT>
__global__ void __launch_bounds__(SPT_BLOCK) k_extend(const DScene sc, const RenderParams P) {
  TravCounters tc = {0, 0, 0};
  uint32_t n_rays = 0;
  for (int slot = blockIdx.x * blockDim.x + threadIdx.x; slot < P.pool.n; slot += gridDim.x * blockDim.x) {
    if (!(P.pool.flags[slot] & 1u)) continue;
    const float4 o = P.pool.ray_o[slot], d = P.pool.ray_d[slot];
    HitRec h; h.t = FLT_MAX; h.u = h.v = 0.f; h.prim = -1; h.inst = -1; // Intersection::default, intersection.rs:86-102
    traverse<false, COUNT>(sc, mk3(o), mk3(d), o.w, h, &tc);
    P.pool.hit[slot] = make_float4(h.t, h.u, h.v, __int_as_float(h.prim));
    P.pool.hit_inst[slot] = h.inst;
    n_rays++;
  }
  warp_sum_to(&P.ctr->ext_rays, n_rays);
  if (COUNT) { warp_sum_to(&P.ctr->nodes, tc.nodes); warp_sum_to(&P.ctr->tris, tc.tris); warp_sum_to(&P.ctr->insts, tc.insts); }
}

// ---- connect: any-hit for the queued shadow rays (pt.rs:81, :169) ---------------------------------------------
template <bool COUNT>
__global__ void __launch_bounds__(SPT_BLOCK) k_connect(const DScene sc, const RenderParams P) {
  TravCounters tc = {0, 0, 0};
  uint32_t nonfinite = 0;
  const uint32_t n = *P.sq.count;
  for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    const float4 o = P.sq.o[i], d = P.sq.d[i];
    HitRec h; h.t = d.w; h.u = h.v = 0.f; h.prim = -1; h.inst = -1;
    if (!traverse<true, COUNT>(sc, mk3(o), mk3(d), o.w, h, &tc)) {
      const float4 c = P.sq.c[i];
      film_add(P, __float_as_uint(c.w), mk3(c), nonfinite);
    }
  }
  if (blockIdx.x == 0 && threadIdx.x == 0) atomicAdd(&P.ctr->shadow_rays, (unsigned long long)n);
  warp_sum_to(&P.ctr->nonfinite, nonfinite);
  if (COUNT) { warp_sum_to(&P.ctr->nodes, tc.nodes); warp_sum_to(&P.ctr->tris, tc.tris); warp_sum_to(&P.ctr->insts, tc.insts); }
}

// ---- shade (+ regenerate): one iteration of the body of trace_ray's loop, pt.rs:48-207 --------------------------
template <bool COUNT>
__global__ void __launch_bounds__(SPT_BLOCK, SPT_SHADE_MINB) k_shade(const DScene sc, const RenderParams P) {
  uint32_t nonfinite = 0, n_probe = 0, n_new = 0, n_live = 0;
  TravCounters tc = {0, 0, 0};
  for (int slot = blockIdx.x * blockDim.x + threadIdx.x; slot < P.pool.n; slot += gridDim.x * blockDim.x) {
    uint32_t flags = P.pool.flags[slot];
    bool alive = flags & 1u;
    bool terminate = !alive;
    bool emit = false;
    float3 sh_o = mk3(0.f), sh_d = mk3(0.f), sh_c = mk3(0.f);
    float sh_tmin = 0.f, sh_tmax = 0.f;
    uint32_t pixel = 0;
    DrawSrc rng; rng.list = nullptr; rng.idx = rng.n = 0; rng.state = 0;
    float3 ro = mk3(0.f), rd = mk3(0.f), thr = mk3(1.f);
    float rtmin = 1e-4f, last_pdf = 0.f;
    int depth = 0, medium = -1;

    if (alive) {
      const float4 o4 = P.pool.ray_o[slot], d4 = P.pool.ray_d[slot], h4 = P.pool.hit[slot], t4 = P.pool.thr[slot];
      ro = mk3(o4); rd = mk3(d4); rtmin = o4.w;
      HitRec hit; hit.t = h4.x; hit.u = h4.y; hit.v = h4.z; hit.prim = __float_as_int(h4.w); hit.inst = P.pool.hit_inst[slot];
      thr = mk3(t4); last_pdf = t4.w;
      rng.state = P.pool.rng[slot];
      pixel = P.pool.pixel[slot];
      depth = (int)((flags >> 8) & 0xffu);
      medium = (int)(flags >> 16) - 1;
      const bool does_hit = hit.inst >= 0;
      bool medium_event = false;

      if (medium >= 0) { // pt.rs:56-96
        const DMedium md = sc.mediums[medium];
        const float3 sig_t = mk3(md.sigma_t[0], md.sigma_t[1], md.sigma_t[2]), sig_s = mk3(md.sigma_s[0], md.sigma_s[1], md.sigma_s[2]);
        const float3 wo = -rd;
        // Homogeneous::sample_pi, homogeneous.rs:31-60
        const float rx = rng.next(), ry = rng.next();
        const float sst = rx < 1.0f / 3.0f ? sig_t.x : (rx < 2.0f / 3.0f ? sig_t.y : sig_t.z);
        const float sample_t = -logf(1.0f - ry) / sst;
        const float tm = fminf(sample_t, hit.t);
        const float3 att = exp3(-sig_t * tm);
        const float3 pi = ro - wo * tm;
        if (sample_t < hit.t) {
          thr *= att * sig_s / avg3(sig_t * att);
          medium_event = true;
          if (!does_hit) terminate = true; // the reference panics here (pt.rs:77 unwrap on None)
          else {
            LightSample ls;
            if (sc.n_lights > 0) {
              sample_one_light(sc, pi, rng, ls);
              const float phase = hg_phase(md.g, dot3(wo, ls.dir));
              // shadow_ray_from_medium, pt.rs:212-233
              float tl, transported, tmin_s = 1e-4f;
              if (leaf_intersect(sc, hit.inst, hit.prim, pi, ls.dir, 1e-4f, ls.dist - 0.001f, tl)) { transported = tl; tmin_s += tl; }
              else { transported = ls.dist; tmin_s += ls.dist - 0.001f; }
              const float3 atten = exp3(-sig_t * transported);
              if (ls.pdf != 0.0f && isfinite(ls.pdf)) {
                float3 li = ls.delta ? atten * phase * ls.L / ls.pdf : atten * phase * ls.L * power_heuristic(ls.pdf, phase) / ls.pdf;
                emit = true; sh_o = pi; sh_d = ls.dir; sh_tmin = tmin_s; sh_tmax = ls.dist - 0.001f; sh_c = thr * li;
              }
            }
            // Homogeneous::sample_wi, homogeneous.rs:62-71
            const float u0 = rng.next(), u1 = rng.next();
            const float ct = hg_cdf_inverse(md.g, u0), st = sqrtf(1.0f - ct * ct), phi = 2.0f * SPT_PI * u1;
            float sp, cp; sincosf(phi, &sp, &cp);
            rd = medium_local_to_world(wo, mk3(st * cp, st * sp, ct));
            last_pdf = hg_phase(md.g, ct);
            ro = pi; rtmin = 1e-4f;
          }
        } else { // left the medium: pt.rs:62-64 `continue` re-finds the same hit; the cached hit is reused (Q21)
          thr *= att / avg3(att);
          medium = -1;
        }
      }

      if (!medium_event && !terminate) {
        if (!does_hit) { // pt.rs:97-111
          if (sc.env_tex) {
            float3 env; float env_pdf;
            env_eval_dir(sc, rd, env, env_pdf);
            const float w = depth == 0 ? 1.0f : power_heuristic(last_pdf, pdf_env_light(sc) * env_pdf);
            film_add(P, pixel, thr * env * w, nonfinite);
          }
          terminate = true;
        } else { // pt.rs:112-193
          HitGeom g;
          load_hit(sc, ro, rd, hit, g);
          if (depth == 0 && sc.has_image_tex) { // ray differentials exist only for camera rays (Q18)
            const uint32_t k = P.pool.sample[slot];
            DrawSrc r2; r2.list = nullptr; r2.idx = r2.n = 0; r2.state = rng_init(P.seed, pixel, k);
            float ox, oy, x, y; pixel_offset(P, pixel, k, r2, ox, oy); film_xy(P, pixel, ox, oy, x, y);
            const float aspect = (float)P.width / (float)P.height, ssi = 1.0f / sqrtf((float)P.spp);
            const float3 eye = mk3(sc.camera.eye[0], sc.camera.eye[1], sc.camera.eye[2]);
            calc_differential(g, eye, camera_dir(sc.camera, x + aspect / (float)P.width * ssi, y), eye, camera_dir(sc.camera, x, y + ssi / (float)P.height));
          }
          const DSurface sf = sc.surfaces[g.surface];
          const TexIn tin = texin_from(g);
          const Bsdf bsdf = make_bsdf(sc, sc.materials[sf.material], tin);
          Frame fr = surface_frame(sc, sf, rd, g, tin);
          const float3 em = surface_emissive(sc, sf, tin);
          if (lum3(em) > 0.0f) {
            const float w = depth == 0 ? 1.0f : power_heuristic(last_pdf, pdf_shape_light(sc, ro, g, sf.double_sided != 0));
            film_add(P, pixel, thr * em * w, nonfinite);
          }
          float3 po = g.p;
          const float3 wo = to_local(fr, -rd);
          BsdfSample bs;
          bsdf_sample(bsdf, wo, rng, bs);
          bool dead_after_nee = false;
          if (bs.need_probe) { // Subsurface::sample, substrate.rs:231-333
            float rand_u = rng.next();
            const float rand_x = rng.next(), rand_y = rng.next();
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
              const float phi = 2.0f * SPT_PI * rand_y;
              float sphi, cphi; sincosf(phi, &sphi, &cphi);
              const float sample_l = sqrtf(r_max * r_max + sample_r * sample_r);
              const float3 start = po + st * cphi * sample_r + sb * sphi * sample_r + sn * sample_l;
              HitRec ph; ph.t = 2.0f * sample_l; ph.u = ph.v = 0.f; ph.prim = -1; ph.inst = -1;
              n_probe++;
              ok = traverse<false, COUNT>(sc, start, -sn, 1e-4f, ph, &tc);
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
                sub_sample_diffuse(bsdf, wo, rng, wi, fs, ps);
                plastic_combine(bsdf, wo, wi, fs, ps, LOBE_DIFFUSE, bs);
                po = pi; fr = pfr;                 // pt.rs:147-151
                thr *= spv / pdf_pi;
              }
            }
            if (!ok) dead_after_nee = true; // sample has wi = 0, bxdf = 0: the path contributes nothing further
          }
          if (!bsdf_is_delta(bsdf) && sc.n_lights > 0) { // pt.rs:158-179
            LightSample ls;
            sample_one_light(sc, po, rng, ls);
            const float3 wi = to_local(fr, ls.dir);
            const float3 f = bsdf_eval(bsdf, wo, wi);
            const float mat_pdf = bsdf_pdf(bsdf, wo, wi);
            if (ls.pdf != 0.0f && isfinite(ls.pdf)) {
              const float w = ls.delta ? 1.0f : power_heuristic(ls.pdf, mat_pdf);
              const float3 li = ls.L * f * fabsf(wi.z) * w / fmaxf(ls.pdf, 0.00001f);
              emit = true; sh_o = po; sh_d = ls.dir; sh_tmin = 1e-4f / fmaxf(fabsf(wi.z), 0.00001f); sh_tmax = ls.dist - 0.001f; sh_c = thr * li;
            }
          }
          if (dead_after_nee) terminate = true;
          else {
            last_pdf = bs.pdf;
            const float3 wi_world = to_world(fr, bs.wi);
            ro = po; rd = wi_world; rtmin = 1e-4f / fmaxf(fabsf(bs.wi.z), 0.00001f);
            thr *= bs.f * fabsf(bs.wi.z) / fmaxf(bs.pdf, 0.00001f);
            const float hd = dot3(wi_world, fr.hemi); // Coordinate::in_expected_hemisphere, coord.rs:53-59
            if (!(bs.dir == DIR_REFLECT ? hd >= 0.f : hd <= 0.f)) terminate = true;
            else if (dot3(wi_world, g.n) < 0.f) medium = sf.double_sided ? -1 : sf.inside_medium; // pt.rs:190-192, surface.rs:105-111
          }
        }
      }
      if (!terminate) { // pt.rs:195-206
        if (!finite3(thr)) terminate = true;
        else {
          const float rr = rng.next();
          const float p = clampf_(lum3(thr), 0.001f, 0.95f);
          if (rr > p) terminate = true;
          else { thr = thr * (1.0f / p); depth++; if (depth >= P.max_depth) terminate = true; }
        }
      }
    }

    // ---- shadow-ray emission (compacted queue, one atomic per warp)
    {
      const uint32_t qi = warp_reserve<uint32_t>(P.sq.count, emit);
      if (emit) {
        P.sq.o[qi] = make_float4(sh_o.x, sh_o.y, sh_o.z, sh_tmin);
        P.sq.d[qi] = make_float4(sh_d.x, sh_d.y, sh_d.z, sh_tmax);
        P.sq.c[qi] = make_float4(sh_c.x, sh_c.y, sh_c.z, __uint_as_float(pixel));
      }
    }
    // ---- regenerate finished / empty slots from the sample cursor
    if (terminate) {
      bool got = false;
      uint32_t k = 0;
      unsigned long long s = warp_reserve<unsigned long long>(&P.ctr->cursor, true);
      for (;;) {
        if (s >= P.total_samples) break;
        if (ordinal_to_sample(P, s, pixel, k)) { got = true; break; }
        s = atomicAdd(&P.ctr->cursor, 1ull);
      }
      if (got) {
        rng.state = rng_init(P.seed, pixel, k);
        float ox, oy, x, y;
        pixel_offset(P, pixel, k, rng, ox, oy);
        film_xy(P, pixel, ox, oy, x, y);
        ro = mk3(sc.camera.eye[0], sc.camera.eye[1], sc.camera.eye[2]);
        rd = camera_dir(sc.camera, x, y);
        rtmin = 1e-4f; thr = mk3(1.f); last_pdf = 0.f; depth = 0; medium = -1;
        P.pool.pixel[slot] = pixel;
        P.pool.sample[slot] = k;
        n_new++;
        alive = true;
      } else alive = false;
    }
    if (alive) {
      P.pool.ray_o[slot] = make_float4(ro.x, ro.y, ro.z, rtmin);
      P.pool.ray_d[slot] = make_float4(rd.x, rd.y, rd.z, 0.f);
      P.pool.thr[slot] = make_float4(thr.x, thr.y, thr.z, last_pdf);
      P.pool.rng[slot] = rng.state;
      P.pool.flags[slot] = 1u | ((uint32_t)depth << 8) | ((uint32_t)(medium + 1) << 16);
      n_live++;
    } else if (flags & 1u) P.pool.flags[slot] = 0u;
  }
  warp_sum_to(&P.ctr->nonfinite, nonfinite);
  warp_sum_to(&P.ctr->probe_rays, n_probe);
  warp_sum_to(&P.ctr->paths, n_new);
  {
    const unsigned active = __activemask();
    const uint32_t s = __reduce_add_sync(active, n_live);
    if (s && lane_id() == (uint32_t)(__ffs(active) - 1)) atomicAdd(&P.ctr->live, s);
  }
  if (COUNT) { warp_sum_to(&P.ctr->nodes, tc.nodes); warp_sum_to(&P.ctr->tris, tc.tris); warp_sum_to(&P.ctr->insts, tc.insts); }
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
struct RayIO { float ox, oy, oz, tmin, dx, dy, dz, tmax; };
struct HitIO { float t, u, v; int prim, inst; };

template <bool COUNT>
__global__ void __launch_bounds__(SPT_BLOCK) k_trace_batch(const DScene sc, const float4* __restrict__ rays, long long n, HitIO* __restrict__ out, DevCounters* ctr) {
  TravCounters tc = {0, 0, 0};
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
    const float4 a = ldg4(rays + 2 * i), b = ldg4(rays + 2 * i + 1);
    HitRec h; h.t = b.w; h.u = h.v = 0.f; h.prim = -1; h.inst = -1;
    const bool found = traverse<false, COUNT>(sc, mk3(a), mk3(b), a.w, h, &tc);
    if (found && sc.inst[h.inst].kind == 0) { // sphere: report texcoords like Intersection::texcoords
      const DInstance I = sc.inst[h.inst];
      const float3 o2 = xmat_point(I.inv0, I.inv1, I.inv2, mk3(a)), d2 = xmat_vec(I.inv0, I.inv1, I.inv2, mk3(b));
      const float4 cr = sc.spheres[I.root];
      const float2 uv = sphere_uv((xpoint_at(o2, d2, h.t) - mk3(cr)) / cr.w);
      h.u = uv.x; h.v = uv.y;
    }
    HitIO o; o.t = h.t; o.u = h.u; o.v = h.v; o.prim = h.prim; o.inst = h.inst;
    out[i] = o;
  }
  if (COUNT && ctr) { warp_sum_to(&ctr->nodes, tc.nodes); warp_sum_to(&ctr->tris, tc.tris); warp_sum_to(&ctr->insts, tc.insts); }
}

template <bool COUNT>
__global__ void __launch_bounds__(SPT_BLOCK) k_occluded_batch(const DScene sc, const float4* __restrict__ rays, long long n, uint8_t* __restrict__ out, DevCounters* ctr) {
  TravCounters tc = {0, 0, 0};
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
    const float4 a = ldg4(rays + 2 * i), b = ldg4(rays + 2 * i + 1);
    HitRec h; h.t = b.w; h.u = h.v = 0.f; h.prim = -1; h.inst = -1;
    out[i] = traverse<true, COUNT>(sc, mk3(a), mk3(b), a.w, h, &tc) ? 1 : 0;
  }
  if (COUNT && ctr) { warp_sum_to(&ctr->nodes, tc.nodes); warp_sum_to(&ctr->tris, tc.tris); warp_sum_to(&ctr->insts, tc.insts); }
}
