// misc_kernels.cu — logic (binning), special (media / environment), generate (camera + pixel sampler), resolve (film) and
// the BxDF test-hook kernel.
#include "launch.h"
#include "shade_kernels.cuh"

// ---- logic: bin the live paths by the kind of work their hit needs ---------------------------------------------------
// Dense pass in slot order, one slot per thread; queue space is reserved with ONE global atomic per class per block
// (shared-memory counters inside the block), so the queues stay nearly slot-ordered and the shade kernels' gathers
// stay mostly coalesced.
__global__ void __launch_bounds__(SPT_LOGIC_BLOCK) k_logic(const DScene sc, const RenderParams P) {
  __shared__ uint32_t s_cnt[SPT_NQ], s_base[SPT_NQ];
  if (threadIdx.x < SPT_NQ) s_cnt[threadIdx.x] = 0;
  const int par = P.parity;
  // A draining pool: the previous iteration's generate left the list of the slots that are still alive (ext_queue[par ^ 1]) — this pass
  // walks that list instead of the whole pool (a 16 M-slot pool holding 3 K paths cost 77 us here, in every one of the last iterations)
  const bool from_list = P.ctr->it.list_valid[par ^ 1] != 0u;
  const uint32_t n_prev = P.ctr->it.n_ext[par ^ 1];
  if (blockIdx.x == 0 && threadIdx.x == 0) { // connect + extend of the last iteration are done
    P.ctr->it.live = 0; P.ctr->it.next_ext = 0; P.ctr->it.sq = 0; P.ctr->it.next_con = 0; P.ctr->it.n_ext[par] = 0;
    P.ctr->it.drain = (P.use_ext_queue && P.ctr->cursor >= P.total_samples) ? 1u : 0u; // the cursor only moves in generate: one value for the whole iteration
  }
  if (from_list && blockIdx.x > 0 && (uint32_t)(blockIdx.x * blockDim.x) >= n_prev) return; // whole block beyond the list
  __syncthreads();
  const uint32_t gid = blockIdx.x * blockDim.x + threadIdx.x;
  int slot = -1;
  if (from_list) { if (gid < n_prev) slot = (int)P.ext_queue[par ^ 1][gid]; }
  else if (gid < (uint32_t)P.pool.n) slot = (int)gid;
  int cls = -1; uint32_t rank = 0;
  if (slot >= 0) {
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

// ---- special: paths inside a medium (pt.rs:56-96) and environment hits (pt.rs:97-111) -----------------------------------------
// A path that leaves its medium is handed on to the material queue of the surface it left through, to be shaded in
// this same iteration with the cached hit (Q21: the reference re-traces the same ray and finds the same hit).
__global__ void __launch_bounds__(SPT_BLOCK) k_special(const DScene sc, const RenderParams P) {
  uint32_t nonfinite = 0, n_culled = 0;
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
    push_shadow(P, sh, s.pixel, n_culled);
    if (to_surface) { // write back what the medium changed, then queue for surface shading
      P.pool.rec[slot].thr = make_float4(s.thr.x, s.thr.y, s.thr.z, s.last_pdf);
      P.pool.rec[slot].aux = make_uint4((uint32_t)s.rng.state, (uint32_t)(s.rng.state >> 32), s.pixel, s.sample);
      P.pool.flags[slot] = 1u | ((uint32_t)s.depth << 8);
      const int cls = P.sort_materials ? (int)sc.inst_mat[hit.inst] : SPT_Q_GENERIC;
      const uint32_t qi = atomicAdd(&P.ctr->it.q[cls], 1u);
      P.queue[(size_t)cls * P.pool.n + qi] = (uint32_t)slot;
    } else finish_bounce(P, slot, s, terminate);
  }
  warp_sum_to(&P.ctr->nonfinite, nonfinite);
  warp_sum_to(&P.ctr->shadow_culled, n_culled);
}

// ---- generate: refill finished / empty slots from the sample cursor (render loop pt.rs:263-280 + pixel sampler + camera) ----
// Camera-sample ordinals are reserved with ONE global atomic per block and trip (block-wide count of free slots via
// ballot + shared memory), and handed out in slot order, so a warp's new paths are consecutive ordinals = an 8x4 pixel
// block of primary rays.  An ordinal that falls into the padding of an edge tile leaves its slot free until the next
// iteration.
__global__ void __launch_bounds__(SPT_LOGIC_BLOCK) k_generate(const DScene sc, const RenderParams P) {
  __shared__ uint32_t s_warp[SPT_LOGIC_BLOCK / 32];
  __shared__ unsigned long long s_base;
  if (blockIdx.x == 0 && threadIdx.x < SPT_NQ) P.ctr->it.q[threadIdx.x] = 0; // the shade kernels of this iteration are done
  uint32_t n_new = 0, n_live = 0;
  const uint32_t lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
  if (P.ctr->it.drain) {
    // Every camera sample has been started: nothing to regenerate.  The pool now empties within max_depth iterations, and a kernel
    // that scans it finds fewer and fewer rays per chunk (at 10 % live a lane of k_extend spends 9 rounds finding a ray that lives 4).
    // So this pass compacts the live slots into ext_queue (one atomic per block and trip) and k_extend fetches from that list.
    // From the second drain iteration on, the candidates are the previous iteration's list, not the pool.
    const int par = P.parity;
    const bool from_list = P.ctr->it.list_valid[par ^ 1] != 0u;
    const uint32_t n_src = from_list ? P.ctr->it.n_ext[par ^ 1] : (uint32_t)P.pool.n;
    const uint32_t* src = P.ext_queue[par ^ 1];
    uint32_t* dst = P.ext_queue[par];
    if (blockIdx.x == 0 && threadIdx.x == 0) P.ctr->it.list_valid[par] = 1u; // read by the NEXT iteration's kernels only
    for (uint32_t base = blockIdx.x * blockDim.x; base < n_src; base += gridDim.x * blockDim.x) {
      const uint32_t idx = base + threadIdx.x;
      const uint32_t slot = idx < n_src ? (from_list ? src[idx] : idx) : 0u;
      const bool live = idx < n_src && (P.pool.flags[slot] & 1u);
      const uint32_t m = __ballot_sync(0xffffffffu, live);
      if (lane == 0) s_warp[warp] = __popc(m);
      __syncthreads();
      if (threadIdx.x == 0) {
        uint32_t total = 0;
        for (int w = 0; w < SPT_LOGIC_BLOCK / 32; w++) { const uint32_t c = s_warp[w]; s_warp[w] = total; total += c; }
        s_base = total ? (unsigned long long)atomicAdd(&P.ctr->it.n_ext[par], total) : 0ull;
      }
      __syncthreads();
      if (live) { dst[(uint32_t)s_base + s_warp[warp] + __popc(m & ((1u << lane) - 1u))] = slot; n_live++; }
      __syncthreads();
    }
    warp_sum_to32(&P.ctr->it.live, n_live);
    return;
  }
  if (blockIdx.x == 0 && threadIdx.x == 0) P.ctr->it.list_valid[P.parity] = 0u; // no list from this iteration
  for (int base_slot = blockIdx.x * blockDim.x; base_slot < P.pool.n; base_slot += gridDim.x * blockDim.x) {
    const int slot = base_slot + (int)threadIdx.x;
    const bool in_range = slot < P.pool.n;
    const bool dead = in_range && !(P.pool.flags[slot] & 1u);
    if (in_range && !dead) n_live++;
    const uint32_t m = __ballot_sync(0xffffffffu, dead);
    if (lane == 0) s_warp[warp] = __popc(m);
    __syncthreads();
    if (threadIdx.x == 0) {
      uint32_t total = 0;
      for (int w = 0; w < SPT_LOGIC_BLOCK / 32; w++) { const uint32_t c = s_warp[w]; s_warp[w] = total; total += c; }
      s_base = total ? atomicAdd(&P.ctr->cursor, (unsigned long long)total) : 0ull;
    }
    __syncthreads();
    const unsigned long long ord = s_base + s_warp[warp] + __popc(m & ((1u << lane) - 1u));
    __syncthreads(); // s_warp / s_base are rewritten by the next trip
    uint32_t k = 0, pixel = 0;
    if (!dead || ord >= P.total_samples || !ordinal_to_sample(P, ord, pixel, k)) continue;
    PathRng rng; rng.state = rng_init(P.seed, pixel, k);
    float ox, oy, x, y;
    pixel_offset(P, pixel, k, rng, ox, oy);
    film_xy(P, pixel, ox, oy, x, y);
    const float3 rd = camera_dir(sc.camera, x, y);
    PathRec* rec = P.pool.rec + slot;
    rec->ray_o = make_float4(sc.camera.eye[0], sc.camera.eye[1], sc.camera.eye[2], 1e-4f);
    rec->ray_d = make_float4(rd.x, rd.y, rd.z, 0.f);
    rec->thr = make_float4(1.f, 1.f, 1.f, 0.f);
    rec->aux = make_uint4((uint32_t)rng.state, (uint32_t)(rng.state >> 32), pixel, k);
    P.pool.flags[slot] = 1u;
    n_new++; n_live++;
  }
  warp_sum_to(&P.ctr->paths, n_new);
  warp_sum_to32(&P.ctr->it.live, n_live);
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


// ---- resolve for any box radius: Film::filter_pixel (film.rs:68-91) with BoxFilter (boxf.rs:10-34), quirk Q14 ------------------------
// The reference keeps every sample (offset - 0.5, colour) and reconstructs pixel (x, y) as
//     sum of the UNWEIGHTED colours of all samples of the (2R+1)^2 neighbour pixels, R = ceil(radius - 0.5)
//   / number of those samples whose position relative to the pixel centre, (i + ox - 0.5, j + oy - 0.5), lies in the box.
// The numerator only needs the per-pixel sums the film already holds.  The denominator needs the sub-pixel offset of every sample,
// which is a closed-form function of (pixel, k) for all three samplers (pixel_offset(): R2 ordinal, or the first draws of the
// path's keyed stream) — so it is recomputed here instead of being stored: no per-sample memory, (2R+1)^2 * spp offset
// evaluations per pixel (C2 at radius 1.5: 4.8 G evaluations, a few ms).
__global__ void k_resolve_box(const RenderParams P, int k0, int n_k, float radius, int radius_int, float* __restrict__ out_mean,
                              uint8_t* __restrict__ out_rgb8) {
  const int n_px = P.width * P.height;
  for (int p = blockIdx.x * blockDim.x + threadIdx.x; p < n_px; p += gridDim.x * blockDim.x) {
    const int x = p % P.width, y = p / P.width;
    long long sum[3] = {0, 0, 0};
    uint32_t in_box = 0;
    for (int j = -radius_int; j <= radius_int; j++) {
      if (y + j < 0 || y + j >= P.height) continue;
      for (int i = -radius_int; i <= radius_int; i++) {
        if (x + i < 0 || x + i >= P.width) continue;
        const uint32_t q = (uint32_t)((y + j) * P.width + (x + i));
        sum[0] += P.film[3 * (size_t)q]; sum[1] += P.film[3 * (size_t)q + 1]; sum[2] += P.film[3 * (size_t)q + 2];
        for (int k = k0; k < k0 + n_k; k++) {
          PathRng rng; rng.state = rng_init(P.seed, q, (uint32_t)k);
          float ox, oy;
          pixel_offset(P, q, (uint32_t)k, rng, ox, oy);
          const float fx = (float)i + (ox - 0.5f), fy = (float)j + (oy - 0.5f); // pt.rs:277-279: the film stores offset - 0.5
          if (fabsf(fx) <= radius && fabsf(fy) <= radius) in_box++;            // BoxFilter::weight, boxf.rs:27-33
        }
      }
    }
    const float inv = 1.0f / (float)in_box; // Color / f32 multiplies by the reciprocal (color.rs:125-139); 0 samples in the box -> inf / NaN as in the reference
#pragma unroll
    for (int c = 0; c < 3; c++) {
      const float m = (float)((double)sum[c] * (1.0 / 16777216.0)) * inv;
      out_mean[3 * (size_t)p + c] = m;
      if (out_rgb8) out_rgb8[3 * (size_t)p + c] = (uint8_t)clampf_(m * 255.0f, 0.0f, 255.0f);
    }
  }
}

// ---- tile-packed film exchange (spt_render_gather): 3 x int64 per pixel of the tiles a rank owns, tile-major ---------------------------------
__global__ void k_pack_tiles(const long long* __restrict__ film, const uint32_t* __restrict__ tiles, int n_tiles, const RenderParams P, long long* __restrict__ out) {
  const int tp = P.tile_w * P.tile_h;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < (long long)n_tiles * tp; i += (long long)gridDim.x * blockDim.x) {
    const int t = (int)(i / tp), q = (int)(i % tp);
    const uint32_t tile = tiles[t];
    const int x = (int)(tile % (uint32_t)P.tiles_x) * P.tile_w + q % P.tile_w, y = (int)(tile / (uint32_t)P.tiles_x) * P.tile_h + q / P.tile_w;
    long long v[3] = {0, 0, 0};
    if (x < P.width && y < P.height) { const size_t px = (size_t)y * P.width + x; v[0] = film[3 * px]; v[1] = film[3 * px + 1]; v[2] = film[3 * px + 2]; }
    out[3 * i] = v[0]; out[3 * i + 1] = v[1]; out[3 * i + 2] = v[2];
  }
}
// root: segment r of `in` holds rank r's tiles in the order of all_tiles[r * per_rank ..] (0xffffffff = padding); the root's own segment is skipped
__global__ void k_unpack_tiles(long long* __restrict__ film, const uint32_t* __restrict__ all_tiles, int per_rank, int world, int root, const RenderParams P,
                               const long long* __restrict__ in) {
  const int tp = P.tile_w * P.tile_h;
  const long long n = (long long)world * per_rank * tp;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
    const long long t = i / tp; const int q = (int)(i % tp);
    if ((int)(t / per_rank) == root) continue;
    const uint32_t tile = all_tiles[t];
    if (tile == 0xffffffffu) continue;
    const int x = (int)(tile % (uint32_t)P.tiles_x) * P.tile_w + q % P.tile_w, y = (int)(tile / (uint32_t)P.tiles_x) * P.tile_h + q / P.tile_w;
    if (x < P.width && y < P.height) { const size_t px = (size_t)y * P.width + x; film[3 * px] = in[3 * i]; film[3 * px + 1] = in[3 * i + 1]; film[3 * px + 2] = in[3 * i + 2]; }
  }
}

// ---- debug_normal AOV (the reference's cargo feature `debug_normal`, pt.rs:113-118) ---------------------------------------------------
// Takes the place of logic + special + shade in the wavefront loop: every live path is a camera ray that k_extend has just traced.
// First surface hit -> colour = 0.5 * inter.normal + 0.5 (the interpolated, instance-transformed normal of Triangle::intersect /
// Sphere::intersect, not the normal-mapped shading frame); miss -> the environment as at depth 0 (pt.rs:97-111); path ends.
__global__ void __launch_bounds__(SPT_LOGIC_BLOCK) k_aov_normal(const DScene sc, const RenderParams P) {
  if (blockIdx.x == 0 && threadIdx.x == 0) { P.ctr->it.live = 0; P.ctr->it.next_ext = 0; P.ctr->it.sq = 0; P.ctr->it.next_con = 0; P.ctr->it.n_ext[P.parity] = 0; P.ctr->it.drain = 0; }
  uint32_t nonfinite = 0;
  for (int slot = blockIdx.x * blockDim.x + threadIdx.x; slot < P.pool.n; slot += gridDim.x * blockDim.x) {
    const uint32_t flags = P.pool.flags[slot];
    if (!(flags & 1u)) continue;
    PathRegs s; HitRec hit;
    load_path(P, slot, flags, s, hit);
    if (hit.inst >= 0) {
      HitGeom g;
      load_hit(sc, s.ro, s.rd, hit, g);
      film_add(P, s.pixel, g.n * 0.5f + mk3(0.5f), nonfinite);
    } else if (sc.env_tex) {
      float3 env; float env_pdf;
      env_eval_dir(sc, s.rd, env, env_pdf);
      film_add(P, s.pixel, env, nonfinite);
    }
    P.pool.flags[slot] = 0u;
  }
  warp_sum_to(&P.ctr->nonfinite, nonfinite);
}

// ---- BxDF test hook = BxdfT::{sample,pdf,bxdf,is_delta} with caller-supplied random numbers -----------------------------------
__global__ void k_bxdf_batch(const DScene sc, const SptBxdfQuery* __restrict__ q, long long n, SptBxdfResult* __restrict__ out) {
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
    const SptBxdfQuery Q = q[i];
    HitGeom g; g.p = mk3(0.f); g.n = mk3(0, 0, 1); g.tan = mk3(1, 0, 0); g.bitan = mk3(0, 1, 0);
    g.uv = make_float2(Q.texcoord[0], Q.texcoord[1]); g.duvdx = g.duvdy = make_float2(0.f, 0.f);
    const Bsdf b = make_bsdf<-1>(sc, sc.materials[Q.material], texin_from(g));
    DrawSrc rng; rng.state = 0; rng.list = q[i].rand; rng.idx = 0; rng.n = 8;
    const float3 wo = mk3(Q.wo[0], Q.wo[1], Q.wo[2]), wi = mk3(Q.wi[0], Q.wi[1], Q.wi[2]);
    BsdfSample s;
    bsdf_sample<-1>(b, wo, rng, s);
    SptBxdfResult R;
    if (s.need_probe) { rng.next(); rng.next(); rng.next(); s.lobe = -1; } // no scene in the hook: probe cannot run
    R.sample_wi[0] = s.wi.x; R.sample_wi[1] = s.wi.y; R.sample_wi[2] = s.wi.z;
    R.sample_f[0] = s.f.x; R.sample_f[1] = s.f.y; R.sample_f[2] = s.f.z;
    R.sample_pdf = s.pdf; R.sample_lobe = s.lobe; R.sample_dir = s.dir; R.n_rand_used = rng.idx;
    const float3 f = bsdf_eval<-1>(b, wo, wi);
    R.eval_f[0] = f.x; R.eval_f[1] = f.y; R.eval_f[2] = f.z; R.eval_pdf = bsdf_pdf<-1>(b, wo, wi);
    R.is_delta = bsdf_is_delta<-1>(b) ? 1 : 0; R.bxdf_tag = b.tag;
    out[i] = R;
  }
}

namespace spt {
void launch_logic(const DScene& sc, const RenderParams& P, int grid, cudaStream_t st) { k_logic<<<grid, SPT_LOGIC_BLOCK, 0, st>>>(sc, P); }
void launch_special(const DScene& sc, const RenderParams& P, int grid, cudaStream_t st) { k_special<<<grid, SPT_BLOCK, 0, st>>>(sc, P); }
void launch_generate(const DScene& sc, const RenderParams& P, int grid, cudaStream_t st) { k_generate<<<grid, SPT_LOGIC_BLOCK, 0, st>>>(sc, P); }
void launch_resolve(const long long* film, int n_pixels, float inv_spp, float* out_mean, uint8_t* out_rgb8, int grid, cudaStream_t st) {
  k_resolve<<<grid, 256, 0, st>>>(film, n_pixels, inv_spp, out_mean, out_rgb8);
}
void launch_resolve_box(const RenderParams& P, int k0, int n_k, float radius, int radius_int, float* out_mean, uint8_t* out_rgb8, int grid, cudaStream_t st) {
  k_resolve_box<<<grid, 128, 0, st>>>(P, k0, n_k, radius, radius_int, out_mean, out_rgb8);
}
void launch_pack_tiles(const long long* film, const uint32_t* tiles, int n_tiles, const RenderParams& P, long long* out, int grid, cudaStream_t st) {
  k_pack_tiles<<<grid, 128, 0, st>>>(film, tiles, n_tiles, P, out);
}
void launch_unpack_tiles(long long* film, const uint32_t* all_tiles, int per_rank, int world, int root, const RenderParams& P, const long long* in, int grid, cudaStream_t st) {
  k_unpack_tiles<<<grid, 128, 0, st>>>(film, all_tiles, per_rank, world, root, P, in);
}
void launch_aov_normal(const DScene& sc, const RenderParams& P, int grid, cudaStream_t st) { k_aov_normal<<<grid, SPT_LOGIC_BLOCK, 0, st>>>(sc, P); }
void launch_bxdf(const DScene& sc, const SptBxdfQuery* q, long long n, SptBxdfResult* out, int grid, cudaStream_t st) {
  k_bxdf_batch<<<grid, SPT_BLOCK, 0, st>>>(sc, q, n, out);
}
void launch_shade(int mat, const DScene& sc, const RenderParams& P, bool count, bool lean, int grid, cudaStream_t st) {
  if (mat >= 0 && mat <= 3) { if (lean) launch_shade_a_lean(mat, sc, P, count, grid, st); else launch_shade_a(mat, sc, P, count, grid, st); }
  else if (mat >= 4 && mat <= 6) { if (lean) launch_shade_b_lean(mat, sc, P, count, grid, st); else launch_shade_b(mat, sc, P, count, grid, st); }
  else if (mat == 7 && lean) launch_shade_c_lean(mat, sc, P, count, grid, st);
  else launch_shade_c(mat, sc, P, count, grid, st);
}
} // namespace spt
