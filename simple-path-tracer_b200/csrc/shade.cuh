// shade.cuh — per-hit shading helpers for the wavefront kernels: hit reconstruction, shading frame, emitters,
// light sampling with MIS pdfs, homogeneous media.  Each function cites the reference code it stands for.
#pragma once
#include "bsdf.cuh"
#include "traverse.cuh"

struct Frame { // Coordinate, core/coord.rs:3-8
  float3 x, y, z, hemi;
};
SPT_D Frame frame_from_tangent_normal(float3 t, float3 n, float3 hemi) { // coord.rs:11-23
  Frame f;
  f.z = n; f.y = norm3(cross3(n, t)); f.x = cross3(f.y, f.z); f.hemi = hemi;
  return f;
}
SPT_D float3 to_local(const Frame& f, float3 w) { return mk3(dot3(f.x, w), dot3(f.y, w), dot3(f.z, w)); }
SPT_D float3 to_world(const Frame& f, float3 l) { return f.x * l.x + f.y * l.y + f.z * l.z; }

struct HitGeom { // the parts of Intersection (core/intersection.rs:6-18) shading needs
  float3 p, n, tan, bitan;
  float2 uv, duvdx, duvdy;
  int inst, surface;
  float prim_pdf; // primitive.pdf(inter) before the instance Jacobian (triangle.rs:272-280 / sphere.rs:118-120, / N)
};

SPT_D float3 mat_rows_vec(float4 r0, float4 r1, float4 r2, float3 v) {
  return mk3(r0.x * v.x + r0.y * v.y + r0.z * v.z, r1.x * v.x + r1.y * v.y + r1.z * v.z, r2.x * v.x + r2.y * v.y + r2.z * v.z);
}

SPT_D void sphere_frame(float3 n, float3& tan, float3& bitan) { // sphere.rs:66-77
  const float st = sqrtf(1.0f - n.y * n.y);
  if (st != 0.0f) { bitan = n * (-n.y / st); bitan.y = st; tan = cross3(bitan, n); }
  else if (n.y > 0.f) { bitan = mk3(1, 0, 0); tan = mk3(0, 0, 1); }
  else { bitan = mk3(-1, 0, 0); tan = mk3(0, 0, -1); }
}

struct TriVerts { float3 p[3], n[3], t[3], b[3]; float2 uv[3]; };
SPT_D void load_tri(const DScene& sc, int mesh, int tri, TriVerts& tv) { // by file-order triangle id (light sampling)
  const DMesh m = sc.meshes[mesh];
  const uint32_t* ix = sc.indices + 3 * ((size_t)m.idx_base + tri);
#pragma unroll
  for (int k = 0; k < 3; k++) {
    const float4* vp = sc.verts + 4 * ((size_t)m.vert_base + ix[k]);
    const float4 a = ldg4(vp), b = ldg4(vp + 1), c = ldg4(vp + 2), d = ldg4(vp + 3);
    tv.p[k] = mk3(a); tv.n[k] = mk3(b); tv.uv[k] = make_float2(a.w, b.w); tv.t[k] = mk3(c); tv.b[k] = mk3(d);
  }
}
SPT_D float tri_area(const TriVerts& tv) { return len3(cross3(tv.p[1] - tv.p[0], tv.p[2] - tv.p[0])) * 0.5f; }

// Rebuilds the Intersection the reference fills during traversal (triangle.rs:188-213, sphere.rs:63-80,
// instance.rs:97-104) from the compact hit record — once per accepted hit instead of on every closer candidate.
// Triangle attributes come from the pre-gathered `shade_tris` record of the hit triangle (9 x 16-byte loads).
SPT_D void load_hit(const DScene& sc, float3 o, float3 d, const HitRec& h, HitGeom& g) {
  const DInstanceShade S = sc.inst_shade[h.inst];
  float3 n, tan, bitan;
  if (h.tri < 0) {
    const DInstance I = sc.inst[h.inst];
    const float3 o2 = xmat_point(I.inv0, I.inv1, I.inv2, o), d2 = xmat_vec(I.inv0, I.inv1, I.inv2, d);
    const float4 cr = sc.spheres[I.root];
    n = (xpoint_at(o2, d2, h.t) - mk3(cr)) / cr.w;
    sphere_frame(n, tan, bitan);
    g.uv = sphere_uv(n);
    g.prim_pdf = 0.25f * SPT_INV_PI;
    g.surface = I.surface;
  } else {
    const float4* q = sc.shade_tris + 9 * (size_t)h.tri;
    const float4 n0 = ldg4(q), n1 = ldg4(q + 1), n2 = ldg4(q + 2), t0 = ldg4(q + 3), t1 = ldg4(q + 4), t2 = ldg4(q + 5);
    const float4 b0 = ldg4(q + 6), b1 = ldg4(q + 7), b2 = ldg4(q + 8);
    const float v = h.u, w = h.v, u = 1.0f - v - w;
    n = norm3(mk3(n0) * u + mk3(n1) * v + mk3(n2) * w);
    g.uv = make_float2(n0.w * u + n2.w * v + t1.w * w, n1.w * u + t0.w * v + t2.w * w);
    tan = mk3(t0) * u + mk3(t1) * v + mk3(t2) * w;
    bitan = mk3(b0) * u + mk3(b1) * v + mk3(b2) * w;
    g.prim_pdf = b0.w; // Q6; bvh.rs:300-302
    g.surface = sc.inst[h.inst].surface;
  }
  g.p = xpoint_at(o, d, h.t);
  g.n = norm3(mat_rows_vec(S.it0, S.it1, S.it2, n)); // Transform::transform_normal3a, transform.rs:29-31
  g.tan = mat_rows_vec(S.fwd0, S.fwd1, S.fwd2, tan);
  g.bitan = mat_rows_vec(S.fwd0, S.fwd1, S.fwd2, bitan);
  g.inst = h.inst;
  g.duvdx = g.duvdy = make_float2(0.f, 0.f);
}

SPT_D bool solve2x2(float a00, float a01, float a10, float a11, float bx, float by, float& x1, float& x2) { // intersection.rs:104-118
  const float det = a00 * a11 - a10 * a01;
  if (det == 0.0f) return false;
  x2 = (by * a00 - bx * a01) / det;
  x1 = fabsf(a00) > fabsf(a01) ? (bx - a10 * x2) / a00 : (by - a11 * x2) / a01;
  return true;
}
// Intersection::calc_differential (intersection.rs:28-83); aux rays = the camera rays one pixel-footprint away.
SPT_D void calc_differential(HitGeom& g, float3 ax_o, float3 ax_d, float3 ay_o, float3 ay_d) {
  const float dd = dot3(g.p, g.n);
  const float tx = (dd - dot3(ax_o, g.n)) / dot3(ax_d, g.n), ty = (dd - dot3(ay_o, g.n)) / dot3(ay_d, g.n);
  const float3 dpdx = ax_o + ax_d * tx - g.p, dpdy = ay_o + ay_d * ty - g.p;
  const float nx = fabsf(g.n.x), ny = fabsf(g.n.y), nz = fabsf(g.n.z);
  float bxx, bxy, byx, byy, a00, a01, a10, a11;
  if (nx >= ny && nx >= nz) { bxx = dpdx.y; bxy = dpdx.z; byx = dpdy.y; byy = dpdy.z; a00 = g.tan.y; a01 = g.tan.z; a10 = g.bitan.y; a11 = g.bitan.z; }
  else if (ny >= nz) { bxx = dpdx.z; bxy = dpdx.x; byx = dpdy.z; byy = dpdy.x; a00 = g.tan.z; a01 = g.tan.x; a10 = g.bitan.z; a11 = g.bitan.x; }
  else { bxx = dpdx.x; bxy = dpdx.y; byx = dpdy.x; byy = dpdy.y; a00 = g.tan.x; a01 = g.tan.y; a10 = g.bitan.x; a11 = g.bitan.y; }
  float x1, x2;
  if (solve2x2(a00, a01, a10, a11, bxx, bxy, x1, x2)) g.duvdx = make_float2(x1, x2);
  if (solve2x2(a00, a01, a10, a11, byx, byy, x1, x2)) g.duvdy = make_float2(x1, x2);
}

SPT_D TexIn texin_from(const HitGeom& g) { // impl From<&Intersection> for TextureInput, texture/mod.rs:149-163
  TexIn t;
  t.specified = mk3(0.f); t.position = g.p; t.normal = g.n; t.tangent = g.tan; t.bitangent = g.bitan;
  t.uv = g.uv; t.duvdx = g.duvdx; t.duvdy = g.duvdy; t.mode = 1; t.wrap = 0;
  return t;
}

SPT_D float3 surface_emissive(const DScene& sc, const DSurface& s, const TexIn& in) { // surface.rs:49-55
  const float3 e = mk3(s.emissive[0], s.emissive[1], s.emissive[2]);
#if SPT_LEAN
  return e;
#else
  return s.emissive_map >= 0 ? e * tex_color(sc, s.emissive_map, in) : e;
#endif
}
SPT_D Frame surface_frame(const DScene& sc, const DSurface& s, float3 ray_d, const HitGeom& g, const TexIn& in) { // surface.rs:65-95
  float3 sn = g.n;
  if (!SPT_LEAN && s.normal_map >= 0) {
    const float3 v = tex_color(sc, s.normal_map, in) * 2.0f - mk3(1.f);
    const float3 nl = norm3(v);
    sn = norm3(nl.x * norm3(g.tan) + nl.y * norm3(g.bitan) + nl.z * g.n);
  }
  const bool hit_back = dot3(ray_d, g.n) > 0.f;
  return frame_from_tangent_normal(g.tan, (s.double_sided && hit_back) ? -sn : sn, hit_back ? -g.n : g.n);
}

// ---- media (medium/homogeneous.rs, medium/util.rs) ---------------------------------------------------------
SPT_D float hg_phase(float g, float cs) { // util.rs:1-7
  const float g2 = g * g;
  float den = 1.0f + g2 + 2.0f * g * cs;
  den = den * sqrtf(den);
  return 0.25f * SPT_INV_PI * (1.0f - g2) / den;
}
SPT_D float hg_cdf_inverse(float g, float u) { // util.rs:10-19
  if (fabsf(g) < 0.01f) return 1.0f - 2.0f * u;
  const float g2 = g * g, t = (1.0f - g2) / (1.0f - g + 2.0f * g * u);
  return 0.5f * (1.0f + g2 - t * t) / g;
}
SPT_D float3 medium_local_to_world(float3 wo, float3 wl) { // util.rs:21-31
  float3 v = fabsf(wo.y) < 0.99f ? mk3(0, 1, 0) : mk3(1, 0, 0);
  const float3 u = norm3(cross3(v, wo));
  v = cross3(wo, u);
  return u * wl.x + v * wl.y + wo * wl.z;
}

// ---- lights (light/*.rs) and light samplers (light_sampler/*.rs) --------------------------------------------
SPT_D void env_lookup(const DScene& sc, float theta, float phi, float3& c, float& pdf) { // environment.rs:52-84 (Q7)
  const int W = sc.env_w, H = sc.env_h;
  const float x = phi * 0.5f * SPT_INV_PI * (float)W;
  int x1 = f2i_sat(roundf(x)), x0 = x1 - 1;
  const float xt = x - (float)x0 - 0.5f;
  x0 = min(max(x0, 0), W - 1); x1 = min(max(x1, 0), W - 1);
  const float y = theta * SPT_INV_PI * (float)H;
  int y1 = f2i_sat(roundf(y)), y0 = y1 - 1;
  const float yt = y - (float)y0 - 0.5f;
  y0 = min(max(y0, 0), H - 1); y1 = min(max(y1, 0), H - 1);
  const float4 c00 = ldg4(sc.env_tex + (size_t)y0 * W + x0), c01 = ldg4(sc.env_tex + (size_t)y1 * W + x0);
  const float4 c10 = ldg4(sc.env_tex + (size_t)y0 * W + x1), c11 = ldg4(sc.env_tex + (size_t)y1 * W + x1);
  const float4 c0 = lerp4(c00, c01, yt), c1 = lerp4(c10, c11, yt); // rgb = texel, w = alias-table probability
  const float4 cc = lerp4(c0, c1, xt);
  c = mk3(cc) * mk3(sc.env_scale[0], sc.env_scale[1], sc.env_scale[2]);
  pdf = c0.w * (1.0f - xt) * c1.w * xt; // product, as written (environment.rs:81)
}
SPT_D void env_eval_dir(const DScene& sc, float3 wi, float3& c, float& pdf) { // environment.rs:128-133
  env_lookup(sc, acosf(wi.y), atan2f(wi.x, wi.z) + SPT_PI, c, pdf);
}
SPT_D int alias_pick(const float2* __restrict__ tab, int n, float u, int& x_out) { // AliasTable::sample, alias_table.rs:58-67
  const float temp = u * (float)n;
  const int x = min((int)temp, n - 1);
  const float y = temp - (float)x;
  const float2 e = __ldg(tab + x);
  x_out = x;
  return y < e.x ? x : __float_as_int(e.y);
}

struct LightSample { float3 dir, L; float pdf, dist; bool delta; };

// LightT::sample for one light.  ShapeLight: shape_light.rs:21-43 with Instance::sample (instance.rs:115-129),
// BvhAccel::sample (bvh.rs:293-298; triangle picked in file order — a permutation of the reference's uniform pick),
// Triangle::sample (triangle.rs:224-270, Q6) and Sphere::sample (sphere.rs:90-116, Q5).
template <class RNG>
SPT_D void light_sample(const DScene& sc, const DLight& l, float3 pos, RNG& rng, LightSample& ls) {
  ls.delta = l.tag <= 2;
  switch (l.tag) {
    case 0: // directional.rs:27-29
      ls.dir = mk3(-l.direction[0], -l.direction[1], -l.direction[2]); ls.pdf = 1.0f;
      ls.L = mk3(l.strength[0], l.strength[1], l.strength[2]); ls.dist = FLT_MAX;
      return;
    case 1: case 2: { // point.rs:24-30, spot.rs:59-65
      float3 s = mk3(l.position[0], l.position[1], l.position[2]) - pos;
      const float d2 = dot3(s, s), d = sqrtf(d2);
      s = s / d;
      float3 L = mk3(l.strength[0], l.strength[1], l.strength[2]);
      if (l.tag == 2) { // SpotLight::strength, spot.rs:50-55
        const float3 dir = mk3(l.direction[0], l.direction[1], l.direction[2]);
        L = L * clampf_((dot3(dir, -s) - l.cos_outer) / fmaxf(l.cos_inner - l.cos_outer, 0.0001f), 0.0f, 1.0f);
      }
      ls.dir = s; ls.pdf = 1.0f; ls.L = L / d2; ls.dist = d;
      return;
    }
    case 3: {
      const DInstance I = sc.inst[l.instance];
      const DInstanceShade S = sc.inst_shade[l.instance];
      float3 p, n, tan, bitan; float2 uv; float ppdf;
      if ((I.kind & 0xff) == 0) {
        const float u0 = rng.next(), u1 = rng.next(); // Rng::uniform_on_sphere, rng.rs:56-63
        float sp, cp; sincos_2pi(u0, sp, cp);
        const float ct = 1.0f - 2.0f * u1, st = sqrtf(1.0f - ct * ct);
        n = mk3(st * cp, st * sp, ct);
        const float4 cr = sc.spheres[I.root];
        p = mk3(cr) + n * cr.w;
        uv = sphere_uv(n);
        sphere_frame(n, tan, bitan);
        ppdf = 0.25f * SPT_INV_PI;
      } else {
        const int nt = sc.meshes[I.mesh].n_tris;
        const int ti = min((int)(rng.next() * (float)nt), nt - 1);
        const float r0 = rng.next(), r1 = rng.next();
        const float r0s = sqrtf(r0);
        const float u = 1.0f - r0s, v = r0s * (1.0f - r1), w = 1.0f - u - v;
        TriVerts tv; load_tri(sc, I.mesh, ti, tv);
        p = tv.p[0] * u + tv.p[1] * v + tv.p[2] * w;
        n = tv.n[0] * u + tv.n[1] * v + tv.n[2] * w;
        tan = tv.t[0] * u + tv.t[1] * v + tv.t[2] * w;
        bitan = tv.b[0] * u + tv.b[1] * v + tv.b[2] * w;
        uv = make_float2(tv.uv[0].x * u + tv.uv[1].x * v + tv.uv[2].x * w, tv.uv[0].y * u + tv.uv[1].y * v + tv.uv[2].y * w);
        ppdf = (1.0f / fmaxf(tri_area(tv), 0.001f)) / (float)nt;
      }
      const float original_area = len3(cross3(tan, bitan));
      HitGeom g;
      g.p = mat_rows_vec(S.fwd0, S.fwd1, S.fwd2, p) + mk3(S.fwd0.w, S.fwd1.w, S.fwd2.w);
      g.n = norm3(mat_rows_vec(S.it0, S.it1, S.it2, n));
      g.bitan = mat_rows_vec(S.fwd0, S.fwd1, S.fwd2, bitan);
      g.tan = mat_rows_vec(S.fwd0, S.fwd1, S.fwd2, tan);
      g.uv = uv; g.duvdx = g.duvdy = make_float2(0.f, 0.f);
      const float transformed_area = len3(cross3(g.tan, g.bitan));
      const float pdf_area = ppdf * original_area / transformed_area;
      const DSurface sf = sc.surfaces[I.surface];
      float3 em = surface_emissive(sc, sf, texin_from(g));
      const float3 lv = g.p - pos;
      const float d2 = dot3(lv, lv), d = sqrtf(d2);
      const float3 dir = lv / d;
      float cs;
      if (sf.double_sided) cs = fabsf(dot3(dir, g.n));
      else { cs = dot3(dir, -g.n); if (!(cs > 0.f)) { cs = 1.0f; em = mk3(0.f); } }
      ls.dir = dir; ls.pdf = pdf_area * d2 / fmaxf(cs, 0.001f); ls.L = em; ls.dist = d;
      return;
    }
    default: { // EnvLight::sample, environment.rs:111-126
      const float u = rng.next();
      int x0;
      const int ind = alias_pick(sc.env_alias, sc.env_w * sc.env_h, u, x0);
      const int x = ind % sc.env_w, y = ind / sc.env_w;
      const float rx = rng.next(), ry = rng.next();
      const float theta = ((float)y + ry) / (float)sc.env_h * SPT_PI;
      const float phi = ((float)x + rx) / (float)sc.env_w * 2.0f * SPT_PI;
      float st, ct, sp, cp; sincospif(((float)y + ry) / (float)sc.env_h, &st, &ct); sincos_2pi(((float)x + rx) / (float)sc.env_w, sp, cp);
      ls.dir = mk3(st * sp, ct, st * cp);
      env_lookup(sc, theta, phi, ls.L, ls.pdf);
      ls.dist = INFINITY;
      return;
    }
  }
}

// LightSamplerT::sample_light: uniform.rs:34-41 / power_is.rs:46-56
template <class RNG>
SPT_D void sample_one_light(const DScene& sc, float3 pos, RNG& rng, LightSample& ls) {
  int index; float pick;
  const float u = rng.next();
  if (sc.light_sampler == 0) { index = min((int)(u * (float)sc.n_lights), sc.n_lights - 1); pick = sc.num_light_inv; }
  else { int x0; index = alias_pick(sc.light_alias, sc.n_lights, u, x0); pick = sc.light_prob[index]; }
  light_sample(sc, sc.lights[index], pos, rng, ls);
  ls.pdf = sc.light_sampler == 0 ? ls.pdf * pick : pick * ls.pdf;
}
// LightSamplerT::pdf_shape_light: uniform.rs:43-67 / power_is.rs:58-88, with Instance::pdf (instance.rs:131-141)
SPT_D float pdf_shape_light(const DScene& sc, float3 from, const HitGeom& g, bool double_sided) {
  const DInstance I = sc.inst[g.inst];
  const float3 t0 = mat_rows_vec(I.inv0, I.inv1, I.inv2, g.tan), b0 = mat_rows_vec(I.inv0, I.inv1, I.inv2, g.bitan);
  const float primitive_pdf = g.prim_pdf * len3(cross3(t0, b0)) / len3(cross3(g.tan, g.bitan));
  const float3 lv = g.p - from;
  const float d2 = dot3(lv, lv);
  const float3 dir = lv / sqrtf(d2);
  float cs;
  if (double_sided) cs = fabsf(dot3(dir, g.n));
  else { cs = dot3(dir, -g.n); if (!(cs > 0.f)) cs = 1.0f; }
  const float local = primitive_pdf * d2 / fmaxf(cs, 0.00001f);
  if (sc.light_sampler == 0) return local * sc.num_light_inv;
  const int li = sc.inst_light[g.inst];
  return local * (li >= 0 ? sc.light_prob[li] : 0.0f);
}
SPT_D float pdf_env_light(const DScene& sc) { // uniform.rs:69-75 / power_is.rs:90-96
  if (sc.env_light_index < 0) return 1.0f;
  return sc.light_sampler == 0 ? sc.num_light_inv : sc.light_prob[sc.env_light_index];
}
SPT_D float power_heuristic(float p0, float p1) { return p0 * p0 / (p0 * p0 + p1 * p1); } // pt.rs:298-302
