// oracle/bxdf.h — TEST INFRASTRUCTURE ONLY.  Restates bxdf/{util,fresnel,microfacet(GGX),lambert,pseudo,
// substrate,microfacet_*,specular_*}.rs.  PARITY UNPINNED (see vec.h).
#pragma once
#include "texture.h"

namespace orc {

struct Surface;
Coordinate surface_coord(const Surface* s, const Ray& ray, const Intersection& inter); // core/surface.rs:65-95

// ---- bxdf/util.rs --------------------------------------------------------------------------------
inline float pow2(float x) { return x * x; }
inline float pow5(float x) { return x * x * x * x * x; }
inline Vec3 reflect(Vec3 i) { return {-i.x, -i.y, i.z}; }                         // :3-5
inline Vec3 reflect_n(Vec3 i, Vec3 n) { return 2.0f * dot(i, n) * n - i; }          // :7-9
inline bool refract(Vec3 i, float ior, Vec3& o) {                                   // :11-24
  float ratio = i.z >= 0.0f ? 1.0f / ior : ior;
  float oz2 = 1.0f - (1.0f - i.z * i.z) * ratio * ratio;
  if (oz2 >= 0.0f) {
    float oz = i.z >= 0.0f ? -std::sqrt(oz2) : std::sqrt(oz2);
    o = {-i.x * ratio, -i.y * ratio, oz};
    return true;
  }
  return false;
}
inline bool refract_n(Vec3 i, Vec3 n, float ior, Vec3& o) { // :26-46
  float cos_i = dot(i, n);
  if (cos_i >= 0.0f) {
    float ratio = 1.0f / ior;
    float oz2 = 1.0f - (1.0f - cos_i * cos_i) * ratio * ratio;
    if (oz2 >= 0.0f) { o = (ratio * cos_i - std::sqrt(oz2)) * n - ratio * i; return true; }
    return false;
  }
  float ratio = ior;
  float oz2 = 1.0f - (1.0f - cos_i * cos_i) * ratio * ratio;
  if (oz2 >= 0.0f) { o = (std::sqrt(oz2) + ratio * cos_i) * n - ratio * i; return true; }
  return false;
}
inline float fresnel_n(float ior, Vec3 i, Vec3 n) { // :56-81
  float i_ior = dot(i, n) >= 0.0f ? 1.0f : ior;
  float o_ior = dot(i, n) >= 0.0f ? ior : 1.0f;
  Vec3 rf;
  if (refract_n(i, n, ior, rf)) {
    float idotn = std::fabs(dot(i, n)), rdotn = std::fabs(dot(rf, n));
    float denom = i_ior * idotn + o_ior * rdotn, num = i_ior * idotn - o_ior * rdotn;
    float rs = num / denom; rs = rs * rs;
    denom = i_ior * rdotn + o_ior * idotn; num = i_ior * rdotn - o_ior * idotn;
    float rp = num / denom; rp = rp * rp;
    return 0.5f * (rs + rp);
  }
  return 1.0f;
}
inline float fresnel(float ior, Vec3 i) { return fresnel_n(ior, i, {0, 0, 1}); } // :52-54
inline Color fresnel_conductor_n(Color ior, Color ior_k, Vec3 i, Vec3 n) {        // :87-112
  float cs = dot(i, n);
  Color ir = cs >= 0.0f ? ior : WHITE / ior;
  Color kr = cs >= 0.0f ? ior_k : WHITE / ior_k;
  float cos2 = cs * cs, sin2 = 1.0f - cos2;
  Color ir2 = ir * ir, kr2 = kr * kr;
  Color t0 = ir2 - kr2 - Color::gray(sin2);
  Color a2_b2 = (t0 * t0 + 4.0f * ir2 * kr2).sqrt();
  Color t1 = a2_b2 + Color::gray(cos2);
  Color a = (0.5f * (a2_b2 + t0)).sqrt();
  Color t2 = 2.0f * cs * a;
  Color rs = (t1 - t2) / (t1 + t2);
  Color t3 = cos2 * a2_b2 + Color::gray(sin2 * sin2);
  Color t4 = t2 * sin2;
  Color rp = rs * (t3 - t4) / (t3 + t4);
  return 0.5f * (rs + rp);
}
inline Color schlick_fresnel_with_r0(Color r0, float cs) { return r0 + (WHITE - r0) * pow5(1.0f - cs); } // :120-122
inline float fresnel_moment1(float eta) { // :124-134
  float e2 = eta * eta, e3 = e2 * eta, e4 = e3 * eta, e5 = e4 * eta;
  if (eta < 1.0f) return 0.45966f - 1.73965f * eta + 3.37668f * e2 - 3.904945f * e3 + 2.49277f * e4 - 0.68441f * e5;
  return -4.61686f + 11.1136f * eta - 10.4646f * e2 + 5.11455f * e3 - 1.27198f * e4 + 0.12746f * e5;
}
inline Vec3 half_from_reflect(Vec3 i, Vec3 o) { return i.z >= 0.0f ? normalize(i + o) : -normalize(i + o); } // :136-142
inline Vec3 half_from_refract(Vec3 i, Vec3 o, float ior) {                                                   // :144-154
  Vec3 h = i.z >= 0.0f ? normalize(i + ior * o) : normalize(ior * i + o);
  if (h.z < 0.0f) h = -h;
  return h;
}
inline float ggx_ndf_aniso(Vec3 h, float ax, float ay) { // :161-164
  return FRAC_1_PI / fmax_(ax * ay * pow2(pow2(h.x / ax) + pow2(h.y / ay) + pow2(h.z)), 0.0001f);
}
inline float smith_g1_aniso(Vec3 v, float ax, float ay) { // :172-174
  return 2.0f / (1.0f + std::sqrt(1.0f + (pow2(ax * v.x) + pow2(ay * v.y)) / fmax_(pow2(v.z), 0.0001f)));
}
inline float smith_separable_visible_aniso(Vec3 v, Vec3 l, float ax, float ay) { // :176-180
  float a = std::fabs(v.z) + std::sqrt(pow2(ax * v.x) + pow2(ay * v.y) + pow2(v.z));
  float b = std::fabs(l.z) + std::sqrt(pow2(ax * l.x) + pow2(ay * l.y) + pow2(l.z));
  return 1.0f / (a * b);
}
inline float ggx_smith_vndf_pdf(Vec3 h, Vec3 v, float ax, float ay) { // :189-194
  if (!(v.z >= 0.0f)) v = -v;
  return smith_g1_aniso(v, ax, ay) * ggx_ndf_aniso(h, ax, ay) * fmax_(dot(v, h), 0.0f) / fmax_(v.z, 0.0001f);
}
inline Vec3 ggx_smith_vndf_sample(Vec3 ve, float ax, float ay, float r0, float r1, float& pdf) { // :196-224
  if (!(ve.z >= 0.0f)) ve = -ve;
  Vec3 vh = normalize(Vec3(ax * ve.x, ay * ve.y, ve.z));
  float len_sqr = vh.x * vh.x + vh.y * vh.y;
  Vec3 t1v = len_sqr > 0.0f ? Vec3(-vh.y, vh.x, 0.0f) / std::sqrt(len_sqr) : Vec3(1, 0, 0);
  Vec3 t2v = cross(vh, t1v);
  float r = std::sqrt(r0);
  float phi = 2.0f * PI * r1;
  float t1 = r * std::cos(phi), t2 = r * std::sin(phi);
  float s = 0.5f * (1.0f + vh.z);
  t2 = (1.0f - s) * std::sqrt(1.0f - t1 * t1) + s * t2;
  Vec3 nh = t1 * t1v + t2 * t2v + std::sqrt(fmax_(1.0f - t1 * t1 - t2 * t2, 0.0f)) * vh;
  Vec3 ne = normalize(Vec3(ax * nh.x, ay * nh.y, fmax_(nh.z, 0.0f)));
  pdf = ggx_smith_vndf_pdf(ne, ve, ax, ay);
  return ne;
}

// ---- bxdf/fresnel.rs -----------------------------------------------------------------------------
struct Fresnel {
  enum Kind { kDielectric, kConductor, kSchlick } kind = kDielectric;
  float ior_v = 1; Color eta, k, r0;
  Color fresnel(Vec3 i, Vec3 n) const {
    switch (kind) {
      case kDielectric: return Color::gray(fresnel_n(ior_v, i, n));  // :32-36
      case kConductor: return fresnel_conductor_n(eta, k, i, n);      // :53-57
      default: return schlick_fresnel_with_r0(r0, dot(i, n));         // :72-75
    }
  }
  float ior() const { // :38,58,76-79
    if (kind == kDielectric) return ior_v;
    if (kind == kConductor) return 1.0f;
    float s = std::sqrt(r0.luminance());
    return (1.0f - s) / (1.0f + s);
  }
};

struct SubsurfaceSample { Vec3 pi; Coordinate coord_pi; Color sp; float pdf_pi = 0; }; // bxdf/mod.rs:65-70
struct BxdfSample {                                                                   // bxdf/mod.rs:72-78
  Vec3 wi; LobeType lobe = kDiffuse; DirType dir = kReflect; bool ty_subsurface = false;
  Color bxdf; float pdf = 0; bool has_ss = false; SubsurfaceSample ss;
};
struct BxdfInputs { Vec3 po; Coordinate coord_po; Vec3 wo; const Aggregate* scene = nullptr; uint64_t* probe_rays = nullptr; };

inline const std::vector<std::pair<float, float>>& ss_cdf_table() { // substrate.rs:187-196
  static const std::vector<std::pair<float, float>> tab = [] {
    std::vector<std::pair<float, float>> t(512);
    for (int i = 0; i < 512; i++) {
      float x = (float)i / 512.0f;
      x = -2.0f * std::log(1.0f - x);
      float y = 1.0f - std::exp(-x) * 0.25f - std::exp(-x / 3.0f) * 0.75f;
      t[i] = {x, y};
    }
    return t;
  }();
  return tab;
}

// ---- bxdf/substrate.rs ---------------------------------------------------------------------------
struct Substrate {
  enum Kind { kLambert, kDiffuseSub, kSubsurface } kind = kLambert;
  Color reflectance_v;
  float ior = 1; Color bxdf_wo_fresnel; // Diffuse :120-124
  Color d;                              // Subsurface :182-185

  static Substrate lambert(Color r) { Substrate s; s.kind = kLambert; s.reflectance_v = r; return s; }
  static Substrate diffuse(Color r, float ior) { // Diffuse::new :127-137
    Substrate s; s.kind = kDiffuseSub; s.reflectance_v = r; s.ior = ior;
    float fdr = 2.0f * fresnel_moment1(1.0f / ior);
    s.bxdf_wo_fresnel = r * FRAC_1_PI / ((WHITE - r * fdr) * ior * ior);
    return s;
  }
  static float powi4(float x) { float x2 = x * x; return x2 * x2; }
  static Substrate subsurface(Color r, float ior, float ld) { // Subsurface::new :199-209
    Substrate s = diffuse(r, ior); s.kind = kSubsurface;
    s.d = Color(ld / (3.5f + 100.0f * powi4(r.r - 0.33f)), ld / (3.5f + 100.0f * powi4(r.g - 0.33f)), ld / (3.5f + 100.0f * powi4(r.b - 0.33f)));
    return s;
  }
  Color reflectance() const { return reflectance_v; }
  Color sp(float r) const { // :211-215
    Color e1 = (-r / d).exp();
    Color e2 = (-r / d / 3.0f).exp();
    return (e1 + e2) * FRAC_1_PI / (8.0f * d * r);
  }
  static float sample_r(float rand) { // :217-228
    const auto& T = ss_cdf_table();
    for (size_t i = 1; i < T.size(); i++) {
      if (T[i].second >= rand) {
        float t = (rand - T[i - 1].second) / (T[i].second - T[i - 1].second);
        return T[i].first * t + T[i - 1].first * (1.0f - t);
      }
    }
    return -1.0f;
  }
  BxdfSample sample_diffuse_like(const BxdfInputs& in, Rng& rng) const { // Lambert lambert.rs:20-36 / Diffuse :141-160
    Vec3 wi = rng.cosine_weighted_on_hemisphere();
    if (in.wo.z < 0.0f) wi.z = -wi.z;
    BxdfSample s; s.wi = wi; s.lobe = kDiffuse; s.dir = kReflect;
    if (kind == kLambert) s.bxdf = reflectance_v * FRAC_1_PI;
    else { float fi = fresnel(ior, wi); s.bxdf = (1.0f - fi) * bxdf_wo_fresnel; }
    s.pdf = std::fabs(wi.z) * FRAC_1_PI;
    return s;
  }
  BxdfSample sample(const BxdfInputs& in, Rng& rng) const {
    if (kind != kSubsurface) return sample_diffuse_like(in, rng);
    // Subsurface::sample :231-333
    BxdfSample samp; samp.wi = {0, 0, 0}; samp.lobe = kDiffuse; samp.dir = kReflect; samp.ty_subsurface = true;
    samp.bxdf = BLACK; samp.pdf = 1.0f;
    float rand_u = rng.uniform_1d();
    float rand_x, rand_y; rng.uniform_2d(rand_x, rand_y);
    Vec3 pt = in.coord_po.to_world({1, 0, 0}), pb = in.coord_po.to_world({0, 1, 0}), pn = in.coord_po.to_world({0, 0, 1});
    Vec3 st, sb, sn;
    if (rand_u < 0.5f) { rand_u = rand_u * 2.0f; st = pt; sb = pb; sn = pn; }
    else if (rand_u < 0.75f) { rand_u = rand_u * 4.0f - 2.0f; st = pb; sb = pn; sn = pt; }
    else { rand_u = rand_u * 4.0f - 3.0f; st = pn; sb = pt; sn = pb; }
    float sp_d;
    if (rand_u < 1.0f / 3.0f) { rand_u = 3.0f * rand_u; sp_d = d.r; }
    else if (rand_u < 2.0f / 3.0f) { rand_u = 3.0f * rand_u - 1.0f; sp_d = d.g; }
    else { rand_u = 3.0f * rand_u - 2.0f; sp_d = d.b; }
    float sample_r_v = sample_r(rand_x) * sp_d;
    float r_max = ss_cdf_table().back().first * sp_d;
    if (sample_r_v < 0.0f) return samp;
    float pihi = 2.0f * PI * rand_y;
    float pc = std::cos(pihi), ps = std::sin(pihi);
    float sample_l = std::sqrt(r_max * r_max + sample_r_v * sample_r_v);
    Vec3 start_p = in.po + st * pc * sample_r_v + sb * ps * sample_r_v + sn * sample_l;
    Ray ray(start_p, -sn);
    Intersection inter = Intersection::with_t_max(2.0f * sample_l);
    struct Hit { Vec3 p, n; Coordinate c; };
    std::vector<Hit> hits;
    for (;;) { // :286-299 — yields at most one hit
      if (!in.scene) break; // bxdf test hook: no scene, no probe
      if (in.probe_rays) (*in.probe_rays)++;
      if (in.scene->intersect(ray, inter)) {
        Coordinate ct = surface_coord(inter.surface, ray, inter);
        hits.push_back({ray.point_at(inter.t), inter.normal, ct});
        ray.t_min = inter.t + T_MIN_EPS;
      } else break;
    }
    if (hits.empty()) return samp;
    size_t si = std::min((size_t)(rand_u * (float)hits.size()), hits.size() - 1);
    const Hit& h = hits[si];
    Color spv = sp(distance(h.p, in.po));
    Vec3 offset = in.coord_po.to_local(h.p - in.po);
    Vec3 nl = in.coord_po.to_local(h.n);
    float r_xy = std::sqrt(offset.x * offset.x + offset.y * offset.y);
    float r_yz = std::sqrt(offset.y * offset.y + offset.z * offset.z);
    float r_zx = std::sqrt(offset.z * offset.z + offset.x * offset.x);
    float pdf_xy = 0.5f * std::fabs(nl.z) * sp(r_xy).avg();
    float pdf_yz = 0.25f * std::fabs(nl.x) * sp(r_yz).avg();
    float pdf_zx = 0.25f * std::fabs(nl.y) * sp(r_zx).avg();
    float pdf_pi = (pdf_xy + pdf_yz + pdf_zx) / (float)hits.size();
    samp.has_ss = true; samp.ss = {h.p, h.c, spv, pdf_pi};
    Substrate dif = *this; dif.kind = kDiffuseSub;
    BxdfSample ds = dif.sample_diffuse_like(in, rng);
    samp.wi = ds.wi; samp.pdf = ds.pdf; samp.bxdf = ds.bxdf;
    return samp;
  }
  float pdf(Vec3 wo, Vec3 wi) const { return wo.z * wi.z >= 0.0f ? std::fabs(wi.z) * FRAC_1_PI : 1.0f; } // Q10
  Color bxdf(Vec3 wo, Vec3 wi) const {
    if (!(wo.z * wi.z >= 0.0f)) return BLACK;
    if (kind == kLambert) return reflectance_v * FRAC_1_PI;
    float fi = fresnel(ior, wi);
    return (1.0f - fi) * bxdf_wo_fresnel;
  }
};

} // namespace orc
