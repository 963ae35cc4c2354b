// bsdf.cuh — device BxDFs: sample / eval / pdf for the 8 `Bxdf` variants of src/bxdf/mod.rs:92-102, the 3 Fresnel
// models (bxdf/fresnel.rs), anisotropic GGX with Heitz-2018 VNDF sampling (bxdf/util.rs:161-224) and the three
// substrates (bxdf/substrate.rs).  A Bsdf is a flat POD value built per hit from the material's textures
// (MaterialT::bxdf_context, src/material/*.rs) — the reference's nested enums flattened into tags.
// Local frame: z = shading normal.  Quirks kept: Q10 (pdf = 1 across the horizon), Q11, Q12.
#pragma once
#include "texture.cuh"
#include "rng.cuh"

enum { BX_LAMBERT = 0, BX_PSEUDO, BX_MF_CONDUCTOR, BX_MF_DIELECTRIC, BX_MF_PLASTIC, BX_SP_CONDUCTOR, BX_SP_DIELECTRIC, BX_SP_PLASTIC };
enum { FR_DIELECTRIC = 0, FR_CONDUCTOR, FR_SCHLICK };
enum { SUB_LAMBERT = 0, SUB_DIFFUSE, SUB_SUBSURFACE };
enum { LOBE_DIFFUSE = 0, LOBE_GLOSSY, LOBE_SPECULAR };
enum { DIR_REFLECT = 0, DIR_TRANSMIT };

struct Bsdf {
  int tag, fr, sub;
  float ax, ay;   // GGX roughness_x / roughness_y
  float ior;      // dielectric fresnel ior
  float3 c0, c1;  // lambert reflectance | conductor (eta, k) | schlick r0
  float3 s_refl;  // substrate reflectance
  float3 s_base;  // Diffuse::bxdf_wo_fresnel (substrate.rs:127-137)
  float3 s_d;     // Subsurface::d (substrate.rs:199-209)
  float s_ior;
};

struct BsdfSample {
  float3 wi, f;
  float pdf;
  int lobe, dir;
  bool need_probe; // substrate is Subsurface: the caller must trace the probe ray and call bsdf_finish_probe()
};


// Compile-time narrowing by material class: shade kernels are instantiated per material tag MAT (SPT_MAT_*), which
// fixes the Fresnel model and the substrate and leaves at most two Bxdf variants; MAT = -1 keeps everything dynamic.
template <int MAT> SPT_D int tag_of(const Bsdf& b) {
  if (MAT == 0) return BX_LAMBERT;
  if (MAT == 1) return BX_PSEUDO;
  if (MAT == 2) return b.tag == BX_MF_CONDUCTOR ? BX_MF_CONDUCTOR : BX_SP_CONDUCTOR;
  if (MAT == 3) return b.tag == BX_MF_DIELECTRIC ? BX_MF_DIELECTRIC : BX_SP_DIELECTRIC;
  if (MAT >= 4) return b.tag == BX_MF_PLASTIC ? BX_MF_PLASTIC : BX_SP_PLASTIC;
  return b.tag;
}
template <int MAT> SPT_D int fr_of(const Bsdf& b) {
  if (MAT == 2) return FR_CONDUCTOR;
  if (MAT == 5 || MAT == 6) return FR_SCHLICK;
  if (MAT >= 0) return FR_DIELECTRIC;
  return b.fr;
}
template <int MAT> SPT_D int sub_of(const Bsdf& b) {
  if (MAT == 4) return SUB_DIFFUSE;
  if (MAT == 7) return SUB_SUBSURFACE;
  if (MAT >= 0) return SUB_LAMBERT;
  return b.sub;
}

// ---- bxdf/util.rs ---------------------------------------------------------------------------------------
SPT_D float3 refl_z(float3 i) { return mk3(-i.x, -i.y, i.z); }
SPT_D float3 refl_n(float3 i, float3 n) { return (2.0f * dot3(i, n)) * n - i; }
SPT_D bool refr_z(float3 i, float ior, float3& o) { // :11-24
  const float ratio = i.z >= 0.f ? 1.0f / ior : ior;
  const float oz2 = 1.0f - (1.0f - i.z * i.z) * ratio * ratio;
  if (oz2 >= 0.f) { o = mk3(-i.x * ratio, -i.y * ratio, i.z >= 0.f ? -sqrtf(oz2) : sqrtf(oz2)); return true; }
  return false;
}
SPT_D bool refr_n(float3 i, float3 n, float ior, float3& o) { // :26-46
  const float ci = dot3(i, n);
  const float ratio = ci >= 0.f ? 1.0f / ior : ior;
  const float oz2 = 1.0f - (1.0f - ci * ci) * ratio * ratio;
  if (!(oz2 >= 0.f)) return false;
  o = (ci >= 0.f ? (ratio * ci - sqrtf(oz2)) : (sqrtf(oz2) + ratio * ci)) * n - ratio * i;
  return true;
}
SPT_D float fresnel_diel(float ior, float3 i, float3 n) { // fresnel_n :56-81
  const float ci = dot3(i, n);
  const float ii = ci >= 0.f ? 1.0f : ior, oi = ci >= 0.f ? ior : 1.0f;
  float3 rf;
  if (!refr_n(i, n, ior, rf)) return 1.0f;
  const float a = fabsf(ci), b = fabsf(dot3(rf, n));
  const float rs = (ii * a - oi * b) / (ii * a + oi * b);
  const float rp = (ii * b - oi * a) / (ii * b + oi * a);
  return 0.5f * (rs * rs + rp * rp);
}
SPT_D float3 fresnel_cond(float3 eta, float3 k, float3 i, float3 n) { // fresnel_conductor_n :87-112
  const float cs = dot3(i, n);
  const float3 ir = cs >= 0.f ? eta : mk3(1.f) / eta, kr = cs >= 0.f ? k : mk3(1.f) / k;
  const float cos2 = cs * cs, sin2 = 1.0f - cos2;
  const float3 ir2 = ir * ir, kr2 = kr * kr;
  const float3 t0 = ir2 - kr2 - mk3(sin2);
  const float3 a2b2 = sqrt3(t0 * t0 + 4.0f * ir2 * kr2);
  const float3 t1 = a2b2 + mk3(cos2);
  const float3 a = sqrt3(0.5f * (a2b2 + t0));
  const float3 t2 = (2.0f * cs) * a;
  const float3 rs = (t1 - t2) / (t1 + t2);
  const float3 t3 = cos2 * a2b2 + mk3(sin2 * sin2);
  const float3 t4 = t2 * sin2;
  const float3 rp = rs * (t3 - t4) / (t3 + t4);
  return 0.5f * (rs + rp);
}
SPT_D float pow5f(float x) { const float x2 = x * x; return x2 * x2 * x; }
SPT_D float fresnel_moment1(float eta) { // :124-134
  const float e2 = eta * eta, e3 = e2 * eta, e4 = e3 * eta, e5 = e4 * eta;
  return eta < 1.0f ? 0.45966f - 1.73965f * eta + 3.37668f * e2 - 3.904945f * e3 + 2.49277f * e4 - 0.68441f * e5
                    : -4.61686f + 11.1136f * eta - 10.4646f * e2 + 5.11455f * e3 - 1.27198f * e4 + 0.12746f * e5;
}
SPT_D float3 half_refl(float3 i, float3 o) { const float3 h = norm3(i + o); return i.z >= 0.f ? h : -h; } // :136-142
SPT_D float3 half_refr(float3 i, float3 o, float ior) {                                                     // :144-154
  float3 h = i.z >= 0.f ? norm3(i + ior * o) : norm3(ior * i + o);
  return h.z < 0.f ? -h : h;
}
SPT_D float ggx_d(float3 h, float ax, float ay) { return SPT_INV_PI / fmaxf(ax * ay * sqr(sqr(h.x / ax) + sqr(h.y / ay) + sqr(h.z)), 0.0001f); } // :161-164
SPT_D float smith_g1(float3 v, float ax, float ay) { return 2.0f / (1.0f + sqrtf(1.0f + (sqr(ax * v.x) + sqr(ay * v.y)) / fmaxf(sqr(v.z), 0.0001f))); } // :172-174
SPT_D float smith_vis(float3 v, float3 l, float ax, float ay) { // :176-180
  const float a = fabsf(v.z) + sqrtf(sqr(ax * v.x) + sqr(ay * v.y) + sqr(v.z));
  const float b = fabsf(l.z) + sqrtf(sqr(ax * l.x) + sqr(ay * l.y) + sqr(l.z));
  return 1.0f / (a * b);
}
SPT_D float vndf_pdf(float3 h, float3 v, float ax, float ay) { // :189-194
  if (!(v.z >= 0.f)) v = -v;
  return smith_g1(v, ax, ay) * ggx_d(h, ax, ay) * fmaxf(dot3(v, h), 0.f) / fmaxf(v.z, 0.0001f);
}
SPT_D float3 vndf_sample(float3 ve, float ax, float ay, float u0, float u1, float& pdf) { // :196-224
  if (!(ve.z >= 0.f)) ve = -ve;
  const float3 vh = norm3(mk3(ax * ve.x, ay * ve.y, ve.z));
  const float l2 = vh.x * vh.x + vh.y * vh.y;
  const float3 T1 = l2 > 0.f ? mk3(-vh.y, vh.x, 0.f) / sqrtf(l2) : mk3(1, 0, 0);
  const float3 T2 = cross3(vh, T1);
  const float r = sqrtf(u0);
  float sp, cp; sincos_2pi(u1, sp, cp);
  const float t1 = r * cp; float t2 = r * sp;
  const float s = 0.5f * (1.0f + vh.z);
  t2 = (1.0f - s) * sqrtf(1.0f - t1 * t1) + s * t2;
  const float3 nh = t1 * T1 + t2 * T2 + sqrtf(fmaxf(1.0f - t1 * t1 - t2 * t2, 0.f)) * vh;
  const float3 ne = norm3(mk3(ax * nh.x, ay * nh.y, fmaxf(nh.z, 0.f)));
  pdf = vndf_pdf(ne, ve, ax, ay);
  return ne;
}
SPT_D float3 cosine_hemisphere(float u0, float u1) { // Rng::cosine_weighted_on_hemisphere, core/rng.rs:72-80
  float sp, cp; sincos_2pi(u0, sp, cp);
  const float st = sqrtf(u1), ct = sqrtf(1.0f - u1);
  return mk3(st * cp, st * sp, ct);
}

// ---- Fresnel dispatch (bxdf/fresnel.rs) -------------------------------------------------------------------
template <int MAT> SPT_D float3 bsdf_fresnel(const Bsdf& b, float3 i, float3 n) {
  if (fr_of<MAT>(b) == FR_DIELECTRIC) return mk3(fresnel_diel(b.ior, i, n));
  if (fr_of<MAT>(b) == FR_CONDUCTOR) return fresnel_cond(b.c0, b.c1, i, n);
  return b.c0 + (mk3(1.f) - b.c0) * pow5f(1.0f - dot3(i, n)); // schlick_fresnel_with_r0 :120-122
}
template <int MAT> SPT_D float bsdf_fresnel_ior(const Bsdf& b) { // FresnelT::ior
  if (fr_of<MAT>(b) == FR_DIELECTRIC) return b.ior;
  if (fr_of<MAT>(b) == FR_CONDUCTOR) return 1.0f;
  const float s = sqrtf(lum3(b.c0));
  return (1.0f - s) / (1.0f + s);
}
SPT_D float ndf_vis(const Bsdf& b, float3 wo, float3 wi, float3 h) { return ggx_d(h, b.ax, b.ay) * smith_vis(wo, wi, b.ax, b.ay); } // microfacet.rs:48-53

// ---- substrates (bxdf/substrate.rs) -------------------------------------------------------------------------
SPT_D float sub_pdf(float3 wo, float3 wi) { return wo.z * wi.z >= 0.f ? fabsf(wi.z) * SPT_INV_PI : 1.0f; }
template <int MAT> SPT_D float3 sub_eval(const Bsdf& b, float3 wo, float3 wi) {
  if (!(wo.z * wi.z >= 0.f)) return mk3(0.f);
  if (sub_of<MAT>(b) == SUB_LAMBERT) return b.s_refl * SPT_INV_PI;
  return (1.0f - fresnel_diel(b.s_ior, wi, mk3(0, 0, 1))) * b.s_base; // Diffuse::bxdf :169-176 (Subsurface delegates, :345-347)
}
template <int MAT, class RNG> SPT_D void sub_sample_diffuse(const Bsdf& b, float3 wo, RNG& rng, float3& wi, float3& f, float& pdf) { // :141-160, lambert.rs:20-36
  const float u0 = rng.next(), u1 = rng.next();
  wi = cosine_hemisphere(u0, u1);
  if (wo.z < 0.f) wi.z = -wi.z;
  f = sub_of<MAT>(b) == SUB_LAMBERT ? b.s_refl * SPT_INV_PI : (1.0f - fresnel_diel(b.s_ior, wi, mk3(0, 0, 1))) * b.s_base;
  pdf = fabsf(wi.z) * SPT_INV_PI;
}
SPT_D float3 ss_sp(const Bsdf& b, float r) { // Subsurface::sp :211-215
  const float3 e1 = exp3(mk3(-r) / b.s_d), e2 = exp3(mk3(-r) / b.s_d * (1.0f / 3.0f));
  return (e1 + e2) * SPT_INV_PI / ((8.0f * b.s_d) * r);
}

template <int MAT> SPT_D bool bsdf_is_delta(const Bsdf& b) { const int t = tag_of<MAT>(b); return t == BX_PSEUDO || t == BX_SP_CONDUCTOR || t == BX_SP_DIELECTRIC; }

SPT_D float plastic_reflect_pdf(const Bsdf& b, float3 fres) {
  const float sw = lum3(fres), bw = lum3((mk3(1.f) - fres) * b.s_refl);
  return sw / (sw + bw);
}

// BxdfT::pdf
template <int MAT> SPT_D float bsdf_pdf(const Bsdf& b, float3 wo, float3 wi) {
  const float3 Z = mk3(0, 0, 1);
  const bool same = wo.z * wi.z >= 0.f;
  switch (tag_of<MAT>(b)) {
    case BX_LAMBERT: return same ? fabsf(wi.z) * SPT_INV_PI : 1.0f; // lambert.rs:38-44
    case BX_PSEUDO: case BX_SP_CONDUCTOR: return 1.0f;
    case BX_MF_CONDUCTOR: { // microfacet_conductor.rs:44-53
      if (!same) return 1.0f;
      const float3 h = half_refl(wo, wi);
      return vndf_pdf(h, wo, b.ax, b.ay) / (4.0f * fabsf(dot3(wo, h)));
    }
    case BX_MF_DIELECTRIC: { // microfacet_dielectric.rs:93-116
      if (same) {
        const float3 h = half_refl(wo, wi);
        return lum3(bsdf_fresnel<MAT>(b, wo, h)) * vndf_pdf(h, wo, b.ax, b.ay) / (4.0f * fabsf(dot3(wo, h)));
      }
      const float ior = bsdf_fresnel_ior<MAT>(b);
      const float3 h = half_refr(wo, wi, ior);
      const float rp = lum3(bsdf_fresnel<MAT>(b, wo, h));
      const float ratio = wo.z >= 0.f ? 1.0f / ior : ior;
      const float den = sqr(ratio * dot3(wo, h) + dot3(wi, h));
      return (1.0f - rp) * vndf_pdf(h, wo, b.ax, b.ay) * fabsf(dot3(wi, h)) / den;
    }
    case BX_MF_PLASTIC: { // microfacet_plastic.rs:81-98
      if (!same) return 1.0f;
      const float rp = plastic_reflect_pdf(b, bsdf_fresnel<MAT>(b, wo, Z));
      const float3 h = half_refl(wo, wi);
      return rp * vndf_pdf(h, wo, b.ax, b.ay) / (4.0f * fabsf(dot3(wo, h))) + (1.0f - rp) * sub_pdf(wo, wi);
    }
    case BX_SP_DIELECTRIC: { const float rp = lum3(bsdf_fresnel<MAT>(b, wo, Z)); return same ? rp : 1.0f - rp; } // specular_dielectric.rs:75-83
    default: { // BX_SP_PLASTIC, specular_plastic.rs:63-77
      if (!same) return 1.0f;
      const float rp = plastic_reflect_pdf(b, bsdf_fresnel<MAT>(b, wo, Z));
      return rp + (1.0f - rp) * sub_pdf(wo, wi);
    }
  }
}

// BxdfT::bxdf
template <int MAT> SPT_D float3 bsdf_eval(const Bsdf& b, float3 wo, float3 wi) {
  const float3 Z = mk3(0, 0, 1), K = mk3(0.f), W = mk3(1.f);
  const bool same = wo.z * wi.z >= 0.f;
  switch (tag_of<MAT>(b)) {
    case BX_LAMBERT: return same ? b.c0 * SPT_INV_PI : K;
    case BX_PSEUDO: return dot3(wo, wi) < -0.999f ? W / fabsf(wi.z) : K; // pseudo.rs:32-38
    case BX_MF_CONDUCTOR: {
      if (!same) return K;
      const float3 h = half_refl(wo, wi);
      return bsdf_fresnel<MAT>(b, wo, h) * ndf_vis(b, wo, wi, h);
    }
    case BX_MF_DIELECTRIC: { // :118-141
      if (same) { const float3 h = half_refl(wo, wi); return bsdf_fresnel<MAT>(b, wo, h) * ndf_vis(b, wo, wi, h); }
      const float ior = bsdf_fresnel_ior<MAT>(b);
      const float3 h = half_refr(wo, wi, ior);
      const float3 f = bsdf_fresnel<MAT>(b, wo, h);
      const float ratio = wo.z >= 0.f ? 1.0f / ior : ior;
      const float den = sqr(ratio * dot3(wo, h) + dot3(wi, h));
      const float num = 4.0f * fabsf(dot3(wo, h)) * fabsf(dot3(wi, h));
      return (W - f) * ndf_vis(b, wo, wi, h) * num / den;
    }
    case BX_MF_PLASTIC: { // :100-114
      if (!same) return K;
      const float3 h = half_refl(wo, wi);
      return bsdf_fresnel<MAT>(b, wo, h) * ndf_vis(b, wo, wi, h) + (W - bsdf_fresnel<MAT>(b, wo, Z)) * sub_eval<MAT>(b, wo, wi);
    }
    case BX_SP_CONDUCTOR: return dot3(wi, refl_z(wo)) > 0.999f ? bsdf_fresnel<MAT>(b, wo, Z) / fabsf(wi.z) : K; // :43-50
    case BX_SP_DIELECTRIC: { // :85-106
      const float3 f = bsdf_fresnel<MAT>(b, wo, Z);
      if (same) return dot3(wi, refl_z(wo)) > 0.999f ? f / fabsf(wi.z) : K;
      const float ior = bsdf_fresnel_ior<MAT>(b);
      float3 e;
      if (refr_z(wo, ior, e) && dot3(wi, e) > 0.999f) {
        const float ratio = wo.z >= 0.f ? 1.0f / ior : ior;
        return (ratio * ratio) * (W - f) / fabsf(wi.z);
      }
      return K;
    }
    default: { // BX_SP_PLASTIC :79-89
      if (!same) return K;
      const float3 f = bsdf_fresnel<MAT>(b, wo, Z);
      return f / fabsf(wi.z) + (W - f) * sub_eval<MAT>(b, wo, wi);
    }
  }
}

// Combination step shared by MicrofacetPlastic / SpecularPlastic when the substrate was sampled
// (microfacet_plastic.rs:59-78, specular_plastic.rs:46-60): adds the specular lobe's value and pdf at samp.wi.
template <int MAT> SPT_D void plastic_combine(const Bsdf& b, float3 wo, float3 wi, float3 fsub, float psub, int lobe, BsdfSample& s) {
  const float3 Z = mk3(0, 0, 1), W = mk3(1.f);
  const float3 fm = bsdf_fresnel<MAT>(b, wo, Z);
  const float rp = plastic_reflect_pdf(b, fm);
  s.wi = wi; s.lobe = lobe; s.dir = DIR_REFLECT;
  if (tag_of<MAT>(b) == BX_MF_PLASTIC) {
    const float3 h = half_refl(wo, wi);
    const float spdf = rp * vndf_pdf(h, wo, b.ax, b.ay) / (4.0f * fabsf(dot3(wo, h)));
    s.f = (W - fm) * fsub + bsdf_fresnel<MAT>(b, wo, h) * ndf_vis(b, wo, wi, h);
    s.pdf = (1.0f - rp) * psub + spdf;
  } else {
    s.f = (W - fm) * fsub + fm / fabsf(wi.z);
    s.pdf = (1.0f - rp) * psub + rp;
  }
}

// BxdfT::sample.  Random numbers are consumed in the reference's order.
template <int MAT, class RNG> SPT_D void bsdf_sample(const Bsdf& b, float3 wo, RNG& rng, BsdfSample& s) {
  const float3 Z = mk3(0, 0, 1), W = mk3(1.f);
  s.need_probe = false;
  switch (tag_of<MAT>(b)) {
    case BX_LAMBERT: { // lambert.rs:20-36
      const float u0 = rng.next(), u1 = rng.next();
      float3 wi = cosine_hemisphere(u0, u1);
      if (wo.z < 0.f) wi.z = -wi.z;
      s.wi = wi; s.f = b.c0 * SPT_INV_PI; s.pdf = fabsf(wi.z) * SPT_INV_PI; s.lobe = LOBE_DIFFUSE; s.dir = DIR_REFLECT;
      return;
    }
    case BX_PSEUDO: // pseudo.rs:14-26
      s.wi = -wo; s.f = W / fabsf(wo.z); s.pdf = 1.0f; s.lobe = LOBE_SPECULAR; s.dir = DIR_TRANSMIT;
      return;
    case BX_MF_CONDUCTOR: { // microfacet_conductor.rs:23-42
      const float u0 = rng.next(), u1 = rng.next();
      float hp; const float3 h = vndf_sample(wo, b.ax, b.ay, u0, u1, hp);
      const float3 wi = refl_n(wo, h);
      s.wi = wi; s.f = bsdf_fresnel<MAT>(b, wo, h) * ndf_vis(b, wo, wi, h); s.pdf = hp / (4.0f * fabsf(dot3(wo, h)));
      s.lobe = LOBE_GLOSSY; s.dir = DIR_REFLECT;
      return;
    }
    case BX_MF_DIELECTRIC: { // microfacet_dielectric.rs:23-91
      const float u0 = rng.next(), u1 = rng.next();
      float hp; const float3 h = vndf_sample(wo, b.ax, b.ay, u0, u1, hp);
      const float3 f = bsdf_fresnel<MAT>(b, wo, h);
      const float rp = lum3(f);
      const float ior = bsdf_fresnel_ior<MAT>(b);
      float3 wi;
      s.lobe = LOBE_GLOSSY;
      if (rng.next() < rp) {
        wi = refl_n(wo, h);
        s.wi = wi; s.f = f * ndf_vis(b, wo, wi, h); s.pdf = rp * hp / (4.0f * fabsf(dot3(wo, h))); s.dir = DIR_REFLECT;
      } else if (refr_n(wo, h, ior, wi)) {
        const float ratio = wo.z >= 0.f ? 1.0f / ior : ior;
        const float den = sqr(ratio * dot3(wo, h) + dot3(wi, h));
        s.pdf = (1.0f - rp) * hp * fabsf(dot3(wi, h)) / den;
        s.f = (W - f) * ndf_vis(b, wo, wi, h) * (4.0f * fabsf(dot3(wo, h)) * fabsf(dot3(wi, h))) / den;
        s.wi = wi; s.dir = DIR_TRANSMIT;
      } else { s.wi = mk3(0.f); s.f = mk3(0.f); s.pdf = 1.0f; s.dir = DIR_TRANSMIT; }
      return;
    }
    case BX_SP_CONDUCTOR: { // specular_conductor.rs:19-37
      const float3 wi = refl_z(wo);
      s.wi = wi; s.f = bsdf_fresnel<MAT>(b, wo, Z) / fabsf(wi.z); s.pdf = 1.0f; s.lobe = LOBE_SPECULAR; s.dir = DIR_REFLECT;
      return;
    }
    case BX_SP_DIELECTRIC: { // specular_dielectric.rs:19-73
      const float3 f = bsdf_fresnel<MAT>(b, wo, Z);
      const float rp = lum3(f);
      const float ior = bsdf_fresnel_ior<MAT>(b);
      float3 wi;
      s.lobe = LOBE_SPECULAR;
      if (rng.next() < rp) { wi = refl_z(wo); s.wi = wi; s.f = f / fabsf(wi.z); s.pdf = rp; s.dir = DIR_REFLECT; }
      else if (refr_z(wo, ior, wi)) {
        const float ratio = wo.z >= 0.f ? 1.0f / ior : ior;
        s.wi = wi; s.f = (ratio * ratio) * (W - f) / fabsf(wi.z); s.pdf = 1.0f - rp; s.dir = DIR_TRANSMIT;
      } else { s.wi = mk3(0.f); s.f = mk3(0.f); s.pdf = 1.0f; s.dir = DIR_TRANSMIT; }
      return;
    }
    default: { // BX_MF_PLASTIC (microfacet_plastic.rs:25-79) and BX_SP_PLASTIC (specular_plastic.rs:20-61)
      const float3 fm = bsdf_fresnel<MAT>(b, wo, Z);
      const float rp = plastic_reflect_pdf(b, fm);
      if (rng.next() < rp) {
        if (tag_of<MAT>(b) == BX_MF_PLASTIC) {
          const float u0 = rng.next(), u1 = rng.next();
          float hp; const float3 h = vndf_sample(wo, b.ax, b.ay, u0, u1, hp);
          const float3 wi = refl_n(wo, h);
          s.wi = wi; s.lobe = LOBE_GLOSSY; s.dir = DIR_REFLECT;
          s.f = bsdf_fresnel<MAT>(b, wo, h) * ndf_vis(b, wo, wi, h) + (W - fm) * sub_eval<MAT>(b, wo, wi);
          s.pdf = rp * hp / (4.0f * fabsf(dot3(wo, h))) + (1.0f - rp) * sub_pdf(wo, wi);
        } else {
          const float3 wi = refl_z(wo);
          s.wi = wi; s.lobe = LOBE_SPECULAR; s.dir = DIR_REFLECT;
          s.f = fm / fabsf(wi.z) + (W - fm) * sub_eval<MAT>(b, wo, wi);
          s.pdf = rp + (1.0f - rp) * sub_pdf(wo, wi);
        }
        return;
      }
      if (sub_of<MAT>(b) == SUB_SUBSURFACE) { s.need_probe = true; s.wi = mk3(0.f); s.f = mk3(0.f); s.pdf = 1.0f; s.lobe = LOBE_DIFFUSE; s.dir = DIR_REFLECT; return; }
      float3 wi, fs; float ps;
      sub_sample_diffuse<MAT>(b, wo, rng, wi, fs, ps);
      plastic_combine<MAT>(b, wo, wi, fs, ps, LOBE_DIFFUSE, s);
      return;
    }
  }
}

// ---- MaterialT::bxdf_context (src/material/*.rs) ------------------------------------------------------------
SPT_D void bsdf_set_rough(Bsdf& b, float rx, float ry, int rough_tag, int spec_tag) {
  if (rx < 0.0001f || ry < 0.0001f) b.tag = spec_tag; else { b.tag = rough_tag; b.ax = rx; b.ay = ry; }
}
SPT_D void bsdf_set_diffuse_substrate(Bsdf& b, float3 refl, float ior) { // Diffuse::new, substrate.rs:127-137
  b.sub = SUB_DIFFUSE; b.s_refl = refl; b.s_ior = ior;
  const float fdr = 2.0f * fresnel_moment1(1.0f / ior);
  b.s_base = refl * SPT_INV_PI / ((mk3(1.f) - refl * fdr) * ior * ior);
}
SPT_D float powi4(float x) { const float x2 = x * x; return x2 * x2; }

template <int MAT> SPT_D Bsdf make_bsdf(const DScene& sc, const DMaterial& m, const TexIn& in) {
  Bsdf b;
  b.tag = BX_LAMBERT; b.fr = FR_DIELECTRIC; b.sub = SUB_LAMBERT; b.ax = b.ay = 0.f; b.ior = m.ior;
  b.c0 = b.c1 = b.s_refl = b.s_base = b.s_d = mk3(0.f); b.s_ior = 1.f;
  switch (MAT >= 0 ? MAT : m.tag) {
    case 0: b.c0 = tex_color(sc, m.tex[0], in); break; // lambert.rs:26-31
    case 1: b.tag = BX_PSEUDO; break;                   // pseudo.rs:20-24
    case 2: { // conductor.rs:50-73
      b.fr = FR_CONDUCTOR; b.c0 = tex_color(sc, m.tex[0], in); b.c1 = tex_color(sc, m.tex[1], in);
      bsdf_set_rough(b, sqr(tex_float(sc, m.tex[2], in, 0)), sqr(tex_float(sc, m.tex[3], in, 0)), BX_MF_CONDUCTOR, BX_SP_CONDUCTOR);
      break;
    }
    case 3: // dielectric.rs:65-89 (Q11: reflectance / transmittance textures have no effect)
      bsdf_set_rough(b, sqr(tex_float(sc, m.tex[2], in, 0)), sqr(tex_float(sc, m.tex[3], in, 0)), BX_MF_DIELECTRIC, BX_SP_DIELECTRIC);
      break;
    case 4: // plastic.rs:60-82 (Q12: roughness NOT squared)
      bsdf_set_diffuse_substrate(b, tex_color(sc, m.tex[0], in), m.ior);
      bsdf_set_rough(b, tex_float(sc, m.tex[2], in, 0), tex_float(sc, m.tex[3], in, 0), BX_MF_PLASTIC, BX_SP_PLASTIC);
      break;
    case 5: { // pbr_metallic.rs:67-98
      const float3 base = tex_color(sc, m.tex[0], in);
      const float rx = sqr(tex_float(sc, m.tex[2], in, m.rough_chan)), ry = sqr(tex_float(sc, m.tex[3], in, m.rough_chan));
      const float metallic = tex_float(sc, m.tex[1], in, m.metal_chan);
      b.fr = FR_SCHLICK; b.c0 = metallic * base + (1.0f - metallic) * mk3(0.04f);
      b.sub = SUB_LAMBERT; b.s_refl = base * (1.0f - metallic);
      bsdf_set_rough(b, rx, ry, BX_MF_PLASTIC, BX_SP_PLASTIC);
      break;
    }
    case 6: { // pbr_specular.rs:60-88
      b.sub = SUB_LAMBERT; b.s_refl = tex_color(sc, m.tex[0], in);
      b.fr = FR_SCHLICK; b.c0 = tex_color(sc, m.tex[1], in);
      bsdf_set_rough(b, sqr(tex_float(sc, m.tex[2], in, m.rough_chan)), sqr(tex_float(sc, m.tex[3], in, m.rough_chan)), BX_MF_PLASTIC, BX_SP_PLASTIC);
      break;
    }
    default: { // subsurface.rs:65-94 + Subsurface::new substrate.rs:199-209
      const float3 albedo = tex_color(sc, m.tex[0], in);
      const float ld = tex_float(sc, m.tex[1], in, 0);
      bsdf_set_diffuse_substrate(b, albedo, m.ior);
      b.sub = SUB_SUBSURFACE;
      b.s_d = mk3(ld / (3.5f + 100.0f * powi4(albedo.x - 0.33f)), ld / (3.5f + 100.0f * powi4(albedo.y - 0.33f)),
                  ld / (3.5f + 100.0f * powi4(albedo.z - 0.33f)));
      bsdf_set_rough(b, sqr(tex_float(sc, m.tex[2], in, 0)), sqr(tex_float(sc, m.tex[3], in, 0)), BX_MF_PLASTIC, BX_SP_PLASTIC);
      break;
    }
  }
  return b;
}
