// oracle/bxdf2.h — TEST INFRASTRUCTURE ONLY.  The `Bxdf` enum (bxdf/mod.rs:92-102) and its 8 BxdfT impls.
// PARITY UNPINNED (see vec.h).
#pragma once
#include "bxdf.h"

namespace orc {

struct Bxdf {
  // order of bxdf/mod.rs:92-102
  enum Tag { kLambert, kPseudo, kMicrofacetConductor, kMicrofacetDielectric, kMicrofacetPlastic,
             kSpecularConductor, kSpecularDielectric, kSpecularPlastic } tag = kLambert;
  Color reflectance;      // Lambert
  float ax = 0, ay = 0;   // GgxMicrofacet roughness_x / roughness_y (microfacet.rs:20-23)
  Fresnel fr;
  Substrate sub;

  // GgxMicrofacet (microfacet.rs:35-54)
  Vec3 sample_half(const BxdfInputs& in, Rng& rng, float& pdf) const { float a, b; rng.uniform_2d(a, b); return ggx_smith_vndf_sample(in.wo, ax, ay, a, b, pdf); }
  float half_pdf(Vec3 wo, Vec3 h) const { return ggx_smith_vndf_pdf(h, wo, ax, ay); }
  float ndf_visible(Vec3 wo, Vec3 wi, Vec3 h) const { return ggx_ndf_aniso(h, ax, ay) * smith_separable_visible_aniso(wo, wi, ax, ay); }

  bool is_delta() const { return tag == kPseudo || tag == kSpecularConductor || tag == kSpecularDielectric; }

  float plastic_reflect_pdf(Color fres) const { // shared by *_plastic.rs
    float sw = fres.luminance();
    float bw = ((WHITE - fres) * sub.reflectance()).luminance();
    return sw / (sw + bw);
  }

  BxdfSample sample(const BxdfInputs& in, Rng& rng) const {
    const Vec3 Z{0, 0, 1};
    BxdfSample s;
    switch (tag) {
      case kLambert: { // lambert.rs:20-36
        Vec3 wi = rng.cosine_weighted_on_hemisphere();
        if (in.wo.z < 0.0f) wi.z = -wi.z;
        s.wi = wi; s.lobe = kDiffuse; s.dir = kReflect; s.bxdf = reflectance * FRAC_1_PI; s.pdf = std::fabs(wi.z) * FRAC_1_PI;
        return s;
      }
      case kPseudo: // pseudo.rs:14-26
        s.wi = -in.wo; s.lobe = kSpecular; s.dir = kTransmit; s.bxdf = WHITE / std::fabs(in.wo.z); s.pdf = 1.0f;
        return s;
      case kMicrofacetConductor: { // microfacet_conductor.rs:23-42
        float hp; Vec3 h = sample_half(in, rng, hp);
        Color f = fr.fresnel(in.wo, h);
        Vec3 wi = reflect_n(in.wo, h);
        s.wi = wi; s.lobe = kGlossy; s.dir = kReflect;
        s.bxdf = f * ndf_visible(in.wo, wi, h);
        s.pdf = hp / (4.0f * std::fabs(dot(in.wo, h)));
        return s;
      }
      case kMicrofacetDielectric: { // microfacet_dielectric.rs:23-91
        float hp; Vec3 h = sample_half(in, rng, hp);
        Color f = fr.fresnel(in.wo, h);
        float rp = f.luminance();
        Vec3 wi;
        if (rng.uniform_1d() < rp) {
          wi = reflect_n(in.wo, h);
          s.wi = wi; s.lobe = kGlossy; s.dir = kReflect;
          s.bxdf = f * ndf_visible(in.wo, wi, h);
          s.pdf = rp * hp / (4.0f * std::fabs(dot(in.wo, h)));
        } else if (refract_n(in.wo, h, fr.ior(), wi)) {
          float ratio = in.wo.z >= 0.0f ? 1.0f / fr.ior() : fr.ior();
          float denom = ratio * dot(in.wo, h) + dot(wi, h); denom = denom * denom;
          float num = std::fabs(dot(wi, h));
          s.pdf = (1.0f - rp) * hp * num / denom;
          num = 4.0f * std::fabs(dot(in.wo, h)) * std::fabs(dot(wi, h));
          s.bxdf = (WHITE - f) * ndf_visible(in.wo, wi, h) * num / denom;
          s.wi = wi; s.lobe = kGlossy; s.dir = kTransmit;
        } else { s.wi = {0, 0, 0}; s.lobe = kGlossy; s.dir = kTransmit; s.bxdf = BLACK; s.pdf = 1.0f; }
        return s;
      }
      case kMicrofacetPlastic: { // microfacet_plastic.rs:25-79
        Color fm = fr.fresnel(in.wo, Z);
        float rp = plastic_reflect_pdf(fm);
        if (rng.uniform_1d() < rp) {
          float hp; Vec3 h = sample_half(in, rng, hp);
          Color f = fr.fresnel(in.wo, h);
          Vec3 wi = reflect_n(in.wo, h);
          Color sb = f * ndf_visible(in.wo, wi, h);
          float sp = rp * hp / (4.0f * std::fabs(dot(in.wo, h)));
          Color ub = (WHITE - fm) * sub.bxdf(in.wo, wi);
          float up = (1.0f - rp) * sub.pdf(in.wo, wi);
          s.wi = wi; s.lobe = kGlossy; s.dir = kReflect; s.bxdf = sb + ub; s.pdf = sp + up;
          return s;
        }
        BxdfSample samp = sub.sample(in, rng);
        float up = (1.0f - rp) * samp.pdf;
        Color ub = (WHITE - fm) * samp.bxdf;
        Vec3 h = half_from_reflect(in.wo, samp.wi);
        float hp = half_pdf(in.wo, h);
        float sp = rp * hp / (4.0f * std::fabs(dot(in.wo, h)));
        Color f = fr.fresnel(in.wo, h);
        Color sb = f * ndf_visible(in.wo, samp.wi, h);
        samp.bxdf = ub + sb; samp.pdf = up + sp;
        return samp;
      }
      case kSpecularConductor: { // specular_conductor.rs:19-37
        Color f = fr.fresnel(in.wo, Z);
        Vec3 wi = reflect(in.wo);
        s.wi = wi; s.lobe = kSpecular; s.dir = kReflect; s.bxdf = f / std::fabs(wi.z); s.pdf = 1.0f;
        return s;
      }
      case kSpecularDielectric: { // specular_dielectric.rs:19-73
        Color f = fr.fresnel(in.wo, Z);
        float rp = f.luminance();
        Vec3 wi;
        if (rng.uniform_1d() < rp) {
          wi = reflect(in.wo);
          s.wi = wi; s.lobe = kSpecular; s.dir = kReflect; s.bxdf = f / std::fabs(wi.z); s.pdf = rp;
        } else if (refract(in.wo, fr.ior(), wi)) {
          float ratio = in.wo.z >= 0.0f ? 1.0f / fr.ior() : fr.ior();
          s.wi = wi; s.lobe = kSpecular; s.dir = kTransmit;
          s.bxdf = ratio * ratio * (WHITE - f) / std::fabs(wi.z); s.pdf = 1.0f - rp;
        } else { s.wi = {0, 0, 0}; s.lobe = kSpecular; s.dir = kTransmit; s.bxdf = BLACK; s.pdf = 1.0f; }
        return s;
      }
      default: { // kSpecularPlastic, specular_plastic.rs:20-61
        Color f = fr.fresnel(in.wo, Z);
        float rp = plastic_reflect_pdf(f);
        if (rng.uniform_1d() < rp) {
          Vec3 wi = reflect(in.wo);
          Color sb = f / std::fabs(wi.z);
          Color ub = (WHITE - f) * sub.bxdf(in.wo, wi);
          float up = (1.0f - rp) * sub.pdf(in.wo, wi);
          s.wi = wi; s.lobe = kSpecular; s.dir = kReflect; s.bxdf = sb + ub; s.pdf = rp + up;
          return s;
        }
        BxdfSample samp = sub.sample(in, rng);
        float up = (1.0f - rp) * samp.pdf;
        Color ub = (WHITE - f) * samp.bxdf;
        Color sb = f / std::fabs(samp.wi.z);
        samp.bxdf = ub + sb; samp.pdf = up + rp;
        return samp;
      }
    }
  }

  float pdf(Vec3 wo, Vec3 wi) const {
    const Vec3 Z{0, 0, 1};
    bool same = wo.z * wi.z >= 0.0f;
    switch (tag) {
      case kLambert: return same ? std::fabs(wi.z) * FRAC_1_PI : 1.0f; // lambert.rs:38-44 (Q10)
      case kPseudo: return 1.0f;
      case kMicrofacetConductor: // :44-53
        if (same) { Vec3 h = half_from_reflect(wo, wi); return half_pdf(wo, h) / (4.0f * std::fabs(dot(wo, h))); }
        return 1.0f;
      case kMicrofacetDielectric: { // :93-116
        if (same) {
          Vec3 h = half_from_reflect(wo, wi);
          float hp = half_pdf(wo, h);
          float rp = fr.fresnel(wo, h).luminance();
          return rp * hp / (4.0f * std::fabs(dot(wo, h)));
        }
        Vec3 h = half_from_refract(wo, wi, fr.ior());
        float hp = half_pdf(wo, h);
        float rp = fr.fresnel(wo, h).luminance();
        float ratio = wo.z >= 0.0f ? 1.0f / fr.ior() : fr.ior();
        float denom = ratio * dot(wo, h) + dot(wi, h); denom = denom * denom;
        float num = std::fabs(dot(wi, h));
        return (1.0f - rp) * hp * num / denom;
      }
      case kMicrofacetPlastic: { // :81-98
        if (!same) return 1.0f;
        float rp = plastic_reflect_pdf(fr.fresnel(wo, Z));
        Vec3 h = half_from_reflect(wo, wi);
        float sp = rp * half_pdf(wo, h) / (4.0f * std::fabs(dot(wo, h)));
        return sp + (1.0f - rp) * sub.pdf(wo, wi);
      }
      case kSpecularConductor: return 1.0f;
      case kSpecularDielectric: { float rp = fr.fresnel(wo, Z).luminance(); return same ? rp : 1.0f - rp; } // :75-83
      default: { // :63-77
        if (!same) return 1.0f;
        float rp = plastic_reflect_pdf(fr.fresnel(wo, Z));
        return rp + (1.0f - rp) * sub.pdf(wo, wi);
      }
    }
  }

  Color bxdf(Vec3 wo, Vec3 wi) const {
    const Vec3 Z{0, 0, 1};
    bool same = wo.z * wi.z >= 0.0f;
    switch (tag) {
      case kLambert: return same ? reflectance * FRAC_1_PI : BLACK;
      case kPseudo: return dot(wo, wi) < -0.999f ? WHITE / std::fabs(wi.z) : BLACK; // pseudo.rs:32-38
      case kMicrofacetConductor:
        if (same) { Vec3 h = half_from_reflect(wo, wi); return fr.fresnel(wo, h) * ndf_visible(wo, wi, h); }
        return BLACK;
      case kMicrofacetDielectric: { // :118-141
        if (same) { Vec3 h = half_from_reflect(wo, wi); return fr.fresnel(wo, h) * ndf_visible(wo, wi, h); }
        Vec3 h = half_from_refract(wo, wi, fr.ior());
        Color f = fr.fresnel(wo, h);
        float ratio = wo.z >= 0.0f ? 1.0f / fr.ior() : fr.ior();
        float denom = ratio * dot(wo, h) + dot(wi, h); denom = denom * denom;
        float num = 4.0f * std::fabs(dot(wo, h)) * std::fabs(dot(wi, h));
        return (WHITE - f) * ndf_visible(wo, wi, h) * num / denom;
      }
      case kMicrofacetPlastic: { // :100-114
        if (!same) return BLACK;
        Vec3 h = half_from_reflect(wo, wi);
        Color refl = fr.fresnel(wo, h) * ndf_visible(wo, wi, h);
        Color subs = (WHITE - fr.fresnel(wo, Z)) * sub.bxdf(wo, wi);
        return refl + subs;
      }
      case kSpecularConductor: { // :43-50
        if (dot(wi, reflect(wo)) > 0.999f) return fr.fresnel(wo, Z) / std::fabs(wi.z);
        return BLACK;
      }
      case kSpecularDielectric: { // :85-106
        Color f = fr.fresnel(wo, Z);
        Vec3 e;
        if (same) return dot(wi, reflect(wo)) > 0.999f ? f / std::fabs(wi.z) : BLACK;
        if (refract(wo, fr.ior(), e)) {
          if (dot(wi, e) > 0.999f) {
            float ratio = wo.z >= 0.0f ? 1.0f / fr.ior() : fr.ior();
            return ratio * ratio * (WHITE - f) / std::fabs(wi.z);
          }
          return BLACK;
        }
        return BLACK;
      }
      default: { // :79-89
        if (!same) return BLACK;
        Color f = fr.fresnel(wo, Z);
        return f / std::fabs(wi.z) + (WHITE - f) * sub.bxdf(wo, wi);
      }
    }
  }
};

} // namespace orc
