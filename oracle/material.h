// oracle/material.h — TEST INFRASTRUCTURE ONLY.  Restates material/*.rs (in-scope variants), core/surface.rs,
// medium/{homogeneous,util}.rs and core/alias_table.rs.  PARITY UNPINNED (see vec.h).
#pragma once
#include "bxdf2.h"

namespace orc {

struct Material { // tags follow include/spt.h SPT_MAT_*
  int tag = 0; int tex[4] = {-1, -1, -1, -1}; float ior = 1; int rough_chan = 0, metal_chan = 0;

  Bxdf bxdf_context(const TextureTable& T, const Intersection& inter) const {
    TextureInput in = TextureInput::from(inter);
    Bxdf b;
    auto ggx_or_spec = [&](float rx, float ry, Bxdf::Tag rough, Bxdf::Tag spec) {
      if (rx < 0.0001f || ry < 0.0001f) b.tag = spec; else { b.tag = rough; b.ax = rx; b.ay = ry; }
    };
    switch (tag) {
      case 0: b.tag = Bxdf::kLambert; b.reflectance = T.color_at(tex[0], in); return b; // lambert.rs:26-31
      case 1: b.tag = Bxdf::kPseudo; return b;                                           // pseudo.rs:20-24
      case 2: { // conductor.rs:50-73
        b.fr.kind = Fresnel::kConductor;
        b.fr.eta = T.color_at(tex[0], in); b.fr.k = T.color_at(tex[1], in);
        float rx = pow2(T.float_at(tex[2], in, 0)), ry = pow2(T.float_at(tex[3], in, 0));
        ggx_or_spec(rx, ry, Bxdf::kMicrofacetConductor, Bxdf::kSpecularConductor);
        return b;
      }
      case 3: { // dielectric.rs:65-89 (Q11: reflectance/transmittance evaluated then ignored)
        b.fr.kind = Fresnel::kDielectric; b.fr.ior_v = ior;
        float rx = pow2(T.float_at(tex[2], in, 0)), ry = pow2(T.float_at(tex[3], in, 0));
        ggx_or_spec(rx, ry, Bxdf::kMicrofacetDielectric, Bxdf::kSpecularDielectric);
        return b;
      }
      case 4: { // plastic.rs:60-82 (Q12: roughness not squared)
        Color albedo = T.color_at(tex[0], in);
        float rx = T.float_at(tex[2], in, 0), ry = T.float_at(tex[3], in, 0);
        b.fr.kind = Fresnel::kDielectric; b.fr.ior_v = ior;
        b.sub = Substrate::diffuse(albedo, ior);
        ggx_or_spec(rx, ry, Bxdf::kMicrofacetPlastic, Bxdf::kSpecularPlastic);
        return b;
      }
      case 5: { // pbr_metallic.rs:67-98
        Color base = T.color_at(tex[0], in);
        float rx = pow2(T.float_at(tex[2], in, rough_chan)), ry = pow2(T.float_at(tex[3], in, rough_chan));
        float metallic = T.float_at(tex[1], in, metal_chan);
        Color specular = metallic * base + (1.0f - metallic) * Color::gray(0.04f);
        Color diffuse = base * (1.0f - metallic);
        b.fr.kind = Fresnel::kSchlick; b.fr.r0 = specular;
        b.sub = Substrate::lambert(diffuse);
        ggx_or_spec(rx, ry, Bxdf::kMicrofacetPlastic, Bxdf::kSpecularPlastic);
        return b;
      }
      case 6: { // pbr_specular.rs:60-88
        Color diffuse = T.color_at(tex[0], in), specular = T.color_at(tex[1], in);
        float rx = pow2(T.float_at(tex[2], in, rough_chan)), ry = pow2(T.float_at(tex[3], in, rough_chan));
        b.fr.kind = Fresnel::kSchlick; b.fr.r0 = specular;
        b.sub = Substrate::lambert(diffuse);
        ggx_or_spec(rx, ry, Bxdf::kMicrofacetPlastic, Bxdf::kSpecularPlastic);
        return b;
      }
      default: { // subsurface.rs:65-94
        Color albedo = T.color_at(tex[0], in);
        float ld = T.float_at(tex[1], in, 0);
        float rx = pow2(T.float_at(tex[2], in, 0)), ry = pow2(T.float_at(tex[3], in, 0));
        b.fr.kind = Fresnel::kDielectric; b.fr.ior_v = ior;
        b.sub = Substrate::subsurface(albedo, ior, ld);
        ggx_or_spec(rx, ry, Bxdf::kMicrofacetPlastic, Bxdf::kSpecularPlastic);
        return b;
      }
    }
  }
};

// ---- medium/util.rs, medium/homogeneous.rs ---------------------------------------------------------
inline float henyey_greenstein(float g, float cs) { // util.rs:1-7
  float g2 = g * g;
  float denom = 1.0f + g2 + 2.0f * g * cs;
  denom = denom * std::sqrt(denom);
  return 0.25f * FRAC_1_PI * (1.0f - g2) / denom;
}
inline float henyey_greenstein_cdf_inverse(float g, float rand) { // util.rs:10-19
  if (std::fabs(g) < 0.01f) return 1.0f - 2.0f * rand;
  float g2 = g * g;
  float temp = (1.0f - g2) / (1.0f - g + 2.0f * g * rand);
  return 0.5f * (1.0f + g2 - temp * temp) / g;
}
inline Vec3 medium_local_to_world(Vec3 wo_world, Vec3 wi_local) { // util.rs:21-31
  Vec3 v = std::fabs(wo_world.y) < 0.99f ? Vec3(0, 1, 0) : Vec3(1, 0, 0);
  Vec3 u = normalize(cross(v, wo_world));
  v = cross(wo_world, u);
  return Mat3{u, v, wo_world} * wi_local;
}
struct Medium { // homogeneous.rs:5-9
  Color sigma_t, sigma_s; float g = 0;
  void init(Color sa, Color ss, float g_) { sigma_t = sa + ss; sigma_s = ss; g = g_; } // :12-19
  Vec3 sample_pi(Vec3 po, Vec3 wo, float t_max, Rng& rng, bool& still_in, Color& atten_out) const { // :31-60
    float rx, ry; rng.uniform_2d(rx, ry);
    float sst = rx < 1.0f / 3.0f ? sigma_t.r : (rx < 2.0f / 3.0f ? sigma_t.g : sigma_t.b);
    float sample_t = -std::log(1.0f - ry) / sst;
    float tm = fmin_(sample_t, t_max);
    Color attenuation = (-sigma_t * tm).exp();
    Vec3 pi = po - wo * tm;
    if (sample_t < t_max) {
      float ap = (sigma_t * attenuation).avg();
      still_in = true; atten_out = attenuation * sigma_s / ap;
    } else {
      float ap = attenuation.avg();
      still_in = false; atten_out = attenuation / ap;
    }
    return pi;
  }
  Vec3 sample_wi(Vec3 wo, Rng& rng, float& pdf) const { // :62-71
    float rx, ry; rng.uniform_2d(rx, ry);
    float ct = henyey_greenstein_cdf_inverse(g, rx);
    float st = std::sqrt(1.0f - ct * ct);
    float phi = 2.0f * PI * ry;
    Vec3 wi{st * std::cos(phi), st * std::sin(phi), ct};
    pdf = henyey_greenstein(g, ct);
    return medium_local_to_world(wo, wi);
  }
  Color transport_attenuation(float dist) const { return (-sigma_t * dist).exp(); } // :73-75
  float phase(Vec3 wo, Vec3 wi) const { return henyey_greenstein(g, dot(wo, wi)); } // :77-80
};

struct Surface { // surface.rs:14-22
  const Material* material = nullptr;
  int normal_map = -1; Color emissive_v; int emissive_map = -1; bool double_sided = false; const Medium* inside_medium_v = nullptr;
  const TextureTable* T = nullptr;
  bool is_emissive() const { return emissive_v.luminance() > 0.0f; } // :45-47
  Color emissive(const Intersection& inter) const {                 // :49-55
    return emissive_v * (emissive_map >= 0 ? T->color_at(emissive_map, TextureInput::from(inter)) : WHITE);
  }
  Color average_emissive() const { return emissive_v * (emissive_map >= 0 ? T->average_color(emissive_map) : WHITE); } // :57-63
  Coordinate coord(const Ray& ray, const Intersection& inter) const { // :65-95
    Vec3 shade_normal;
    if (normal_map >= 0) {
      Color value = T->color_at(normal_map, TextureInput::from(inter));
      Color nc = value * 2.0f - WHITE;
      Vec3 nl = normalize(Vec3(nc.r, nc.g, nc.b));
      shade_normal = normalize(nl.x * normalize(inter.tangent) + nl.y * normalize(inter.bitangent) + nl.z * inter.normal);
    } else shade_normal = inter.normal;
    bool hit_back = dot(ray.direction, inter.normal) > 0.0f;
    return Coordinate::from_tangent_normal(inter.tangent, (double_sided && hit_back) ? -shade_normal : shade_normal,
                                           hit_back ? -inter.normal : inter.normal);
  }
  const Medium* inside_medium() const { return double_sided ? nullptr : inside_medium_v; } // :105-111
};
inline Coordinate surface_coord(const Surface* s, const Ray& ray, const Intersection& inter) { return s->coord(ray, inter); }

// ---- core/alias_table.rs -----------------------------------------------------------------------------
struct AliasTable {
  std::vector<float> props, u; std::vector<size_t> k;
  void build(std::vector<float> p) { // AliasTable::new :8-56
    props = std::move(p);
    size_t n = props.size();
    u.resize(n); k.resize(n);
    for (size_t i = 0; i < n; i++) { u[i] = props[i] * (float)n; k[i] = i; }
    long poor = -1, rich = -1;
    for (size_t i = 0; i < n; i++) if (u[i] < 1.0f) { poor = (long)i; break; }
    long poor_max = poor;
    for (size_t i = 0; i < n; i++) if (u[i] > 1.0f) { rich = (long)i; break; }
    while (poor >= 0 && rich >= 0) {
      float diff = 1.0f - u[poor];
      u[rich] -= diff;
      k[poor] = (size_t)rich;
      if (u[rich] < 1.0f && rich < poor_max) poor = rich;
      else {
        poor = -1;
        for (size_t i = (size_t)poor_max + 1; i < n; i++) if (u[i] < 1.0f) { poor = (long)i; poor_max = (long)i; break; }
      }
      long r0 = rich; rich = -1;
      for (size_t i = (size_t)r0; i < n; i++) if (u[i] > 1.0f) { rich = (long)i; break; }
    }
  }
  size_t sample(float rand, float& prob) const { // :58-67
    float temp = rand * (float)props.size();
    size_t x = (size_t)temp;
    float y = temp - (float)x;
    if (y < u[x]) { prob = props[x]; return x; }
    prob = props[k[x]]; return k[x];
  }
  float probability(size_t i) const { return props[i]; }
};

} // namespace orc
