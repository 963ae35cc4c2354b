// oracle/lights.h — TEST INFRASTRUCTURE ONLY.  Restates light/*.rs and light_sampler/*.rs.
// PARITY UNPINNED (see vec.h).
#pragma once
#include "material.h"

namespace orc {

struct EnvLight { // environment.rs:9-16
  std::vector<Color> texture; // row-major [y][x]
  Color scale; size_t height = 0, width = 0; AliasTable alias; float avg_power = 0;
  void init(std::vector<Color> tex, size_t w, size_t h, Color sc) { // EnvLight::new :18-50 (Q7: theta lacks pi)
    texture = std::move(tex); width = w; height = h; scale = sc;
    std::vector<float> props; props.reserve(w * h);
    float sum = 0.0f;
    float height_inv = 1.0f / (float)h;
    for (size_t row = 0; row < h; row++) for (size_t x = 0; x < w; x++) {
      float theta = ((float)row + 0.5f) * height_inv;
      float prop = texture[row * w + x].luminance() * std::sin(theta);
      sum += prop; props.push_back(prop);
    }
    float sum_inv = 1.0f / sum;
    for (auto& p : props) p *= sum_inv;
    avg_power = sum / (float)props.size();
    alias.build(std::move(props));
  }
  void lookup(float theta, float phi, Color& c_out, float& dist, float& pdf) const { // :52-84 (Q7: product pdf)
    int W = (int)width, H = (int)height;
    float x = phi * 0.5f * FRAC_1_PI * (float)width;
    int x1 = TextureInput::as_i32(std::round(x)), x0 = x1 - 1;
    float xt = x - (float)x0 - 0.5f;
    x0 = std::clamp(x0, 0, W - 1); x1 = std::clamp(x1, 0, W - 1);
    float y = theta * FRAC_1_PI * (float)height;
    int y1 = TextureInput::as_i32(std::round(y)), y0 = y1 - 1;
    float yt = y - (float)y0 - 0.5f;
    y0 = std::clamp(y0, 0, H - 1); y1 = std::clamp(y1, 0, H - 1);
    Color c00 = texture[y0 * width + x0], c01 = texture[y1 * width + x0], c10 = texture[y0 * width + x1], c11 = texture[y1 * width + x1];
    Color c0 = c00 * (1.0f - yt) + c01 * yt, c1 = c10 * (1.0f - yt) + c11 * yt;
    Color c = c0 * (1.0f - xt) + c1 * xt;
    float p00 = alias.probability(y0 * width + x0), p01 = alias.probability(y1 * width + x0);
    float p10 = alias.probability(y0 * width + x1), p11 = alias.probability(y1 * width + x1);
    float p0 = p00 * (1.0f - yt) + p01 * yt, p1 = p10 * (1.0f - yt) + p11 * yt;
    float p = p0 * (1.0f - xt) * p1 * xt;
    c_out = c * scale; dist = INFINITY; pdf = p;
  }
  void strength_dist_pdf(Vec3 wi, Color& c, float& dist, float& pdf) const { // :128-133
    float theta = std::acos(wi.y);
    float phi = std::atan2(wi.x, wi.z) + PI;
    lookup(theta, phi, c, dist, pdf);
  }
  Vec3 sample(Rng& rng, float& pdf, Color& strength, float& dist) const { // :111-126
    float rand = rng.uniform_1d();
    float dummy; size_t ind = alias.sample(rand, dummy);
    size_t x = ind % width, y = ind / width;
    float rx, ry; rng.uniform_2d(rx, ry);
    float theta = ((float)y + ry) / (float)height * PI;
    float phi = ((float)x + rx) / (float)width * 2.0f * PI;
    float st = std::sin(theta);
    Vec3 wi{st * std::sin(phi), std::cos(theta), st * std::cos(phi)};
    lookup(theta, phi, strength, dist, pdf);
    return wi;
  }
  float power() const { return avg_power * 4.0f * PI; } // :139-141
};

struct Light { // light/mod.rs:28-35, tags follow SPT_LIGHT_*
  int tag = 0;
  Vec3 position, direction; Color strength;
  float cos_inner = 1, cos_outer = 0;
  const Instance* shape = nullptr;
  const EnvLight* env = nullptr;

  Color spot_strength(Vec3 wi) const { // spot.rs:50-55
    float atten = clampf((dot(direction, -wi) - cos_outer) / fmax_(cos_inner - cos_outer, 0.0001f), 0.0f, 1.0f);
    return strength * atten;
  }
  // return (dir, pdf, strength, dist)
  Vec3 sample(Vec3 p, Rng& rng, float& pdf, Color& L, float& dist) const {
    switch (tag) {
      case 0: pdf = 1.0f; L = strength; dist = FLT_MAX; return -direction; // directional.rs:27-29 (direction pre-normalised)
      case 1: { // point.rs:24-30
        Vec3 s = position - p; float d2 = length_squared(s); float d = std::sqrt(d2); s = s / d;
        pdf = 1.0f; L = strength / d2; dist = d; return s;
      }
      case 2: { // spot.rs:59-65
        Vec3 s = position - p; float d2 = length_squared(s); float d = std::sqrt(d2); s = s / d;
        pdf = 1.0f; L = spot_strength(s) / d2; dist = d; return s;
      }
      case 3: { // shape_light.rs:21-43
        float ppdf; Intersection inter = shape->sample(rng, ppdf);
        Color em = inter.surface->emissive(inter);
        Vec3 lv = inter.position - p;
        float d2 = length_squared(lv); float d = std::sqrt(d2);
        Vec3 dir = lv / d;
        float cs;
        if (inter.surface->double_sided) cs = std::fabs(dot(dir, inter.normal));
        else { cs = dot(dir, -inter.normal); if (!(cs > 0.0f)) { cs = 1.0f; em = BLACK; } }
        pdf = ppdf * d2 / fmax_(cs, 0.001f);
        L = em; dist = d; return dir;
      }
      default: return env->sample(rng, pdf, L, dist);
    }
  }
  bool is_delta() const { return tag <= 2; }
  float power(const TextureTable&) const {
    switch (tag) {
      case 0: case 1: case 2: return strength.luminance();
      case 3: return shape->surface_area(Transform()) * shape->surface->average_emissive().luminance(); // shape_light.rs:78-81
      default: return env->power();
    }
  }
};

struct LightSamplerInputs { Vec3 position, normal; }; // light_sampler/mod.rs:12-15

struct LightSampler { // uniform.rs / power_is.rs
  int kind = 0; // SPT_LS_*
  std::vector<Light> lights;
  float num_light_inv = 0;
  int env_light_index = -1;
  AliasTable alias;
  std::vector<int> instance_light; // instance id -> light index (power_is.rs:15, HashMap<InstancePtr,usize>)

  void init(const TextureTable& T) {
    num_light_inv = 1.0f / (float)lights.size(); // uniform.rs:19
    if (kind == 1) {                              // power_is.rs:24-34
      std::vector<float> props(lights.size(), 0.0f);
      float sum = 0.0f;
      for (size_t i = 0; i < lights.size(); i++) { props[i] = lights[i].power(T); sum += props[i]; }
      float sum_inv = 1.0f / sum;
      for (auto& p : props) p *= sum_inv;
      alias.build(std::move(props));
    }
  }
  // returns dir; outputs pdf, strength, dist, is_delta
  Vec3 sample_light(const LightSamplerInputs& in, Rng& rng, float& pdf, Color& L, float& dist, bool& delta) const {
    const Light* light; float pick;
    if (kind == 0) { // uniform.rs:34-41
      float fi = rng.uniform_1d() * (float)lights.size();
      size_t index = std::min((size_t)fi, lights.size() - 1);
      light = &lights[index]; pick = num_light_inv;
      float lp; Vec3 dir = light->sample(in.position, rng, lp, L, dist);
      pdf = lp * pick; delta = light->is_delta(); return dir;
    }
    size_t index = alias.sample(rng.uniform_1d(), pick); // power_is.rs:51-55
    light = &lights[index];
    float lp; Vec3 dir = light->sample(in.position, rng, lp, L, dist);
    pdf = pick * lp; delta = light->is_delta(); return dir;
  }
  float pdf_shape_light(const LightSamplerInputs& in, const Instance* inst, const Intersection& inter) const { // uniform.rs:43-67, power_is.rs:58-88
    float primitive_pdf = inst->pdf(inter);
    Vec3 lv = inter.position - in.position;
    float d2 = length_squared(lv);
    Vec3 dir = lv / std::sqrt(d2);
    float cs;
    if (inter.surface->double_sided) cs = std::fabs(dot(dir, inter.normal));
    else { cs = dot(dir, -inter.normal); if (!(cs > 0.0f)) cs = 1.0f; }
    float local_pdf = primitive_pdf * d2 / fmax_(cs, 0.00001f);
    if (kind == 0) return local_pdf * num_light_inv;
    return local_pdf * alias.probability((size_t)instance_light[inst->inst_id]);
  }
  float pdf_env_light() const { // uniform.rs:69-75, power_is.rs:90-96
    if (env_light_index < 0) return 1.0f;
    return kind == 0 ? num_light_inv : alias.probability((size_t)env_light_index);
  }
};

} // namespace orc
