// oracle/texture.h — TEST INFRASTRUCTURE ONLY.  Restates core/coord.rs and texture/*.rs.
// PARITY UNPINNED (see vec.h).
#pragma once
#include "prims.h"

namespace orc {

enum DirType { kReflect = 0, kTransmit = 1 };           // BxdfDirType, bxdf/mod.rs:45-49
enum LobeType { kDiffuse = 0, kGlossy = 1, kSpecular = 2 }; // BxdfLobeType, bxdf/mod.rs:38-43

struct Coordinate { // coord.rs:3-8
  Mat3 local_to_world, world_to_local;
  Vec3 hemisphere;
  static Coordinate from_tangent_normal(Vec3 t, Vec3 n, Vec3 hemi) { // :11-23
    Vec3 z = n;
    Vec3 y = normalize(cross(z, t));
    Vec3 x = cross(y, z);
    Coordinate c;
    c.local_to_world = Mat3{x, y, z};
    c.world_to_local = transpose(c.local_to_world);
    c.hemisphere = hemi;
    return c;
  }
  Vec3 to_local(Vec3 w) const { return world_to_local * w; }
  Vec3 to_world(Vec3 l) const { return local_to_world * l; }
  bool in_expected_hemisphere(Vec3 dir, DirType ty) const { // :53-59
    return ty == kReflect ? dot(dir, hemisphere) >= 0.0f : dot(dir, hemisphere) <= 0.0f;
  }
};

struct TextureInput { // texture/mod.rs:51-63
  Vec3 specified, position, normal, tangent, bitangent;
  Vec2 texcoords, duvdx, duvdy;
  int mode = 0; // TextureInputMode::Specified (Default)
  int wrap = 0; // Repeat
  static TextureInput from(const Intersection& i) { // :149-163
    TextureInput t;
    t.texcoords = i.texcoords; t.duvdx = i.duvdx; t.duvdy = i.duvdy; t.mode = 1;
    t.position = i.position; t.normal = i.normal; t.tangent = i.tangent; t.bitangent = i.bitangent;
    return t;
  }
  Vec2 value_vec2() const { // :74-83
    switch (mode) {
      case 0: return {specified.x, specified.y};
      case 1: return texcoords;
      case 2: return {position.x, position.y};
      case 3: return {normal.x, normal.y};
      case 4: return {tangent.x, tangent.y};
      default: return {bitangent.x, bitangent.y};
    }
  }
  static float fract(float x) { return x - std::trunc(x); }
  static int as_i32(float x) { // Rust `as i32`: saturating, NaN -> 0
    if (std::isnan(x)) return 0;
    if (x >= 2147483648.0f) return INT32_MAX;
    if (x <= -2147483648.0f) return INT32_MIN;
    return (int)x;
  }
  Vec2 value_vec2_wrapped() const { // :85-135
    Vec2 v = value_vec2();
    switch (wrap) {
      case 0: {
        float xf = fract(v.x), yf = fract(v.y);
        return {v.x >= 0.0f ? xf : 1.0f + xf, v.y >= 0.0f ? yf : 1.0f + yf};
      }
      case 1: {
        float xf = fract(v.x), yf = fract(v.y);
        float xn = v.x >= 0.0f ? xf : 1.0f + xf;
        xn = (as_i32(v.x) % 2 == 0) ? xn : 1.0f - xn;
        float yn = v.y >= 0.0f ? yf : 1.0f + yf;
        yn = (as_i32(v.y) % 2 == 0) ? yn : 1.0f - yn;
        return {xn, yn};
      }
      case 2: return {clampf(v.x, 0.0f, 1.0f), clampf(v.y, 0.0f, 1.0f)};
      default: return {std::fabs(clampf(v.x, 0.0f, 1.0f)), std::fabs(clampf(v.y, 0.0f, 1.0f))};
    }
  }
};

struct Image8 { int w = 0, h = 0; std::vector<uint8_t> px; const uint8_t* at(int x, int y) const { return &px[4 * ((size_t)y * w + x)]; } };

inline std::vector<Image8> generate_mipmap(Image8 base) { // image_tex.rs:63-98
  int width = base.w, height = base.h;
  std::vector<Image8> images;
  images.push_back(std::move(base));
  while (width > 1 || height > 1) {
    width = (width + 1) >> 1; height = (height + 1) >> 1;
    Image8 img; img.w = width; img.h = height; img.px.resize((size_t)4 * width * height);
    const Image8& last = images.back();
    for (int i = 0; i < width; i++) for (int j = 0; j < height; j++) {
      int x0 = 2 * i, x1 = std::min(2 * i + 1, last.w - 1), y0 = 2 * j, y1 = std::min(2 * j + 1, last.h - 1);
      const uint8_t *p0 = last.at(x0, y0), *p1 = last.at(x0, y1), *p2 = last.at(x1, y0), *p3 = last.at(x1, y1);
      for (int c = 0; c < 4; c++) {
        float f = ((float)p0[c] + (float)p1[c] + (float)p2[c] + (float)p3[c]) * 0.25f;
        img.px[4 * ((size_t)j * width + i) + c] = (uint8_t)f; // `as u8` truncation
      }
    }
    images.push_back(std::move(img));
  }
  return images;
}

struct Vec4 { float x, y, z, w; };
inline Vec4 operator*(Vec4 a, float s) { return {a.x * s, a.y * s, a.z * s, a.w * s}; }
inline Vec4 operator+(Vec4 a, Vec4 b) { return {a.x + b.x, a.y + b.y, a.z + b.z, a.w + b.w}; }
inline Vec4 rgba_to_vec4(const uint8_t* p) { return {p[0] / 255.0f, p[1] / 255.0f, p[2] / 255.0f, p[3] / 255.0f}; } // :155-162

inline Vec4 sample_blinear(const Image8& im, float u, float v) { // image_tex.rs:100-123
  float x = u * (float)im.w;
  int x1 = TextureInput::as_i32(std::round(x)), x0 = x1 - 1;
  float xt = x - (float)x0 - 0.5f;
  x0 = std::clamp(x0, 0, im.w - 1); x1 = std::clamp(x1, 0, im.w - 1);
  float y = v * (float)im.h;
  int y1 = TextureInput::as_i32(std::round(y)), y0 = y1 - 1;
  float yt = y - (float)y0 - 0.5f;
  y0 = std::clamp(y0, 0, im.h - 1); y1 = std::clamp(y1, 0, im.h - 1);
  Vec4 c00 = rgba_to_vec4(im.at(x0, y0)), c01 = rgba_to_vec4(im.at(x0, y1));
  Vec4 c10 = rgba_to_vec4(im.at(x1, y0)), c11 = rgba_to_vec4(im.at(x1, y1));
  Vec4 c0 = c00 * (1.0f - yt) + c01 * yt;
  Vec4 c1 = c10 * (1.0f - yt) + c11 * yt;
  return c0 * (1.0f - xt) + c1 * xt;
}

inline Vec4 sample_trilinear(const std::vector<Image8>& images, float u, float v, Vec2 duvdx, Vec2 duvdy) { // :125-153
  if (images.empty()) return {0, 0, 0, 0};
  Vec2 scale{(float)images[0].w, (float)images[0].h};
  duvdx = duvdx * scale; duvdy = duvdy * scale;
  float level = clampf(std::log2(fmax_(length(duvdx), length(duvdy)) + 0.001f), 0.0f, (float)(images.size() - 1));
  size_t l0 = (size_t)std::floor(level);
  if (l0 + 1 == images.size()) return sample_blinear(images[l0], u, v);
  size_t l1 = l0 + 1;
  float lt = level - (float)l0;
  return sample_blinear(images[l0], u, v) * (1.0f - lt) + sample_blinear(images[l1], u, v) * lt;
}

inline float srgb_to_linear(float s) { // srgb_tex.rs:55-61
  return s <= 0.04045f ? s / 12.92f : std::pow((s + 0.055f) / 1.055f, 2.4f);
}

struct Texture { // texture/mod.rs:198-208 (tags follow include/spt.h)
  int tag = 0;
  int child[2] = {-1, -1};
  Color value;
  Vec3 tiling{1, 1, 1}, offset{0, 0, 0};
  int mode = -1, wrap = -1;
  std::vector<Image8> images;
};

struct TextureTable {
  std::vector<Texture> tex;
  static float chan_of(Vec4 v, int chan) { return chan == 0 ? v.x : chan == 1 ? v.y : chan == 2 ? v.z : v.w; }
  static TextureInput apply_modifier(const Texture& t, const TextureInput& in) { // input_modifier.rs:34-47
    TextureInput o;
    o.specified = in.specified * t.tiling + t.offset;
    o.position = in.position * t.tiling + t.offset;
    o.normal = in.normal * t.tiling + t.offset;
    o.tangent = in.tangent * t.tiling + t.offset;
    o.bitangent = in.bitangent * t.tiling + t.offset;
    Vec2 t2{t.tiling.x, t.tiling.y}, o2{t.offset.x, t.offset.y};
    o.texcoords = in.texcoords * t2 + o2;
    o.duvdx = in.duvdx * t2; o.duvdy = in.duvdy * t2;
    o.mode = t.mode >= 0 ? t.mode : in.mode;
    o.wrap = t.wrap >= 0 ? t.wrap : in.wrap;
    return o;
  }
  Color color_at(int id, const TextureInput& in) const {
    const Texture& t = tex[id];
    switch (t.tag) {
      case 0: return t.value;
      case 1: { Vec2 uv = in.value_vec2_wrapped(); Vec4 v = sample_trilinear(t.images, uv.x, uv.y, in.duvdx, in.duvdy); return {v.x, v.y, v.z}; }
      case 2: return color_at(t.child[0], in) + color_at(t.child[1], in);
      case 3: return color_at(t.child[0], in) - color_at(t.child[1], in);
      case 4: return color_at(t.child[0], in) * color_at(t.child[1], in);
      case 5: return color_at(t.child[0], in) / color_at(t.child[1], in);
      case 6: { Color c = color_at(t.child[0], in); return {srgb_to_linear(c.r), srgb_to_linear(c.g), srgb_to_linear(c.b)}; }
      default: return color_at(t.child[0], apply_modifier(t, in));
    }
  }
  float float_at(int id, const TextureInput& in, int chan) const {
    const Texture& t = tex[id];
    switch (t.tag) {
      case 0: return chan == 0 ? t.value.r : chan == 1 ? t.value.g : chan == 2 ? t.value.b : 1.0f; // scalar.rs:26-33
      case 1: { Vec2 uv = in.value_vec2_wrapped(); return chan_of(sample_trilinear(t.images, uv.x, uv.y, in.duvdx, in.duvdy), chan); }
      case 2: return float_at(t.child[0], in, chan) + float_at(t.child[1], in, chan);
      case 3: return float_at(t.child[0], in, chan) - float_at(t.child[1], in, chan);
      case 4: return float_at(t.child[0], in, chan) * float_at(t.child[1], in, chan);
      case 5: return float_at(t.child[0], in, chan) / float_at(t.child[1], in, chan);
      case 6: { float f = float_at(t.child[0], in, chan); return chan == 3 ? f : srgb_to_linear(f); }
      default: return float_at(t.child[0], apply_modifier(t, in), chan);
    }
  }
  Color average_color(int id) const { // TextureT::average_color
    const Texture& t = tex[id];
    switch (t.tag) {
      case 0: return t.value;
      case 1: { Vec4 v = rgba_to_vec4(t.images.back().at(0, 0)); return {v.x, v.y, v.z}; } // image_tex.rs:41-44
      case 2: return average_color(t.child[0]) + average_color(t.child[1]);
      case 3: return average_color(t.child[0]) - average_color(t.child[1]);
      case 4: return average_color(t.child[0]) * average_color(t.child[1]);
      case 5: return average_color(t.child[0]) / average_color(t.child[1]);
      case 6: { Color c = average_color(t.child[0]); return {srgb_to_linear(c.r), srgb_to_linear(c.g), srgb_to_linear(c.b)}; }
      default: return average_color(t.child[0]);
    }
  }
};

} // namespace orc
