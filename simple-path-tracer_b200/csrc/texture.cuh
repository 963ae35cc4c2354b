// texture.cuh — device evaluation of the reference's texture expression trees.
//   TextureT::{color_at,float_at}: scalar.rs:21-33, image_tex.rs:24-37,100-153, binary_op.rs:47-57,
//   srgb_tex.rs:17-28,55-69, input_modifier.rs:34-47,70-77; wrap modes texture/mod.rs:85-135.
// Image textures use the reference's own bilinear convention (round(), -0.5, clamp) on a software RGBA8 mip
// pyramid in global memory — not the texture unit, whose filtering convention differs.
#pragma once
#include "scene.h"
#include "vecmath.cuh"

struct TexIn { // TextureInput, texture/mod.rs:51-63
  float3 specified, position, normal, tangent, bitangent;
  float2 uv, duvdx, duvdy;
  int mode, wrap;
};

SPT_D float fractf_(float x) { return x - truncf(x); }
SPT_D int f2i_sat(float x) { return isnan(x) ? 0 : __float2int_rz(x); } // Rust `as i32` (cvt.rzi saturates)

SPT_D float2 texin_value2(const TexIn& in) {
  switch (in.mode) {
    case 0: return make_float2(in.specified.x, in.specified.y);
    case 1: return in.uv;
    case 2: return make_float2(in.position.x, in.position.y);
    case 3: return make_float2(in.normal.x, in.normal.y);
    case 4: return make_float2(in.tangent.x, in.tangent.y);
    default: return make_float2(in.bitangent.x, in.bitangent.y);
  }
}
SPT_D float2 texin_wrapped(const TexIn& in) {
  const float2 v = texin_value2(in);
  switch (in.wrap) {
    case 0: {
      const float xf = fractf_(v.x), yf = fractf_(v.y);
      return make_float2(v.x >= 0.f ? xf : 1.f + xf, v.y >= 0.f ? yf : 1.f + yf);
    }
    case 1: {
      const float xf = fractf_(v.x), yf = fractf_(v.y);
      float xn = v.x >= 0.f ? xf : 1.f + xf, yn = v.y >= 0.f ? yf : 1.f + yf;
      if (f2i_sat(v.x) % 2 != 0) xn = 1.f - xn;
      if (f2i_sat(v.y) % 2 != 0) yn = 1.f - yn;
      return make_float2(xn, yn);
    }
    case 2: return make_float2(clampf_(v.x, 0.f, 1.f), clampf_(v.y, 0.f, 1.f));
    default: return make_float2(fabsf(clampf_(v.x, 0.f, 1.f)), fabsf(clampf_(v.y, 0.f, 1.f)));
  }
}

SPT_D float4 texel4(const DScene& sc, const DMip& m, int x, int y) {
  const uchar4 p = sc.texels[m.texel_base + (uint32_t)y * (uint32_t)m.w + (uint32_t)x];
  return make_float4(p.x / 255.0f, p.y / 255.0f, p.z / 255.0f, p.w / 255.0f);
}
SPT_D float4 lerp4(float4 a, float4 b, float t) {
  const float s = 1.0f - t;
  return make_float4(a.x * s + b.x * t, a.y * s + b.y * t, a.z * s + b.z * t, a.w * s + b.w * t);
}
SPT_D float4 tex_bilinear(const DScene& sc, const DMip& m, float u, float v) { // image_tex.rs:100-123
  const float x = u * (float)m.w;
  int x1 = f2i_sat(roundf(x)), x0 = x1 - 1;
  const float xt = x - (float)x0 - 0.5f;
  x0 = min(max(x0, 0), m.w - 1); x1 = min(max(x1, 0), m.w - 1);
  const float y = v * (float)m.h;
  int y1 = f2i_sat(roundf(y)), y0 = y1 - 1;
  const float yt = y - (float)y0 - 0.5f;
  y0 = min(max(y0, 0), m.h - 1); y1 = min(max(y1, 0), m.h - 1);
  const float4 c0 = lerp4(texel4(sc, m, x0, y0), texel4(sc, m, x0, y1), yt);
  const float4 c1 = lerp4(texel4(sc, m, x1, y0), texel4(sc, m, x1, y1), yt);
  return lerp4(c0, c1, xt);
}
SPT_D float4 tex_trilinear(const DScene& sc, const DTexture& t, float u, float v, float2 dx, float2 dy) { // :125-153
  if (t.n_mips <= 0) return make_float4(0, 0, 0, 0);
  const DMip m0 = sc.mips[t.mip_base];
  dx.x *= (float)m0.w; dx.y *= (float)m0.h; dy.x *= (float)m0.w; dy.y *= (float)m0.h;
  const float l = fmaxf(sqrtf(dx.x * dx.x + dx.y * dx.y), sqrtf(dy.x * dy.x + dy.y * dy.y)) + 0.001f;
  const float level = clampf_(log2f(l), 0.0f, (float)(t.n_mips - 1));
  const int l0 = (int)floorf(level);
  if (l0 + 1 == t.n_mips) return tex_bilinear(sc, sc.mips[t.mip_base + l0], u, v);
  const float lt = level - (float)l0;
  return lerp4(tex_bilinear(sc, sc.mips[t.mip_base + l0], u, v), tex_bilinear(sc, sc.mips[t.mip_base + l0 + 1], u, v), lt);
}

SPT_D float srgb_lin(float s) { return s <= 0.04045f ? s / 12.92f : powf((s + 0.055f) / 1.055f, 2.4f); }

SPT_D TexIn tex_modify(const DTexture& t, const TexIn& in) { // input_modifier.rs:34-47
  const float3 ti = mk3(t.tiling[0], t.tiling[1], t.tiling[2]), of = mk3(t.offset[0], t.offset[1], t.offset[2]);
  TexIn o;
  o.specified = in.specified * ti + of; o.position = in.position * ti + of; o.normal = in.normal * ti + of;
  o.tangent = in.tangent * ti + of; o.bitangent = in.bitangent * ti + of;
  o.uv = make_float2(in.uv.x * ti.x + of.x, in.uv.y * ti.y + of.y);
  o.duvdx = make_float2(in.duvdx.x * ti.x, in.duvdx.y * ti.y);
  o.duvdy = make_float2(in.duvdy.x * ti.x, in.duvdy.y * ti.y);
  o.mode = t.mode >= 0 ? t.mode : in.mode;
  o.wrap = t.wrap >= 0 ? t.wrap : in.wrap;
  return o;
}

// Evaluates the texture as RGBA (alpha = 1 for scalars, scalar.rs:26-33).  Unary wrappers (sRGB, input modifier) are
// peeled iteratively; a binary operator evaluates its two children with the next-lower template depth, so there is
// no device recursion and the stack frame is known statically.  SPT_TEX_MAX_DEPTH nested binary operators are
// supported (the host flattener rejects deeper trees); every shipped scene needs at most 2.
#define SPT_TEX_MAX_DEPTH 3
template <int D>
__device__ __noinline__ float4 tex_eval_d(const DScene& sc, int id, TexIn in) {
  bool srgb = false;
  for (int guard = 0; guard < 8; guard++) {
    const DTexture& t = sc.textures[id];
    if (t.tag == 6) { srgb = true; id = t.child0; continue; }
    if (t.tag == 7) { in = tex_modify(t, in); id = t.child0; continue; }
    break;
  }
  const DTexture& t = sc.textures[id];
  float4 r = make_float4(0, 0, 0, 0);
  if (t.tag == 0) r = make_float4(t.value[0], t.value[1], t.value[2], 1.0f);
  else if (t.tag == 1) { const float2 uv = texin_wrapped(in); r = tex_trilinear(sc, t, uv.x, uv.y, in.duvdx, in.duvdy); }
  else if (t.tag >= 2 && t.tag <= 5) {
    if constexpr (D > 0) {
      const float4 a = tex_eval_d<D - 1>(sc, t.child0, in), b = tex_eval_d<D - 1>(sc, t.child1, in);
      if (t.tag == 2) r = make_float4(a.x + b.x, a.y + b.y, a.z + b.z, a.w + b.w);
      else if (t.tag == 3) r = make_float4(a.x - b.x, a.y - b.y, a.z - b.z, a.w - b.w);
      else if (t.tag == 4) r = make_float4(a.x * b.x, a.y * b.y, a.z * b.z, a.w * b.w);
      else r = make_float4(a.x / b.x, a.y / b.y, a.z / b.z, a.w / b.w);
    }
  }
  if (srgb) r = make_float4(srgb_lin(r.x), srgb_lin(r.y), srgb_lin(r.z), r.w);
  return r;
}
// SPT_LEAN (a second compilation of the shade kernels, launched for scenes whose textures are all plain scalars and whose surfaces have
// no normal / emissive map — every benchmark config): a texture IS its value, so the expression-tree evaluator, the TextureInput
// record it takes by value and the camera-ray differentials that feed it drop out of the kernels.
#ifndef SPT_LEAN
#define SPT_LEAN 0
#endif
SPT_D float4 tex_eval(const DScene& sc, int id, const TexIn& in) {
  const DTexture& t = sc.textures[id];
#if SPT_LEAN
  return make_float4(t.value[0], t.value[1], t.value[2], 1.0f);
#else
  if (t.tag == 0) return make_float4(t.value[0], t.value[1], t.value[2], 1.0f); // fast path: every benchmark scene
  return tex_eval_d<SPT_TEX_MAX_DEPTH>(sc, id, in);
#endif
}
SPT_D float3 tex_color(const DScene& sc, int id, const TexIn& in) { return mk3(tex_eval(sc, id, in)); }
SPT_D float tex_float(const DScene& sc, int id, const TexIn& in, int chan) {
  const float4 v = tex_eval(sc, id, in);
  return chan == 0 ? v.x : chan == 1 ? v.y : chan == 2 ? v.z : v.w;
}
