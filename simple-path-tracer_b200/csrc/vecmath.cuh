// vecmath.cuh — float3 arithmetic for the device code.
//
// The leaf tests (ray/triangle, ray/sphere, instance transform) must return the same f32 `t` as the
// reference's SSE2 glam arithmetic (no FMA), so everything on that path is written with the explicit
// round-to-nearest intrinsics (__fmul_rn/__fadd_rn/__fsub_rn never contract into FFMA).  Shading math uses
// ordinary operators (FMA contraction allowed; tolerance is statistical there).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <math.h>
#include <float.h>

#define SPT_HD __host__ __device__ __forceinline__
#define SPT_D __device__ __forceinline__

#define SPT_PI 3.14159265358979323846f
#define SPT_INV_PI 0.31830988618379067154f

SPT_HD float3 mk3(float x, float y, float z) { return make_float3(x, y, z); }
SPT_HD float3 mk3(float s) { return make_float3(s, s, s); }
SPT_HD float3 mk3(float4 v) { return make_float3(v.x, v.y, v.z); }
SPT_HD float3 operator+(float3 a, float3 b) { return mk3(a.x + b.x, a.y + b.y, a.z + b.z); }
SPT_HD float3 operator-(float3 a, float3 b) { return mk3(a.x - b.x, a.y - b.y, a.z - b.z); }
SPT_HD float3 operator*(float3 a, float3 b) { return mk3(a.x * b.x, a.y * b.y, a.z * b.z); }
SPT_HD float3 operator/(float3 a, float3 b) { return mk3(a.x / b.x, a.y / b.y, a.z / b.z); }
SPT_HD float3 operator*(float3 a, float s) { return mk3(a.x * s, a.y * s, a.z * s); }
SPT_HD float3 operator*(float s, float3 a) { return mk3(a.x * s, a.y * s, a.z * s); }
#ifndef SPT_VEC_RCP
#define SPT_VEC_RCP 1 // vector / scalar as one reciprocal and three products (the shading math is 2-ulp anyway: -prec-div=false)
#endif
#if SPT_VEC_RCP
SPT_HD float3 operator/(float3 a, float s) { const float r = 1.0f / s; return mk3(a.x * r, a.y * r, a.z * r); }
#else
SPT_HD float3 operator/(float3 a, float s) { return mk3(a.x / s, a.y / s, a.z / s); }
#endif
SPT_HD float3 operator-(float3 a) { return mk3(-a.x, -a.y, -a.z); }
SPT_HD float3& operator+=(float3& a, float3 b) { a = a + b; return a; }
SPT_HD float3& operator*=(float3& a, float3 b) { a = a * b; return a; }
SPT_HD float3& operator*=(float3& a, float s) { a = a * s; return a; }
SPT_HD float dot3(float3 a, float3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
SPT_HD float3 cross3(float3 a, float3 b) { return mk3(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x); }
SPT_HD float len3(float3 a) { return sqrtf(dot3(a, a)); }
SPT_HD float3 norm3(float3 a) { return a / len3(a); }
SPT_HD float lum3(float3 c) { return 0.299f * c.x + 0.587f * c.y + 0.114f * c.z; } // core/color.rs:30-32
SPT_HD float avg3(float3 c) { return (c.x + c.y + c.z) / 3.0f; }
SPT_HD bool finite3(float3 c) { return isfinite(c.x) && isfinite(c.y) && isfinite(c.z); }
SPT_HD float3 exp3(float3 c) { return mk3(expf(c.x), expf(c.y), expf(c.z)); }
SPT_HD float3 sqrt3(float3 c) { return mk3(sqrtf(c.x), sqrtf(c.y), sqrtf(c.z)); }
SPT_HD float clampf_(float x, float lo, float hi) { return x < lo ? lo : (x > hi ? hi : x); }
SPT_HD float sqr(float x) { return x * x; }

#ifdef __CUDACC__
// sin/cos of 2*pi*u for u in [0,1]: sincospif has an exact, branch-free range reduction (no Payne-Hanek slow path,
// which sincosf would inline at every call site).
SPT_D void sincos_2pi(float u, float& s, float& c) { sincospif(2.0f * u, &s, &c); }
// ---- exact (non-contracted) arithmetic for the geometric leaf tests --------------------------------------
SPT_D float xmul(float a, float b) { return __fmul_rn(a, b); }
SPT_D float xadd(float a, float b) { return __fadd_rn(a, b); }
SPT_D float xsub(float a, float b) { return __fsub_rn(a, b); }
SPT_D float3 xsub3(float3 a, float3 b) { return mk3(xsub(a.x, b.x), xsub(a.y, b.y), xsub(a.z, b.z)); }
// glam Vec3A::dot on SSE2: (x*x' + y*y') + z*z'
SPT_D float xdot3(float3 a, float3 b) { return xadd(xadd(xmul(a.x, b.x), xmul(a.y, b.y)), xmul(a.z, b.z)); }
// glam Vec3A::cross on SSE2: (a.zxy*b - a*b.zxy).zxy
SPT_D float3 xcross3(float3 a, float3 b) {
  return mk3(xsub(xmul(a.y, b.z), xmul(b.y, a.z)), xsub(xmul(a.z, b.x), xmul(b.z, a.x)), xsub(xmul(a.x, b.y), xmul(b.x, a.y)));
}
// glam Mat3A * Vec3A: (x_axis*v.x + y_axis*v.y) + z_axis*v.z, with rows r0..r2 holding (m_i0, m_i1, m_i2, t_i)
SPT_D float3 xmat_vec(float4 r0, float4 r1, float4 r2, float3 v) {
  return mk3(xadd(xadd(xmul(r0.x, v.x), xmul(r0.y, v.y)), xmul(r0.z, v.z)),
             xadd(xadd(xmul(r1.x, v.x), xmul(r1.y, v.y)), xmul(r1.z, v.z)),
             xadd(xadd(xmul(r2.x, v.x), xmul(r2.y, v.y)), xmul(r2.z, v.z)));
}
SPT_D float3 xmat_point(float4 r0, float4 r1, float4 r2, float3 p) {
  float3 m = xmat_vec(r0, r1, r2, p);
  return mk3(xadd(m.x, r0.w), xadd(m.y, r1.w), xadd(m.z, r2.w));
}
// origin + direction * t  (core/ray.rs:29-31)
SPT_D float3 xpoint_at(float3 o, float3 d, float t) { return mk3(xadd(o.x, xmul(d.x, t)), xadd(o.y, xmul(d.y, t)), xadd(o.z, xmul(d.z, t))); }
#endif
