// oracle/vec.h — TEST INFRASTRUCTURE ONLY (CPU restatement of the reference; never shipped,
// never linked into the product library).  PARITY UNPINNED: the reference has no tests or
// golden vectors and cannot be built here (no cargo/rustc).
//
// Restates the third-party arithmetic the reference's hot path uses: glam 0.20 (Cargo.toml:13)
// Vec3A / Vec2 / Mat3A / Affine3A on the SSE2 path — plain IEEE f32, no FMA (compile with
// -ffp-contract=off), dot = (x*x + y*y) + z*z, normalize = v / length — and the reference's own
// Color type (src/core/color.rs:3-182).
#pragma once
#include <cmath>
#include <cfloat>
#include <cstdint>
#include <algorithm>

namespace orc {

constexpr float PI = 3.14159265358979323846264338327950288f;        // std::f32::consts::PI
constexpr float FRAC_1_PI = 0.318309886183790671537767526745028724f; // std::f32::consts::FRAC_1_PI

// Rust f32::min/max ignore a NaN operand (return the other one) == fminf/fmaxf.
inline float fmin_(float a, float b) { return std::fmin(a, b); }
inline float fmax_(float a, float b) { return std::fmax(a, b); }
// Rust f32::clamp: NaN stays NaN; otherwise max(min) with plain comparisons.
inline float clampf(float x, float lo, float hi) {
  if (x < lo) x = lo;
  if (x > hi) x = hi;
  return x;
}

struct Vec2 {
  float x = 0, y = 0;
  Vec2() = default;
  Vec2(float x_, float y_) : x(x_), y(y_) {}
};
inline Vec2 operator+(Vec2 a, Vec2 b) { return {a.x + b.x, a.y + b.y}; }
inline Vec2 operator-(Vec2 a, Vec2 b) { return {a.x - b.x, a.y - b.y}; }
inline Vec2 operator*(Vec2 a, Vec2 b) { return {a.x * b.x, a.y * b.y}; }
inline Vec2 operator*(Vec2 a, float s) { return {a.x * s, a.y * s}; }
inline float length(Vec2 a) { return std::sqrt(a.x * a.x + a.y * a.y); }

struct Vec3 {
  float x = 0, y = 0, z = 0;
  Vec3() = default;
  Vec3(float x_, float y_, float z_) : x(x_), y(y_), z(z_) {}
  float operator[](int i) const { return i == 0 ? x : (i == 1 ? y : z); }
};
inline Vec3 operator+(Vec3 a, Vec3 b) { return {a.x + b.x, a.y + b.y, a.z + b.z}; }
inline Vec3 operator-(Vec3 a, Vec3 b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
inline Vec3 operator*(Vec3 a, Vec3 b) { return {a.x * b.x, a.y * b.y, a.z * b.z}; }
inline Vec3 operator/(Vec3 a, Vec3 b) { return {a.x / b.x, a.y / b.y, a.z / b.z}; }
inline Vec3 operator*(Vec3 a, float s) { return {a.x * s, a.y * s, a.z * s}; }
inline Vec3 operator*(float s, Vec3 a) { return {s * a.x, s * a.y, s * a.z}; }
inline Vec3 operator/(Vec3 a, float s) { return {a.x / s, a.y / s, a.z / s}; }
inline Vec3 operator-(Vec3 a) { return {-a.x, -a.y, -a.z}; }
inline Vec3& operator+=(Vec3& a, Vec3 b) { a = a + b; return a; }
inline float dot(Vec3 a, Vec3 b) { return (a.x * b.x + a.y * b.y) + a.z * b.z; }
inline Vec3 cross(Vec3 a, Vec3 b) {
  return {a.y * b.z - b.y * a.z, a.z * b.x - b.z * a.x, a.x * b.y - b.x * a.y};
}
inline float length_squared(Vec3 a) { return dot(a, a); }
inline float length(Vec3 a) { return std::sqrt(dot(a, a)); }
inline Vec3 normalize(Vec3 a) { return a / length(a); }
inline float distance(Vec3 a, Vec3 b) { return length(a - b); }
inline float distance_squared(Vec3 a, Vec3 b) { return length_squared(a - b); }
inline Vec3 vmin(Vec3 a, Vec3 b) { return {std::min(a.x, b.x), std::min(a.y, b.y), std::min(a.z, b.z)}; } // _mm_min_ps
inline Vec3 vmax(Vec3 a, Vec3 b) { return {std::max(a.x, b.x), std::max(a.y, b.y), std::max(a.z, b.z)}; }

struct Mat3 { // glam::Mat3A, column vectors
  Vec3 x{1, 0, 0}, y{0, 1, 0}, z{0, 0, 1};
};
inline Vec3 operator*(const Mat3& m, Vec3 v) { return (m.x * v.x + m.y * v.y) + m.z * v.z; }
inline Mat3 operator*(const Mat3& a, const Mat3& b) { return {a * b.x, a * b.y, a * b.z}; }
inline Mat3 transpose(const Mat3& m) {
  return {{m.x.x, m.y.x, m.z.x}, {m.x.y, m.y.y, m.z.y}, {m.x.z, m.y.z, m.z.z}};
}
inline float determinant(const Mat3& m) { return dot(m.z, cross(m.x, m.y)); }
inline Mat3 inverse(const Mat3& m) { // glam Mat3A::inverse
  Vec3 t0 = cross(m.y, m.z), t1 = cross(m.z, m.x), t2 = cross(m.x, m.y);
  float det = dot(m.z, t2);
  float inv = 1.0f / det;
  Mat3 r{t0 * inv, t1 * inv, t2 * inv};
  return transpose(r);
}

struct Affine3 { // glam::Affine3A
  Mat3 m;
  Vec3 t;
};
inline Vec3 transform_point(const Affine3& a, Vec3 p) { return a.m * p + a.t; }
inline Vec3 transform_vector(const Affine3& a, Vec3 v) { return a.m * v; }
inline Affine3 inverse(const Affine3& a) {
  Mat3 mi = inverse(a.m);
  return {mi, -(mi * a.t)};
}
inline Affine3 operator*(const Affine3& a, const Affine3& b) { return {a.m * b.m, a.m * b.t + a.t}; }

// ---- src/core/color.rs ------------------------------------------------------------------------
struct Color {
  float r = 0, g = 0, b = 0;
  Color() = default;
  Color(float r_, float g_, float b_) : r(r_), g(g_), b(b_) {}
  static Color gray(float v) { return {v, v, v}; }
  float luminance() const { return 0.299f * r + 0.587f * g + 0.114f * b; } // color.rs:30-32
  float avg() const { return (r + g + b) / 3.0f; }                         // color.rs:34-36
  bool is_finite() const { return std::isfinite(r) && std::isfinite(g) && std::isfinite(b); }
  Color exp() const { return {std::exp(r), std::exp(g), std::exp(b)}; }
  Color sqrt() const { return {std::sqrt(r), std::sqrt(g), std::sqrt(b)}; }
};
const Color BLACK{0, 0, 0}, WHITE{1, 1, 1};
inline Color operator+(Color a, Color b) { return {a.r + b.r, a.g + b.g, a.b + b.b}; }
inline Color operator-(Color a, Color b) { return {a.r - b.r, a.g - b.g, a.b - b.b}; }
inline Color operator*(Color a, Color b) { return {a.r * b.r, a.g * b.g, a.b * b.b}; }
inline Color operator/(Color a, Color b) { return {a.r / b.r, a.g / b.g, a.b / b.b}; }
inline Color operator*(Color a, float s) { return {a.r * s, a.g * s, a.b * s}; }
inline Color operator*(float s, Color a) { return {s * a.r, s * a.g, s * a.b}; }
inline Color operator/(Color a, float s) { float i = 1.0f / s; return {a.r * i, a.g * i, a.b * i}; } // color.rs:125-131 (Q24)
inline Color operator/(float s, Color a) { return {s / a.r, s / a.g, s / a.b}; } // color.rs:140-146
inline Color operator-(Color a) { return {-a.r, -a.g, -a.b}; }
inline Color& operator+=(Color& a, Color b) { a = a + b; return a; }
inline Color& operator*=(Color& a, Color b) { a = a * b; return a; }
inline Color& operator/=(Color& a, float s) { a = a / s; return a; }

} // namespace orc
