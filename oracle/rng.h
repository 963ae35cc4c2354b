// oracle/rng.h — TEST INFRASTRUCTURE ONLY.  Restates src/core/rng.rs:7-80.
//
// The reference seeds rand 0.8 `SmallRng` (xoshiro256++ on 64-bit targets; Cargo.toml:14, not
// vendored) from OS entropy (rng.rs:10), so only the *distribution* of uniform_1d() is part of
// its behaviour: `Standard` f32 = (next_u32() >> 8) * 2^-24, next_u32 = upper half of next_u64.
// Two generators are provided:
//   kXoshiro — xoshiro256++ seeded by SplitMix64 from an explicit seed (what SmallRng::seed_from_u64 does);
//              one stream per render thread like pt.rs:262.  The "reference-faithful" mode.
//   kKeyed   — the counter-keyed PCG-XSH-RR stream the CUDA core uses (one stream per
//              (seed, pixel, sample)), restated here independently so that a path traced by the
//              oracle and by the GPU consumes the same numbers ("paired" parity mode).
#pragma once
#include <cstdint>
#include "vec.h"

namespace orc {

inline uint64_t splitmix64(uint64_t& s) {
  s += 0x9E3779B97F4A7C15ull;
  uint64_t z = s;
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
  return z ^ (z >> 31);
}
inline uint64_t mix64(uint64_t z) { return splitmix64(z); }

struct Rng {
  enum Mode { kXoshiro, kKeyed, kList } mode = kXoshiro;
  const float* list = nullptr; int list_n = 0, list_i = 0; // kList: caller-supplied numbers (bxdf test hook)
  uint64_t s[4] = {1, 2, 3, 4}; // xoshiro256++
  uint64_t pcg = 0;             // keyed stream state
  uint64_t draws = 0;

  static Rng xoshiro(uint64_t seed) {
    Rng r;
    r.mode = kXoshiro;
    for (int i = 0; i < 4; i++) r.s[i] = splitmix64(seed);
    return r;
  }
  // Stream key: must match simple-path-tracer_b200/csrc/rng.cuh (rng_init).
  void key(uint64_t seed, uint64_t pixel, uint64_t sample) {
    mode = kKeyed;
    pcg = mix64(mix64(seed ^ (pixel * 0xD1B54A32D192ED03ull)) + sample);
  }
  static inline uint64_t rotl(uint64_t x, int k) { return (x << k) | (x >> (64 - k)); }
  uint32_t next_u32() {
    draws++;
    if (mode == kXoshiro) {
      uint64_t result = rotl(s[0] + s[3], 23) + s[0];
      uint64_t t = s[1] << 17;
      s[2] ^= s[0]; s[3] ^= s[1]; s[1] ^= s[2]; s[0] ^= s[3];
      s[2] ^= t;
      s[3] = rotl(s[3], 45);
      return (uint32_t)(result >> 32);
    }
    uint64_t old = pcg;
    pcg = old * 6364136223846793005ull + 1442695040888963407ull;
    uint32_t xs = (uint32_t)(((old >> 18u) ^ old) >> 27u);
    uint32_t rot = (uint32_t)(old >> 59u);
    return (xs >> rot) | (xs << ((32 - rot) & 31));
  }
  float uniform_1d() { // rng.rs:14-16
    if (mode == kList) { draws++; return list_i < list_n ? list[list_i++] : 0.5f; }
    return (float)(next_u32() >> 8) * (1.0f / 16777216.0f);
  }
  void uniform_2d(float& a, float& b) { a = uniform_1d(); b = uniform_1d(); }    // rng.rs:18-20

  Vec3 uniform_on_sphere() { // rng.rs:56-63
    float rx, ry; uniform_2d(rx, ry);
    float phi = rx * 2.0f * PI;
    float sp = std::sin(phi), cp = std::cos(phi);
    float cos_theta = 1.0f - 2.0f * ry;
    float sin_theta = std::sqrt(1.0f - cos_theta * cos_theta);
    return {sin_theta * cp, sin_theta * sp, cos_theta};
  }
  Vec3 cosine_weighted_on_hemisphere() { // rng.rs:72-80
    float rx, ry; uniform_2d(rx, ry);
    float phi = rx * 2.0f * PI;
    float sp = std::sin(phi), cp = std::cos(phi);
    float sin_theta_sqr = ry;
    float sin_theta = std::sqrt(sin_theta_sqr);
    float cos_theta = std::sqrt(1.0f - sin_theta_sqr);
    return {sin_theta * cp, sin_theta * sp, cos_theta};
  }
};

} // namespace orc
