// oracle/o_math.h -- TEST INFRASTRUCTURE (CPU oracle). Not part of the product; only tests/,
// __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may use it.
//
// Minimal GLSL-like vector math for the CPU restatement of prosper's RtReference shaders.
// Compiled with -ffp-contract=off: every + - * / below is one IEEE binary32 operation, which
// is what makes the "exact" subset (primary ray, instance transform, ray/triangle test,
// DESIGN.md section 4) reproducible bit for bit by the CUDA kernels' __f*_rn intrinsics.
#pragma once
#include <cmath>
#include <cstdint>
#include <cstring>
#include <algorithm>

namespace orc {

constexpr float PI = 3.14159265f; // common/math.glsl:4

struct vec2 { float x, y; };
struct vec3 { float x, y, z; };
struct vec4 { float x, y, z, w; };
struct uvec3 { uint32_t x, y, z; };

inline vec2 operator+(vec2 a, vec2 b) { return {a.x + b.x, a.y + b.y}; }
inline vec2 operator*(vec2 a, float s) { return {a.x * s, a.y * s}; }

inline vec3 operator+(vec3 a, vec3 b) { return {a.x + b.x, a.y + b.y, a.z + b.z}; }
inline vec3 operator-(vec3 a, vec3 b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
inline vec3 operator-(vec3 a) { return {-a.x, -a.y, -a.z}; }
inline vec3 operator*(vec3 a, vec3 b) { return {a.x * b.x, a.y * b.y, a.z * b.z}; }
inline vec3 operator*(vec3 a, float s) { return {a.x * s, a.y * s, a.z * s}; }
inline vec3 operator*(float s, vec3 a) { return {a.x * s, a.y * s, a.z * s}; }
// vector / scalar as the GLSL text says it (sampling.glsl:32, lighting.glsl:29, main.rgen:293): a division per component.
// The oracle never changes for performance: the CUDA kernels lower this to one reciprocal and three multiplies
// (prosper_b200/csrc/hd.h) and have to pass the parity tests WITHIN THEIR STATED TOLERANCE against this, not bit for bit.
inline vec3 operator/(vec3 a, float s) { return {a.x / s, a.y / s, a.z / s}; }
inline vec3 operator/(vec3 a, vec3 b) { return {a.x / b.x, a.y / b.y, a.z / b.z}; }
inline vec3 &operator+=(vec3 &a, vec3 b) { a = a + b; return a; }
inline vec3 &operator*=(vec3 &a, vec3 b) { a = a * b; return a; }
inline vec4 operator+(vec4 a, vec4 b) { return {a.x + b.x, a.y + b.y, a.z + b.z, a.w + b.w}; }
inline vec4 operator*(vec4 a, float s) { return {a.x * s, a.y * s, a.z * s, a.w * s}; }
inline vec4 operator*(vec4 a, vec4 b) { return {a.x * b.x, a.y * b.y, a.z * b.z, a.w * b.w}; }

// dot with the fixed association ((x*x + y*y) + z*z): part of the exact contract
inline float dot(vec3 a, vec3 b) { return ((a.x * b.x) + (a.y * b.y)) + (a.z * b.z); }
inline float dot(vec4 a, vec4 b) { return (((a.x * b.x) + (a.y * b.y)) + (a.z * b.z)) + (a.w * b.w); }
inline vec3 cross(vec3 a, vec3 b)
{
    return {(a.y * b.z) - (a.z * b.y), (a.z * b.x) - (a.x * b.z), (a.x * b.y) - (a.y * b.x)};
}
inline float length(vec3 a) { return std::sqrt(dot(a, a)); }
// normalize = v / sqrt(dot(v,v)) (division, not rsqrt). The primary-ray direction is part of the exact contract (the
// device's xnormalize does the same three divisions); shading normals are compared within tolerance.
inline vec3 normalize(vec3 a) { return a / std::sqrt(dot(a, a)); }
inline vec3 normalizeExact(vec3 a) { return normalize(a); }
// math.glsl:5. GLSL clamp() of a NaN is undefined; contract: NaN -> 0 (fmaxf drops the NaN), which is
// what __saturatef does on the device.
inline float saturate(float x) { return std::fmin(std::fmax(x, 0.0f), 1.0f); }
inline float max3(vec3 v) { return std::max(std::max(v.x, v.y), v.z); }      // math.glsl:8
inline vec3 vmax(vec3 a, vec3 b) { return {std::max(a.x, b.x), std::max(a.y, b.y), std::max(a.z, b.z)}; }
inline vec3 vmin(vec3 a, vec3 b) { return {std::min(a.x, b.x), std::min(a.y, b.y), std::min(a.z, b.z)}; }
inline vec3 clamp(vec3 v, float lo, float hi) // NaN -> lo (same contract as saturate)
{
    return {std::fmin(std::fmax(v.x, lo), hi), std::fmin(std::fmax(v.y, lo), hi), std::fmin(std::fmax(v.z, lo), hi)};
}
inline vec3 mix(vec3 a, vec3 b, float t) { return a * (1.0f - t) + b * t; }
inline vec3 reflect(vec3 i, vec3 n) { return i - n * (2.0f * dot(n, i)); }
inline float sign(float x) { return x > 0.f ? 1.f : (x < 0.f ? -1.f : 0.f); }
inline vec3 xyz(vec4 v) { return {v.x, v.y, v.z}; }

inline uint32_t floatBits(float f) { uint32_t u; std::memcpy(&u, &f, 4); return u; }
inline float bitsFloat(uint32_t u) { float f; std::memcpy(&f, &u, 4); return f; }

// IEEE binary16 -> binary32 (unpackHalf2x16), exact
inline float halfToFloat(uint16_t h)
{
    const uint32_t s = (uint32_t)(h & 0x8000u) << 16;
    uint32_t e = (h >> 10) & 0x1Fu;
    uint32_t m = h & 0x3FFu;
    if (e == 0)
    {
        if (m == 0) return bitsFloat(s);
        // subnormal: normalise
        int shift = 0;
        while ((m & 0x400u) == 0) { m <<= 1; shift++; }
        m &= 0x3FFu;
        return bitsFloat(s | ((uint32_t)(113 - shift) << 23) | (m << 13));
    }
    if (e == 31) return bitsFloat(s | 0x7F800000u | (m << 13));
    return bitsFloat(s | ((e + 112u) << 23) | (m << 13));
}

// binary32 -> binary16, round to nearest even (packHalf2x16 / the fp16 blit)
inline uint16_t floatToHalf(float f)
{
    const uint32_t x = floatBits(f);
    const uint32_t s = (x >> 16) & 0x8000u;
    const uint32_t a = x & 0x7FFFFFFFu;
    if (a >= 0x7F800000u) return (uint16_t)(s | 0x7C00u | ((a > 0x7F800000u) ? 0x200u : 0u));
    if (a >= 0x477FF000u) return (uint16_t)(s | 0x7C00u); // overflows to inf after rounding
    if (a < 0x33000001u) return (uint16_t)s;              // rounds to zero
    int e = (int)(a >> 23) - 127;
    uint32_t m = (a & 0x7FFFFFu) | 0x800000u;
    int shift;
    uint32_t base;
    if (e < -14) { shift = 13 + (-14 - e); base = 0; }
    else { shift = 13; base = (uint32_t)(e + 15) << 10; m &= 0x7FFFFFu; }
    uint32_t r = m >> shift;
    const uint32_t rem = m & ((1u << shift) - 1u);
    const uint32_t half = 1u << (shift - 1);
    if (rem > half || (rem == half && (r & 1u))) r++;
    return (uint16_t)(s | (base + r));
}

} // namespace orc
