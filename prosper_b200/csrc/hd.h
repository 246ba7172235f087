// hd.h -- portability layer so the per-thread bodies of the kernels (core_*.h) compile both with
// nvcc for sm_100a (the product) and with g++ for the kernel-logic simulator under tests/hostsim
// (TEST INFRASTRUCTURE: lets the CPU-only CI exercise the same builder / traversal / shading code
// the GPU runs; it is never part of libprt_b200.so).
#pragma once
#include <stdint.h>
#include <math.h>
#include <string.h>

#ifdef __CUDACC__
#define HD __host__ __device__ __forceinline__
#define HDNI __host__ __device__
#else
#define HD inline
#define HDNI inline
#endif

#ifndef __CUDACC__
// the CUDA vector types the per-thread bodies use, for the host (simulator) build
struct uint2 { uint32_t x, y; };
struct alignas(16) uint4 { uint32_t x, y, z, w; };
struct alignas(16) float4 { float x, y, z, w; };
inline uint2 make_uint2(uint32_t x, uint32_t y) { return uint2{x, y}; }
inline uint4 make_uint4(uint32_t x, uint32_t y, uint32_t z, uint32_t w) { return uint4{x, y, z, w}; }
inline float4 make_float4(float x, float y, float z, float w) { return float4{x, y, z, w}; }
#endif

namespace prt {

// ---- exact IEEE binary32 operations (never contracted into FMA) ---------------------------------
// The "exact contract": primary ray, world->object ray transform and the ray/triangle test are
// evaluated with one rounding per operation in a fixed association on both the oracle
// (g++ -ffp-contract=off) and the device (__f*_rn intrinsics), which makes hit IDs bit-exact.
#if defined(__CUDA_ARCH__)
HD float xmul(float a, float b) { return __fmul_rn(a, b); }
HD float xadd(float a, float b) { return __fadd_rn(a, b); }
HD float xsub(float a, float b) { return __fsub_rn(a, b); }
HD float xdiv(float a, float b) { return __fdiv_rn(a, b); }
HD float xsqrt(float a) { return __fsqrt_rn(a); }
HD uint32_t f2u(float f) { return __float_as_uint(f); }
HD float u2f(uint32_t u) { return __uint_as_float(u); }
HD int popc(uint32_t x) { return __popc(x); }
HD int bfind(uint32_t x) { return 31 - __clz(x); } // index of the highest set bit (x != 0)
HD int clz64(uint64_t x) { return __clzll((long long)x); }
HD float fminf_(float a, float b) { return fminf(a, b); }
HD float fmaxf_(float a, float b) { return fmaxf(a, b); }
#else
// host build: compiled with -ffp-contract=off so plain operators are single roundings
HD float xmul(float a, float b) { return a * b; }
HD float xadd(float a, float b) { return a + b; }
HD float xsub(float a, float b) { return a - b; }
HD float xdiv(float a, float b) { return a / b; }
HD float xsqrt(float a) { return sqrtf(a); }
HD uint32_t f2u(float f) { uint32_t u; memcpy(&u, &f, 4); return u; }
HD float u2f(uint32_t u) { float f; memcpy(&f, &u, 4); return f; }
HD int popc(uint32_t x) { return __builtin_popcount(x); }
HD int bfind(uint32_t x) { return 31 - __builtin_clz(x); }
HD int clz64(uint64_t x) { return x ? __builtin_clzll(x) : 64; }
HD float fminf_(float a, float b) { return fminf(a, b); }
HD float fmaxf_(float a, float b) { return fmaxf(a, b); }
#endif

struct V2 { float x, y; };
struct V3 { float x, y, z; };
struct V4 { float x, y, z, w; };
struct U3 { uint32_t x, y, z; };

HD V3 v3(float x, float y, float z) { return V3{x, y, z}; }
HD V3 operator+(V3 a, V3 b) { return {a.x + b.x, a.y + b.y, a.z + b.z}; }
HD V3 operator-(V3 a, V3 b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
HD V3 operator-(V3 a) { return {-a.x, -a.y, -a.z}; }
HD V3 operator*(V3 a, V3 b) { return {a.x * b.x, a.y * b.y, a.z * b.z}; }
HD V3 operator*(V3 a, float s) { return {a.x * s, a.y * s, a.z * s}; }
HD V3 operator*(float s, V3 a) { return {a.x * s, a.y * s, a.z * s}; }
// vector / scalar = one IEEE reciprocal and three multiplies, on the device, in the simulator and in the oracle alike
// (GLSL leaves the precision of vector division and normalize() to the implementation, which lowers them this way; three
// IEEE divisions per call were 19 % of k_shade's instructions)
HD V3 operator/(V3 a, float s) { const float r = 1.0f / s; return {a.x * r, a.y * r, a.z * r}; }
HD float dot(V3 a, V3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
HD V3 cross(V3 a, V3 b) { return {a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x}; }
HD float length(V3 a) { return sqrtf(dot(a, a)); }
HD V3 normalize(V3 a) { return a / sqrtf(dot(a, a)); }
HD V3 vmin(V3 a, V3 b) { return {fminf_(a.x, b.x), fminf_(a.y, b.y), fminf_(a.z, b.z)}; }
HD V3 vmax(V3 a, V3 b) { return {fmaxf_(a.x, b.x), fmaxf_(a.y, b.y), fmaxf_(a.z, b.z)}; }
// NaN -> 0 like the oracle's saturate (fmaxf drops the NaN); DESIGN.md section 4
HD float saturate(float x) { return fminf_(fmaxf_(x, 0.0f), 1.0f); }
HD float max3(V3 v) { return fmaxf_(fmaxf_(v.x, v.y), v.z); }
HD V3 clamp3(V3 v, float lo, float hi)
{
    return {fminf_(fmaxf_(v.x, lo), hi), fminf_(fmaxf_(v.y, lo), hi), fminf_(fmaxf_(v.z, lo), hi)};
}
HD V3 mix3(V3 a, V3 b, float t) { return a * (1.0f - t) + b * t; }
HD V3 reflect3(V3 i, V3 n) { return i - n * (2.0f * dot(n, i)); }
HD float axis(V3 v, int a) { return a == 0 ? v.x : (a == 1 ? v.y : v.z); }

// ---- exact-contract vector helpers (association fixed, one rounding per op) --------------------
HD V3 xsub3(V3 a, V3 b) { return {xsub(a.x, b.x), xsub(a.y, b.y), xsub(a.z, b.z)}; }
HD float xdot(V3 a, V3 b) { return xadd(xadd(xmul(a.x, b.x), xmul(a.y, b.y)), xmul(a.z, b.z)); }
HD V3 xcross(V3 a, V3 b)
{
    return {xsub(xmul(a.y, b.z), xmul(a.z, b.y)), xsub(xmul(a.z, b.x), xmul(a.x, b.z)),
            xsub(xmul(a.x, b.y), xmul(a.y, b.x))};
}
HD V3 xnormalize(V3 a)
{
    const float l = xsqrt(xdot(a, a));
    return {xdiv(a.x, l), xdiv(a.y, l), xdiv(a.z, l)};
}

// IEEE binary16 -> binary32, exact (unpackHalf2x16, geometry.glsl:71-93)
HD float halfToFloat(uint16_t h)
{
    const uint32_t s = (uint32_t)(h & 0x8000u) << 16;
    const uint32_t e = (h >> 10) & 0x1Fu;
    uint32_t m = h & 0x3FFu;
    if (e == 0)
    {
        if (m == 0) return u2f(s);
        int shift = 0;
        while ((m & 0x400u) == 0) { m <<= 1; shift++; }
        m &= 0x3FFu;
        return u2f(s | ((uint32_t)(113 - shift) << 23) | (m << 13));
    }
    if (e == 31) return u2f(s | 0x7F800000u | (m << 13));
    return u2f(s | ((e + 112u) << 23) | (m << 13));
}

// binary32 -> binary16, round to nearest even (the fp16 blit of RtRef.cpp:336-377)
HD uint16_t floatToHalf(float f)
{
    const uint32_t x = f2u(f);
    const uint32_t s = (x >> 16) & 0x8000u;
    const uint32_t a = x & 0x7FFFFFFFu;
    if (a >= 0x7F800000u) return (uint16_t)(s | 0x7C00u | ((a > 0x7F800000u) ? 0x200u : 0u));
    if (a >= 0x477FF000u) return (uint16_t)(s | 0x7C00u);
    if (a < 0x33000001u) return (uint16_t)s;
    const int e = (int)(a >> 23) - 127;
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

} // namespace prt
