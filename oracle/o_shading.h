// oracle/o_shading.h -- TEST INFRASTRUCTURE (CPU oracle).
// Restatement of res/shader/common/{random,sampling}.glsl, brdf.glsl and rt/ray.glsl.
// One function per GLSL function, same names, same operation order.
#pragma once
#include "o_math.h"

namespace orc {

// ---- common/random.glsl -------------------------------------------------------------------
inline uint32_t pcg(uint32_t v) // random.glsl:7-12
{
    const uint32_t state = v * 747796405u + 2891336453u;
    const uint32_t word = ((state >> ((state >> 28u) + 4u)) ^ state) * 277803737u;
    return (word >> 22u) ^ word;
}

inline uvec3 pcg3d(uvec3 v) // random.glsl:17-28
{
    v.x = v.x * 1664525u + 1013904223u;
    v.y = v.y * 1664525u + 1013904223u;
    v.z = v.z * 1664525u + 1013904223u;
    v.x += v.y * v.z;
    v.y += v.z * v.x;
    v.z += v.x * v.y;
    v.x ^= v.x >> 16u;
    v.y ^= v.y >> 16u;
    v.z ^= v.z >> 16u;
    v.x += v.y * v.z;
    v.y += v.z * v.x;
    v.z += v.x * v.y;
    return v;
}

// random.glsl:42. float(0xFFFFFFFFu) rounds to 2^32, so this is an exact scaling and 1.0 is
// reachable (u >= 0xFFFFFF80 converts to 2^32).
inline float rngTo01(uint32_t u) { return (float)u / 4294967296.0f; }

struct Rng // the shader-global pcg_state, random.glsl:49-68
{
    uvec3 state;
    float rnd01()
    {
        state = pcg3d(state);
        return rngTo01(state.x);
    }
    vec2 rnd2d01()
    {
        state = pcg3d(state);
        return {rngTo01(state.x), rngTo01(state.y)};
    }
};

// ---- brdf.glsl ----------------------------------------------------------------------------
struct Material // scene/material.glsl:4-16
{
    vec3 albedo;
    vec3 normal;
    float roughness;
    float metallic;
    float alpha;
};

struct VisibleSurface // scene/visible_surface.glsl:6-15
{
    vec3 positionWS;
    vec3 normalWS;
    vec3 invViewRayWS;
    vec2 uv;
    float NoV;
    float linearDepth;
    Material material;
};

inline vec3 lambertBRFD(vec3 c_diff) { return c_diff / PI; } // brdf.glsl:9

inline float trowbridgeReitz(float NoH, float alpha) // brdf.glsl:12-19
{
    const float a2 = alpha * alpha;
    const float denom = NoH * NoH * (a2 - 1.0f) + 1.0f;
    return a2 / (PI * denom * denom);
}

inline vec3 schlickFresnel(float VoH, vec3 f0) // brdf.glsl:21-24
{
    const float p = std::pow(1.0f - VoH, 5.0f);
    return f0 + (vec3{1.f, 1.f, 1.f} - f0) * p;
}

inline float schlickTrowbridgeReitz(float NoL, float NoV, float alpha) // brdf.glsl:35-43
{
    float k = alpha * 0.5f;
    k = std::max(k, 0.0001f);
    const float gl = NoL / (NoL * (1.0f - k) + k);
    const float gv = NoV / (NoV * (1.0f - k) + k);
    return gl * gv;
}

inline vec3 cookTorranceBRDF(float NoL, float NoV, float NoH, float VoH, vec3 f0, float roughness)
{ // brdf.glsl:46-58
    const float alpha = roughness * roughness;
    const float D = trowbridgeReitz(NoH, alpha);
    const vec3 F = schlickFresnel(VoH, f0);
    const float G = schlickTrowbridgeReitz(NoL, NoV, alpha);
    const float denom = 4.0f * NoL * NoV + 0.0001f;
    return F * D * G / denom;
}

inline vec3 fresnelZero(const VisibleSurface &s) // brdf.glsl:60-64
{
    return mix(vec3{0.04f, 0.04f, 0.04f}, s.material.albedo, s.material.metallic);
}

inline vec3 evalBRDFTimesNoL(vec3 l, const VisibleSurface &s) // brdf.glsl:67-87
{
    const vec3 h = normalize(s.invViewRayWS + l);
    const float NoL = saturate(dot(s.normalWS, l));
    const float NoH = saturate(dot(s.normalWS, h));
    const float VoH = saturate(dot(s.invViewRayWS, h));
    const vec3 f0 = fresnelZero(s);
    const vec3 c_diff = mix(s.material.albedo * (1.f - 0.04f), vec3{0.f, 0.f, 0.f}, s.material.metallic);
    return (lambertBRFD(c_diff) + cookTorranceBRDF(NoL, s.NoV, NoH, VoH, f0, s.material.roughness)) * NoL;
}

// ---- common/sampling.glsl -----------------------------------------------------------------
inline vec3 cosineSampleHemisphere(vec3 n, vec2 u) // sampling.glsl:18-33
{
    float a = 1.0f - 2.0f * u.x;
    a *= 0.99999f;
    float b = std::sqrt(1.0f - a * a);
    b *= 0.99999f;
    const float phi = 2.0f * PI * u.y;
    const float x = b * std::cos(phi);
    const float y = b * std::sin(phi);
    const float z = a;
    return normalize(n + vec3{x, y, z});
}

inline float cosineHemispherePdf(float NoL) { return NoL / PI; } // sampling.glsl:35

// sampling.glsl:37-47. Returns the ROWS of the GLSL result: transpose(mat3(b1,b2,n)) has rows
// b1,b2,n as columns transposed, so (M * v) = (dot(b1,v), dot(b2,v), dot(n,v)) and
// (transpose(M) * v) = b1*v.x + b2*v.y + n*v.z.
struct Basis { vec3 b1, b2, n; };
inline Basis orthonormalBasis(vec3 n)
{
    const float s = sign(n.z);
    const float a = -1.f / (s + n.z);
    const float b = n.x * n.y * a;
    const vec3 b1 = {1.f + s * n.x * n.x * a, s * b, -s * n.x};
    const vec3 b2 = {b, s + n.y * n.y * a, -n.y};
    return {b1, b2, n};
}
inline vec3 toBasis(const Basis &m, vec3 v) { return {dot(m.b1, v), dot(m.b2, v), dot(m.n, v)}; }
inline vec3 fromBasis(const Basis &m, vec3 v) { return m.b1 * v.x + m.b2 * v.y + m.n * v.z; }

inline vec3 sampleVisibleTrowbridgeReitz(vec3 Ve, float alpha, vec2 Us) // sampling.glsl:53-79
{
    const float alphaVx = alpha * Ve.x;
    const float alphaVy = alpha * Ve.y;
    const vec3 Vh = normalize(vec3{alphaVx, alphaVy, Ve.z});
    const float lensq = Vh.x * Vh.x + Vh.y * Vh.y;
    const vec3 T1 = lensq > 0 ? vec3{-Vh.y, Vh.x, 0.f} * (1.0f / std::sqrt(lensq)) : vec3{1.f, 0.f, 0.f};
    const vec3 T2 = cross(Vh, T1);
    const float r = std::sqrt(Us.x);
    const float phi = 2.0f * PI * Us.y;
    const float t1 = r * std::cos(phi);
    float t2 = r * std::sin(phi);
    const float s = 0.5f * (1.0f + Vh.z);
    t2 = (1.0f - s) * std::sqrt(1.0f - t1 * t1) + s * t2;
    const vec3 Nh = T1 * t1 + T2 * t2 + Vh * std::sqrt(std::max(0.0f, 1.0f - t1 * t1 - t2 * t2));
    const vec3 Ne = normalize(vec3{alpha * Nh.x, alpha * Nh.y, std::max(0.0f, Nh.z)});
    return reflect(-Ve, Ne);
}

inline float visibleTrowbridgeReitzPdf(vec3 Ve, vec3 Le, float alpha) // sampling.glsl:81-93
{
    const vec3 N = {0.f, 0.f, 1.f};
    const vec3 Ne = normalize(Ve + Le);
    const float NoV = saturate(dot(N, Ve));
    const float NoL = saturate(dot(N, Le));
    const float NoH = saturate(dot(N, Ne));
    const float VNDF = schlickTrowbridgeReitz(NoL, NoV, alpha) * NoV * trowbridgeReitz(NoH, alpha) / Ve.z;
    return VNDF / (4 * NoV);
}

// ---- rt/ray.glsl + scene/camera.glsl --------------------------------------------------------
struct Ray { vec3 o; vec3 d; float tMin; float tMax; }; // ray.glsl:4-10

struct CameraUniforms // ss/scene/camera.h:13-35 (532 B), mat4 column-major m[col][row]
{
    float worldToCamera[4][4];
    float cameraToWorld[4][4];
    float cameraToClip[4][4];
    float clipToWorld[4][4];
    float previousWorldToCamera[4][4];
    float previousCameraToClip[4][4];
    float eye[4];
    float planes[6][4];
    uint32_t resolution[2];
    float currentJitter[2];
    float previousJitter[2];
    float near_, far_, maxViewScale;
};
static_assert(sizeof(CameraUniforms) == 532, "CameraUniforms");

inline vec3 cameraWorldFwd(const CameraUniforms &c) // camera.glsl:46-51
{
    return -vec3{c.worldToCamera[0][2], c.worldToCamera[1][2], c.worldToCamera[2][2]};
}

// ray.glsl:15-43. EXACT contract: association as written below, IEEE div and sqrt.
inline Ray pinholeCameraRay(const CameraUniforms &c, vec2 uv)
{
    const float ndx = uv.x * 2.f - 1.f;
    const float ndy = uv.y * 2.f - 1.f;
    Ray ray;
    ray.o = {c.eye[0], c.eye[1], c.eye[2]};
    ray.tMin = 0.f;
    ray.tMax = INFINITY;
    const float aspect = c.cameraToClip[1][1] / c.cameraToClip[0][0];
    const float tanHalfFovY = 1.0f / c.cameraToClip[1][1];
    const vec3 right = {c.worldToCamera[0][0], c.worldToCamera[1][0], c.worldToCamera[2][0]};
    const vec3 up = {c.worldToCamera[0][1], c.worldToCamera[1][1], c.worldToCamera[2][1]};
    const vec3 fwd = cameraWorldFwd(c);
    ray.d = normalizeExact((((right * ndx) * tanHalfFovY) * aspect + ((up * ndy) * tanHalfFovY)) + fwd);
    return ray;
}

inline Ray thinLensCameraRay(const CameraUniforms &c, vec2 uv, vec2 lensOffset, float apertureDiameter,
                             float focusDistance, float focalLength) // ray.glsl:46-78
{
    const Ray pinholeRay = pinholeCameraRay(c, uv);
    const float theta = lensOffset.x * 2.f * PI;
    const float radius = lensOffset.y;
    const float u = std::cos(theta) * std::sqrt(radius);
    const float v = std::sin(theta) * std::sqrt(radius);
    const vec3 focusPoint = pinholeRay.o + pinholeRay.d * (focusDistance / dot(pinholeRay.d, cameraWorldFwd(c)));
    const float fStop = focalLength / apertureDiameter;
    const float circleOfConfusionRadius = focalLength / (2.f * fStop);
    const vec3 lensPos = vec3{1, 0, 0} * (u * circleOfConfusionRadius) + vec3{0, 1, 0} * (v * circleOfConfusionRadius);
    Ray ray;
    // (camera.cameraToWorld * vec4(lensPos, 1)).xyz, column-major
    const auto &m = c.cameraToWorld;
    ray.o = {m[0][0] * lensPos.x + m[1][0] * lensPos.y + m[2][0] * lensPos.z + m[3][0],
             m[0][1] * lensPos.x + m[1][1] * lensPos.y + m[2][1] * lensPos.z + m[3][1],
             m[0][2] * lensPos.x + m[1][2] * lensPos.y + m[2][2] * lensPos.z + m[3][2]};
    ray.d = normalize(focusPoint - ray.o);
    ray.tMin = 0.f;
    ray.tMax = INFINITY;
    return ray;
}

inline vec3 offsetRay(vec3 p, vec3 n) // ray.glsl:83-103 (Waechter & Binder)
{
    const float origin = 1.0f / 32.0f;
    const float float_scale = 1.0f / 65536.0f;
    const float int_scale = 256.0f;
    const int32_t ofI[3] = {(int32_t)(int_scale * n.x), (int32_t)(int_scale * n.y), (int32_t)(int_scale * n.z)};
    const float pv[3] = {p.x, p.y, p.z};
    const float nv[3] = {n.x, n.y, n.z};
    float out[3];
    for (int i = 0; i < 3; ++i)
    {
        const int32_t bits = (int32_t)floatBits(pv[i]) + ((pv[i] < 0) ? -ofI[i] : ofI[i]);
        const float pI = bitsFloat((uint32_t)bits);
        out[i] = std::fabs(pv[i]) < origin ? pv[i] + float_scale * nv[i] : pI;
    }
    return {out[0], out[1], out[2]};
}

} // namespace orc
