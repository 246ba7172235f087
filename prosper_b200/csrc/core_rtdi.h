// core_rtdi.h -- per-pixel bodies of the ReSTIR DI trace pass: res/shader/rt/direct_illumination/main.rgen
// with restir_di/reservoir.glsl, restir_di/resampling_phat.glsl and the G-buffer / depth helpers of scene/material.glsl:18-48
// and scene/camera.glsl:11-30; plus restir_di/initial_reservoirs.comp, the pass that feeds it its reservoirs.
// The pass reads a G-buffer (RGBA8 albedo+roughness, A2B10G10R10 encoded normal+metallic, D32 non-linear depth: the formats
// of GBufferRenderer.cpp:31-32) and one light reservoir per pixel (R32G32: light index bits, unbiased contribution weight),
// and traces ONE shadow ray per pixel through the same TLAS, any-hit and miss logic as the reference pass.
#pragma once
#include "core_shade.h"

namespace prt {

struct RtdiTracePC // shader_structs/push_constants/restir_di/trace.h
{
    uint32_t drawType;
    uint32_t frameIndex;
    uint32_t flags; // bit 0 skipHistory, bit 1 accumulate (rt/direct_illumination/main.rgen:27-28)
};

struct GBufferView
{
    const uint32_t *albedoRoughness; // R8G8B8A8_UNORM
    const uint32_t *normalMetallic;  // A2B10G10R10_UNORM_PACK32: r, g = encoded normal xy, b = metallic, a = encoded normal z
    const float *nonLinearDepth;     // reverse-Z clip depth, 0 = nothing drawn
    uint32_t width, height;
};

HD V4 unpackUnorm8x4(uint32_t p)
{
    return {(float)(p & 255u) / 255.f, (float)((p >> 8) & 255u) / 255.f, (float)((p >> 16) & 255u) / 255.f, (float)(p >> 24) / 255.f};
}
HD V4 unpackA2B10G10R10(uint32_t p)
{
    return {(float)(p & 1023u) / 1023.f, (float)((p >> 10) & 1023u) / 1023.f, (float)((p >> 20) & 1023u) / 1023.f, (float)(p >> 30) / 3.f};
}
HD uint32_t unormCode(float v, float top) // UNORM conversion: clamp, scale, round to nearest
{
    const float c = fminf_(fmaxf_(v, 0.f), 1.f);
    return (uint32_t)(c * top + 0.5f);
}
HD uint32_t packUnorm8x4(V4 v) { return unormCode(v.x, 255.f) | (unormCode(v.y, 255.f) << 8) | (unormCode(v.z, 255.f) << 16) | (unormCode(v.w, 255.f) << 24); }
HD uint32_t packA2B10G10R10(V4 v)
{
    return unormCode(v.x, 1023.f) | (unormCode(v.y, 1023.f) << 10) | (unormCode(v.z, 1023.f) << 20) | (unormCode(v.w, 3.f) << 30);
}

HD V3 signedOctDecode(V3 n) // scene/material.glsl:20-31
{
    V3 OutN;
    OutN.x = (n.x - n.y);
    OutN.y = (n.x + n.y) - 1.0f;
    OutN.z = n.z * 2.0f - 1.0f;
    OutN.z = OutN.z * (1.0f - fabsf(OutN.x) - fabsf(OutN.y));
    return normalize(OutN);
}
HD V3 signedOctEncode(V3 n) // gbuffer.frag:41-55
{
    V3 OutN;
    n = n / (fabsf(n.x) + fabsf(n.y) + fabsf(n.z));
    OutN.y = n.y * 0.5f + 0.5f;
    OutN.x = n.x * 0.5f + OutN.y;
    OutN.y = n.x * -0.5f + OutN.y;
    const float fltMax = 3.40282e+38f;
    OutN.z = saturate(n.z * fltMax);
    return OutN;
}

HD Material loadFromGbuffer(const GBufferView &gb, uint32_t px, uint32_t py) // scene/material.glsl:33-48
{
    const size_t i = (size_t)py * gb.width + px;
    const V4 albedoRoughness = unpackUnorm8x4(gb.albedoRoughness[i]);
    const V4 normalMetallic = unpackA2B10G10R10(gb.normalMetallic[i]);
    Material m;
    m.albedo = {albedoRoughness.x, albedoRoughness.y, albedoRoughness.z};
    m.roughness = albedoRoughness.w;
    m.normal = signedOctDecode(V3{normalMetallic.x, normalMetallic.y, normalMetallic.w});
    m.metallic = normalMetallic.z;
    m.alpha = -1.0f;
    return m;
}

HD V4 mulMat4(const float m[4][4], V4 v) // glm column-major m[col][row]
{
    return {m[0][0] * v.x + m[1][0] * v.y + m[2][0] * v.z + m[3][0] * v.w, m[0][1] * v.x + m[1][1] * v.y + m[2][1] * v.z + m[3][1] * v.w,
            m[0][2] * v.x + m[1][2] * v.y + m[2][2] * v.z + m[3][2] * v.w, m[0][3] * v.x + m[1][3] * v.y + m[2][3] * v.z + m[3][3] * v.w};
}
HD V3 worldPos(const prt_camera_uniforms &c, V2 uv, float nonLinearDepth) // scene/camera.glsl:24-30
{
    const V4 s = mulMat4(c.clipToWorld, V4{uv.x * 2.f - 1.f, uv.y * 2.f - 1.f, nonLinearDepth, 1.f});
    return V3{s.x, s.y, s.z} / s.w;
}

// The VisibleSurface both passes rebuild from the G-buffer (rt/direct_illumination/main.rgen:117-131,
// initial_reservoirs.comp:70-83); uv = px / resolution as there.
HD VisibleSurface gbufferSurface(const SceneView &sv, const GBufferView &gb, uint32_t px, uint32_t py)
{
    const V2 uv = {(float)px / (float)gb.width, (float)py / (float)gb.height};
    const float nonLinearDepth = gb.nonLinearDepth[(size_t)py * gb.width + px];
    VisibleSurface surface;
    surface.positionWS = worldPos(*sv.camera, uv, nonLinearDepth);
    surface.invViewRayWS = normalize(V3{sv.camera->eye[0], sv.camera->eye[1], sv.camera->eye[2]} - surface.positionWS);
    surface.material = loadFromGbuffer(gb, px, py);
    surface.normalWS = surface.material.normal;
    surface.uv = {0.f, 0.f};
    surface.NoV = saturate(dot(surface.normalWS, surface.invViewRayWS));
    return surface;
}

struct LightReservoir // restir_di/reservoir.glsl:4-9
{
    int32_t lightIndex; // < 0 signals an invalid sample
    float unbiasedContributionWeight;
};
HD float luminance(V3 c) { return dot(V3{0.299f, 0.587f, 0.114f}, c); } // common/math.glsl:15

HD float pHatLight(const SceneView &sv, const VisibleSurface &surface, uint32_t lightIndex) // resampling_phat.glsl:4-19
{
    V3 l, irradiance;
    float d;
    sampleLight(sv, surface, lightIndex, l, d, irradiance);
    const V3 brdf = evalBRDFTimesNoL(l, surface);
    return luminance(irradiance * brdf);
}

HD LightReservoir initialLightCandidate(const SceneView &sv, const VisibleSurface &surface, Rng &rng) // initial_reservoirs.comp:31-60
{
    const int lightCount = 1 + (int)(sv.pointLightCount + sv.spotLightCount);
    LightReservoir reservoir = {-1, 0.f};
    float sumResamplingWeights = 0;
    const uint32_t sampleCount = 5;
    for (uint32_t i = 0; i < sampleCount; ++i)
    {
        int lightIndex = (int)(rng.rnd01() * (float)lightCount);
        if (lightIndex > lightCount - 1) lightIndex = lightCount - 1;
        const float misWeight = 1.f / (float)sampleCount;
        const float unbiasedContributionWeight = (float)lightCount; // one of M lights sampled uniformly: W = 1 / (1 / M)
        const float resamplingWeight = misWeight * pHatLight(sv, surface, (uint32_t)lightIndex) * unbiasedContributionWeight;
        sumResamplingWeights += resamplingWeight;
        if (rng.rnd01() < resamplingWeight / sumResamplingWeights) reservoir.lightIndex = lightIndex;
    }
    if (reservoir.lightIndex >= 0) reservoir.unbiasedContributionWeight = sumResamplingWeights / pHatLight(sv, surface, (uint32_t)reservoir.lightIndex);
    return reservoir;
}

// evaluateDirectLightingReSTIR (rt/direct_illumination/main.rgen:89-110) cut at its traceRayEXT: `contribution` is what the
// pixel gets when the shadow ray towards (l, d) is unoccluded. Returns false when there is nothing to trace.
HD bool restirShade(const SceneView &sv, const VisibleSurface &surface, LightReservoir reservoir, V3 &l, float &d, V3 &contribution)
{
    if (surface.material.alpha == 0 || reservoir.lightIndex < 0) return false;
    V3 irradiance;
    sampleLight(sv, surface, (uint32_t)reservoir.lightIndex, l, d, irradiance);
    if (dot(l, surface.normalWS) <= 0) return false;
    contribution = irradiance * evalBRDFTimesNoL(l, surface) * reservoir.unbiasedContributionWeight;
    return true;
}

// The G-buffer texel a raster pass would have written for a primary hit (gbuffer.frag:57-110), from the tracer's own hit:
// albedo + roughness, encoded shading normal + metallic, and the non-linear (reverse-Z) depth of the hit point.
HD void gbufferTexel(const SceneView &sv, const VisibleSurface &surface, uint32_t &albedoRoughness, uint32_t &normalMetallic, float &depth)
{
    const Material &m = surface.material;
    albedoRoughness = packUnorm8x4(V4{m.albedo.x, m.albedo.y, m.albedo.z, m.roughness});
    const V3 e = signedOctEncode(surface.normalWS);
    normalMetallic = packA2B10G10R10(V4{e.x, e.y, m.metallic, e.z});
    const prt_camera_uniforms &c = *sv.camera;
    const V4 view = mulMat4(c.worldToCamera, V4{surface.positionWS.x, surface.positionWS.y, surface.positionWS.z, 1.f});
    const V4 clip = mulMat4(c.cameraToClip, view);
    depth = clip.z / clip.w;
}

} // namespace prt
