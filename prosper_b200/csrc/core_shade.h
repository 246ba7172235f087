// core_shade.h -- per-path bodies of the wavefront stages: ray generation and material shading.
// Restates res/shader/rt/reference/main.rgen with brdf.glsl, common/{random,sampling}.glsl,
// rt/ray.glsl, scene/{lighting,skybox}.glsl; one function per GLSL function, same names.
#pragma once
#include "core_scene.h"
#include "core_trace.h"

namespace prt {

constexpr float PI = 3.14159265f;      // common/math.glsl:4
constexpr uint32_t MAX_BOUNCES = 6;    // RtRef.hpp:22, RtRef.cpp:410

// ---- common/random.glsl ---------------------------------------------------------------------------
HD U3 pcg3d(U3 v) // random.glsl:17-28
{
    v.x = v.x * 1664525u + 1013904223u;
    v.y = v.y * 1664525u + 1013904223u;
    v.z = v.z * 1664525u + 1013904223u;
    v.x += v.y * v.z; v.y += v.z * v.x; v.z += v.x * v.y;
    v.x ^= v.x >> 16u; v.y ^= v.y >> 16u; v.z ^= v.z >> 16u;
    v.x += v.y * v.z; v.y += v.z * v.x; v.z += v.x * v.y;
    return v;
}
HD float rngTo01(uint32_t u) { return (float)u / 4294967296.0f; } // random.glsl:42: float(0xFFFFFFFFu) == 2^32
struct Rng // the shader-global pcg_state (random.glsl:49-68)
{
    U3 state;
    HD float rnd01() { state = pcg3d(state); return rngTo01(state.x); }
    HD V2 rnd2d01() { state = pcg3d(state); return {rngTo01(state.x), rngTo01(state.y)}; }
};

// ---- rt/ray.glsl + scene/camera.glsl -----------------------------------------------------------------
struct Ray { V3 o; V3 d; float tMin; float tMax; }; // ray.glsl:4-10

HD V3 cameraWorldFwd(const prt_camera_uniforms &c) // camera.glsl:46-51
{
    return {-c.worldToCamera[0][2], -c.worldToCamera[1][2], -c.worldToCamera[2][2]};
}

// ray.glsl:15-43. EXACT CONTRACT (primary rays decide the bit-exact hit IDs): oracle association.
HD Ray pinholeCameraRay(const prt_camera_uniforms &c, V2 uv)
{
    const float ndx = xsub(xmul(uv.x, 2.f), 1.f);
    const float ndy = xsub(xmul(uv.y, 2.f), 1.f);
    Ray ray;
    ray.o = {c.eye[0], c.eye[1], c.eye[2]};
    ray.tMin = 0.f;
    ray.tMax = INFINITY;
    const float aspect = xdiv(c.cameraToClip[1][1], c.cameraToClip[0][0]);
    const float tanHalfFovY = xdiv(1.0f, c.cameraToClip[1][1]);
    const V3 right = {c.worldToCamera[0][0], c.worldToCamera[1][0], c.worldToCamera[2][0]};
    const V3 up = {c.worldToCamera[0][1], c.worldToCamera[1][1], c.worldToCamera[2][1]};
    const V3 fwd = cameraWorldFwd(c);
    auto term = [&](float r, float u, float f) {
        const float a = xmul(xmul(xmul(r, ndx), tanHalfFovY), aspect);
        const float b = xmul(xmul(u, ndy), tanHalfFovY);
        return xadd(xadd(a, b), f);
    };
    ray.d = xnormalize(V3{term(right.x, up.x, fwd.x), term(right.y, up.y, fwd.y), term(right.z, up.z, fwd.z)});
    return ray;
}

HD Ray thinLensCameraRay(const prt_camera_uniforms &c, V2 uv, V2 lensOffset, float apertureDiameter, float focusDistance,
                         float focalLength) // ray.glsl:46-78
{
    const Ray pinholeRay = pinholeCameraRay(c, uv);
    const float theta = lensOffset.x * 2.f * PI;
    const float radius = lensOffset.y;
    const float u = cosf(theta) * sqrtf(radius);
    const float v = sinf(theta) * sqrtf(radius);
    const V3 focusPoint = pinholeRay.o + pinholeRay.d * (focusDistance / dot(pinholeRay.d, cameraWorldFwd(c)));
    const float fStop = focalLength / apertureDiameter;
    const float circleOfConfusionRadius = focalLength / (2.f * fStop);
    const V3 lensPos = {u * circleOfConfusionRadius, v * circleOfConfusionRadius, 0.f};
    Ray ray;
    const float(*m)[4] = c.cameraToWorld;
    ray.o = {m[0][0] * lensPos.x + m[1][0] * lensPos.y + m[2][0] * lensPos.z + m[3][0],
             m[0][1] * lensPos.x + m[1][1] * lensPos.y + m[2][1] * lensPos.z + m[3][1],
             m[0][2] * lensPos.x + m[1][2] * lensPos.y + m[2][2] * lensPos.z + m[3][2]};
    ray.d = normalize(focusPoint - ray.o);
    ray.tMin = 0.f;
    ray.tMax = INFINITY;
    return ray;
}

HD V3 offsetRay(V3 p, V3 n) // ray.glsl:83-103 (Waechter & Binder)
{
    const float origin = 1.0f / 32.0f;
    const float float_scale = 1.0f / 65536.0f;
    const float int_scale = 256.0f;
    const float pv[3] = {p.x, p.y, p.z};
    const float nv[3] = {n.x, n.y, n.z};
    float out[3];
    for (int i = 0; i < 3; ++i)
    {
        const int32_t ofI = (int32_t)(int_scale * nv[i]);
        const int32_t bits = (int32_t)f2u(pv[i]) + ((pv[i] < 0) ? -ofI : ofI);
        const float pI = u2f((uint32_t)bits);
        out[i] = fabsf(pv[i]) < origin ? pv[i] + float_scale * nv[i] : pI;
    }
    return {out[0], out[1], out[2]};
}

// ---- brdf.glsl -----------------------------------------------------------------------------------------
struct VisibleSurface // scene/visible_surface.glsl:6-15
{
    V3 positionWS;
    V3 normalWS;
    V3 invViewRayWS;
    V2 uv;
    float NoV;
    Material material;
};

HD V3 lambertBRFD(V3 c_diff) { return c_diff / PI; } // brdf.glsl:9
HD float trowbridgeReitz(float NoH, float alpha) // brdf.glsl:12-19
{
    const float a2 = alpha * alpha;
    const float denom = NoH * NoH * (a2 - 1.0f) + 1.0f;
    return a2 / (PI * denom * denom);
}
HD V3 schlickFresnel(float VoH, V3 f0) // brdf.glsl:21-24
{
    const float p = powf(1.0f - VoH, 5.0f);
    return f0 + (V3{1.f, 1.f, 1.f} - f0) * p;
}
HD float schlickTrowbridgeReitz(float NoL, float NoV, float alpha) // brdf.glsl:35-43
{
    float k = alpha * 0.5f;
    k = fmaxf_(k, 0.0001f);
    const float gl = NoL / (NoL * (1.0f - k) + k);
    const float gv = NoV / (NoV * (1.0f - k) + k);
    return gl * gv;
}
HD V3 cookTorranceBRDF(float NoL, float NoV, float NoH, float VoH, V3 f0, float roughness) // brdf.glsl:46-58
{
    const float alpha = roughness * roughness;
    const float D = trowbridgeReitz(NoH, alpha);
    const V3 F = schlickFresnel(VoH, f0);
    const float G = schlickTrowbridgeReitz(NoL, NoV, alpha);
    const float denom = 4.0f * NoL * NoV + 0.0001f;
    return F * D * G / denom;
}
HD V3 fresnelZero(const VisibleSurface &s) // brdf.glsl:60-64
{
    return mix3(V3{0.04f, 0.04f, 0.04f}, s.material.albedo, s.material.metallic);
}
HD V3 evalBRDFTimesNoL(V3 l, const VisibleSurface &s) // brdf.glsl:67-87
{
    const V3 h = normalize(s.invViewRayWS + l);
    const float NoL = saturate(dot(s.normalWS, l));
    const float NoH = saturate(dot(s.normalWS, h));
    const float VoH = saturate(dot(s.invViewRayWS, h));
    const V3 f0 = fresnelZero(s);
    const V3 c_diff = mix3(s.material.albedo * (1.f - 0.04f), V3{0.f, 0.f, 0.f}, s.material.metallic);
    return (lambertBRFD(c_diff) + cookTorranceBRDF(NoL, s.NoV, NoH, VoH, f0, s.material.roughness)) * NoL;
}

// ---- common/sampling.glsl --------------------------------------------------------------------------------
HD V3 cosineSampleHemisphere(V3 n, V2 u) // sampling.glsl:18-33
{
    float a = 1.0f - 2.0f * u.x;
    a *= 0.99999f;
    float b = sqrtf(1.0f - a * a);
    b *= 0.99999f;
    const float phi = 2.0f * PI * u.y;
    return normalize(n + V3{b * cosf(phi), b * sinf(phi), a});
}
HD float cosineHemispherePdf(float NoL) { return NoL / PI; } // sampling.glsl:35

struct Basis { V3 b1, b2, n; }; // rows of the GLSL transpose(mat3(b1, b2, n)), sampling.glsl:37-47
HD Basis orthonormalBasis(V3 n)
{
    const float s = n.z > 0.f ? 1.f : (n.z < 0.f ? -1.f : 0.f); // GLSL sign()
    const float a = -1.f / (s + n.z);
    const float b = n.x * n.y * a;
    return {V3{1.f + s * n.x * n.x * a, s * b, -s * n.x}, V3{b, s + n.y * n.y * a, -n.y}, n};
}
HD V3 toBasis(const Basis &m, V3 v) { return {dot(m.b1, v), dot(m.b2, v), dot(m.n, v)}; }
HD V3 fromBasis(const Basis &m, V3 v) { return m.b1 * v.x + m.b2 * v.y + m.n * v.z; }

HD V3 sampleVisibleTrowbridgeReitz(V3 Ve, float alpha, V2 Us) // sampling.glsl:53-79 (Heitz 2018)
{
    const V3 Vh = normalize(V3{alpha * Ve.x, alpha * Ve.y, Ve.z});
    const float lensq = Vh.x * Vh.x + Vh.y * Vh.y;
    const V3 T1 = lensq > 0 ? V3{-Vh.y, Vh.x, 0.f} * (1.0f / sqrtf(lensq)) : V3{1.f, 0.f, 0.f};
    const V3 T2 = cross(Vh, T1);
    const float r = sqrtf(Us.x);
    const float phi = 2.0f * PI * Us.y;
    const float t1 = r * cosf(phi);
    float t2 = r * sinf(phi);
    const float s = 0.5f * (1.0f + Vh.z);
    t2 = (1.0f - s) * sqrtf(1.0f - t1 * t1) + s * t2;
    const V3 Nh = T1 * t1 + T2 * t2 + Vh * sqrtf(fmaxf_(0.0f, 1.0f - t1 * t1 - t2 * t2));
    const V3 Ne = normalize(V3{alpha * Nh.x, alpha * Nh.y, fmaxf_(0.0f, Nh.z)});
    return reflect3(-Ve, Ne);
}
HD float visibleTrowbridgeReitzPdf(V3 Ve, V3 Le, float alpha) // sampling.glsl:81-93
{
    const V3 Ne = normalize(Ve + Le);
    const float NoV = saturate(Ve.z);
    const float NoL = saturate(Le.z);
    const float NoH = saturate(Ne.z);
    const float VNDF = schlickTrowbridgeReitz(NoL, NoV, alpha) * NoV * trowbridgeReitz(NoH, alpha) / Ve.z;
    return VNDF / (4 * NoV);
}

// ---- scene/skybox.glsl:4 + main.rgen:249-253: textureLod(skybox, d, 0).rgb --------------------------------
HD V3 sampleSky(const SceneView &sv, V3 d)
{
    const uint32_t n = sv.skyFaceSize;
    if (n == 0 || sv.sky == nullptr) return {0.f, 0.f, 0.f};
    // NaN direction: undefined lookup in the reference; contract: black
    if (!(fabsf(d.x) <= 3.4e38f) || !(fabsf(d.y) <= 3.4e38f) || !(fabsf(d.z) <= 3.4e38f)) return {0.f, 0.f, 0.f};
    const float ax = fabsf(d.x), ay = fabsf(d.y), az = fabsf(d.z);
    uint32_t face;
    float sc, tc, ma;
    if (ax >= ay && ax >= az)
    {
        ma = ax;
        if (d.x >= 0) { face = 0; sc = -d.z; tc = -d.y; }
        else { face = 1; sc = d.z; tc = -d.y; }
    }
    else if (ay >= az)
    {
        ma = ay;
        if (d.y >= 0) { face = 2; sc = d.x; tc = d.z; }
        else { face = 3; sc = d.x; tc = -d.z; }
    }
    else
    {
        ma = az;
        if (d.z >= 0) { face = 4; sc = d.x; tc = -d.y; }
        else { face = 5; sc = -d.x; tc = -d.y; }
    }
    const float s = 0.5f * (sc / ma + 1.f), t = 0.5f * (tc / ma + 1.f);
    const uint16_t *px = sv.sky + (size_t)face * n * n * 4;
    const float fx = s * n - 0.5f, fy = t * n - 0.5f;
    const float x0f = floorf(fx), y0f = floorf(fy);
    const float wx = fx - x0f, wy = fy - y0f;
    const int hi = (int)n - 1;
    auto cl = [&](int i) { return i < 0 ? 0 : (i > hi ? hi : i); };
    const int x0 = cl((int)x0f), x1 = cl((int)x0f + 1), y0 = cl((int)y0f), y1 = cl((int)y0f + 1);
    auto ld = [&](int x, int y) {
        const uint16_t *p = px + ((size_t)y * n + x) * 4;
        return V3{halfToFloat(p[0]), halfToFloat(p[1]), halfToFloat(p[2])};
    };
    return (ld(x0, y0) * (1.f - wx) + ld(x1, y0) * wx) * (1.f - wy) + (ld(x0, y1) * (1.f - wx) + ld(x1, y1) * wx) * wy;
}

// ---- scene/lighting.glsl:15-89 ---------------------------------------------------------------------------
HD void sampleLight(const SceneView &sv, const VisibleSurface &surface, uint32_t lightIndex, V3 &l, float &d, V3 &irradiance)
{
    if (lightIndex == 0)
    {
        const prt_directional_light &dl = *sv.directional;
        l = -normalize(V3{dl.direction[0], dl.direction[1], dl.direction[2]});
        d = 100.f;
        irradiance = {dl.irradiance[0], dl.irradiance[1], dl.irradiance[2]};
        return;
    }
    lightIndex -= 1;
    if (lightIndex < sv.pointLightCount)
    {
        const prt_point_light &light = sv.pointLights[lightIndex];
        const V3 pos = {light.position[0], light.position[1], light.position[2]};
        const V3 radiance = {light.radianceAndRadius[0], light.radianceAndRadius[1], light.radianceAndRadius[2]};
        const float radius = light.radianceAndRadius[3];
        const V3 toLight = pos - surface.positionWS;
        const float d2 = dot(toLight, toLight);
        d = sqrtf(d2);
        l = toLight / d;
        const float dPerR = d / radius;
        const float dPerR2 = dPerR * dPerR;
        const float dPerR4 = dPerR2 * dPerR2;
        const float radialAttenuation = fmaxf_(fminf_(1.0f - dPerR4, 1.f), 0.f);
        irradiance = radiance * radialAttenuation / d2;
        return;
    }
    lightIndex -= sv.pointLightCount;
    if (lightIndex < sv.spotLightCount)
    {
        const prt_spot_light &light = sv.spotLights[lightIndex];
        const V3 toLight = V3{light.positionAndAngleOffset[0], light.positionAndAngleOffset[1], light.positionAndAngleOffset[2]} -
                           surface.positionWS;
        const float d2 = dot(toLight, toLight);
        d = sqrtf(d2);
        l = toLight / d;
        const float cd = dot(-V3{light.direction[0], light.direction[1], light.direction[2]}, l);
        float angularAttenuation = saturate(cd * light.radianceAndAngleScale[3] + light.positionAndAngleOffset[3]);
        angularAttenuation *= angularAttenuation;
        irradiance = V3{light.radianceAndAngleScale[0], light.radianceAndAngleScale[1], light.radianceAndAngleScale[2]} *
                     angularAttenuation / d2;
        return;
    }
    l = {0.f, 1.f, 0.f};
    d = 1.f;
    irradiance = {0.f, 0.f, 0.f};
}

// ---- main.rgen ------------------------------------------------------------------------------------------------
HD bool flagSkipHistory(const prt_reference_pc &pc) { return (pc.flags >> 0) & 1u; }
HD bool flagAccumulate(const prt_reference_pc &pc) { return (pc.flags >> 1) & 1u; }
HD bool flagIBL(const prt_reference_pc &pc) { return (pc.flags >> 2) & 1u; }
HD bool flagDepthOfField(const prt_reference_pc &pc) { return (pc.flags >> 3) & 1u; }
HD bool flagClampIndirect(const prt_reference_pc &pc) { return (pc.flags >> 4) & 1u; }

HD V3 mappedNormal(V3 tangentSpaceNormal, V3 normal, V3 tangent, float sgn) // main.rgen:37-45
{
    const V3 vB = cross(normal, tangent) * sgn;
    return normalize(tangent * tangentSpaceNormal.x + vB * tangentSpaceNormal.y + normal * tangentSpaceNormal.z);
}

HD void addBounce(const prt_reference_pc &PC, V3 &accColor, V3 color, uint32_t bounce) // main.rgen:83-88
{
    if (bounce > 0 && flagClampIndirect(PC)) color = clamp3(color, 0.f, 2.f);
    accColor = accColor + color;
}

// main.rgen:225-240: RNG seed, jittered uv, camera ray of pixel (px, py) for frame `frameIndex`.
HD Ray generatePrimaryRay(const SceneView &sv, const prt_reference_pc &PC, uint32_t frameIndex, uint32_t px, uint32_t py,
                          uint32_t width, uint32_t height, Rng &rng)
{
    rng.state = {px, py, frameIndex};
    const V2 r = rng.rnd2d01();
    const V2 jitteredUV = {xdiv(xadd((float)px, r.x), (float)width), xdiv(xadd((float)py, r.y), (float)height)};
    if (flagDepthOfField(PC))
    {
        const V2 lens = rng.rnd2d01();
        return thinLensCameraRay(*sv.camera, jitteredUV, lens, PC.apertureDiameter, PC.focusDistance, PC.focalLength);
    }
    return pinholeCameraRay(*sv.camera, jitteredUV);
}

// evaluateSurface, main.rgen:146-179, in two halves so that the shade kernel can fetch the material's
// textures between them at a warp-convergent point (kernels_render.cuh).
struct SurfaceGeom
{
    Vertex v; // interpolated, in world space
    uint32_t materialIndex;
};

HD SurfaceGeom surfaceGeometry(const SceneView &sv, uint32_t drawInstanceIndex, uint32_t primitiveID, float bx, float by)
{
    const prt_draw_instance &instance = sv.drawInstances[drawInstanceIndex];
    const prt_model_instance_transforms &trfn = sv.transforms[instance.modelInstanceIndex];
    const prt_geometry_metadata &metadata = sv.metadatas[instance.meshIndex];
    const Vertex v0 = loadVertexThroughIndexBuffer(sv, metadata, primitiveID * 3 + 0);
    const Vertex v1 = loadVertexThroughIndexBuffer(sv, metadata, primitiveID * 3 + 1);
    const Vertex v2 = loadVertexThroughIndexBuffer(sv, metadata, primitiveID * 3 + 2);
    SurfaceGeom g;
    g.v = transformVertex(interpolateVertex(v0, v1, v2, bx, by), trfn);
    g.materialIndex = instance.materialIndex;
    return g;
}

HD VisibleSurface evaluateSurface(const SceneView &sv, V3 rayD, const SurfaceGeom &g, const MaterialTaps &taps)
{
    const Vertex &v = g.v;
    VisibleSurface surface;
    surface.positionWS = v.Position;
    surface.invViewRayWS = -rayD;
    surface.uv = v.TexCoord0;
    surface.material = sampleMaterial(sv, g.materialIndex, taps);
    if (surface.material.normal.x != -2 && v.Tangent.w != 0)
        surface.normalWS = mappedNormal(surface.material.normal, v.Normal, V3{v.Tangent.x, v.Tangent.y, v.Tangent.z}, v.Tangent.w);
    else
        surface.normalWS = normalize(v.Normal);
    surface.NoV = saturate(dot(surface.normalWS, surface.invViewRayWS));
    return surface;
}

HD V3 uintToColor(uint32_t x) // random.glsl:30-40
{
    const uint32_t xr = pcg(x);
    return V3{(float)((xr >> 20) & 0x3FF), (float)((xr >> 10) & 0x3FF), (float)(xr & 0x3FF)} / (float)0x3FF;
}

HD V3 debugColor(const SceneView &sv, uint32_t drawType, uint32_t drawInstanceIndex, uint32_t primitiveID,
                 const VisibleSurface &surface) // main.rgen:181-193; debug.glsl:17-38
{
    const prt_draw_instance &instance = sv.drawInstances[drawInstanceIndex];
    const Material &m = surface.material;
    switch (drawType)
    {
    case PRT_DRAW_PRIMITIVE_ID: return uintToColor(primitiveID);
    case PRT_DRAW_MESH_ID: return uintToColor(instance.meshIndex);
    case PRT_DRAW_MATERIAL_ID: return uintToColor(instance.materialIndex);
    case PRT_DRAW_POSITION: return surface.positionWS;
    case PRT_DRAW_TEXCOORD0: return {surface.uv.x, surface.uv.y, 0.f};
    case PRT_DRAW_ALBEDO: return m.albedo;
    case PRT_DRAW_SHADING_NORMAL: return surface.normalWS * 0.5f + V3{0.5f, 0.5f, 0.5f};
    case PRT_DRAW_ROUGHNESS: return {m.roughness, m.roughness, m.roughness};
    case PRT_DRAW_METALLIC: return {m.metallic, m.metallic, m.metallic};
    default: return {1.f, 0.f, 1.f};
    }
}

HD void importanceSampleBounce(Rng &rng, const VisibleSurface &surface, V3 &throughput, V3 &rd) // main.rgen:90-144
{
    const bool specularOnly = surface.material.metallic > 0.999f;
    const float specularWeight = specularOnly ? 1.f : 0.5f;
    const float diffuseWeight = 1.f - specularWeight;
    const Basis normalBasis = orthonormalBasis(surface.normalWS);
    const V3 vInBasis = toBasis(normalBasis, surface.invViewRayWS);
    const float alpha = surface.material.roughness * surface.material.roughness;
    const bool pickDiffuse = rng.rnd01() < diffuseWeight;
    V3 brdf;
    float NoL, pdf;
    if (pickDiffuse)
    {
        rd = cosineSampleHemisphere(surface.normalWS, rng.rnd2d01());
        NoL = saturate(dot(surface.normalWS, rd));
        brdf = lambertBRFD(surface.material.albedo);
        pdf = cosineHemispherePdf(NoL) * diffuseWeight;
    }
    else
    {
        rd = fromBasis(normalBasis, sampleVisibleTrowbridgeReitz(vInBasis, alpha, rng.rnd2d01()));
        NoL = saturate(dot(surface.normalWS, rd));
        const V3 h = normalize(surface.invViewRayWS + rd);
        const float NoH = saturate(dot(surface.normalWS, h));
        const float VoH = saturate(dot(surface.invViewRayWS, h));
        brdf = cookTorranceBRDF(NoL, surface.NoV, NoH, VoH, fresnelZero(surface), surface.material.roughness);
        pdf = visibleTrowbridgeReitzPdf(vInBasis, toBasis(normalBasis, rd), alpha) * specularWeight;
    }
    // max(x, 0) with NaN x -> 0 (fmaxf), DESIGN.md section 4
    const V3 w = brdf * NoL / pdf;
    throughput = throughput * V3{fmaxf_(w.x, 0.f), fmaxf_(w.y, 0.f), fmaxf_(w.z, 0.f)};
}

// What one pass of the shade stage decides for a path (one iteration of the loop main.rgen:241-283,
// with the two traceRayEXT calls cut out into the trace stages).
struct ShadeResult
{
    bool shadowRay;     // evaluateDirectLighting wants a visibility test
    V3 shadowO, shadowD;
    float shadowTMax;
    uint32_t shadowSeed;
    V3 shadowContribution; // added to the path colour if the shadow ray misses (already clamped)
    bool alive;            // path continues with `next`
    Ray next;
};

// `geom` / `taps` are only read when the ray hit something (hitDrawInstance != MISS_INDEX).
HDNI void shadePath(const SceneView &sv, const prt_reference_pc &PC, uint32_t bounce, V3 rayD, uint32_t hitDrawInstance,
                    uint32_t hitPrimitive, const SurfaceGeom &geom, const MaterialTaps &taps, Rng &rng, V3 &throughput, V3 &color,
                    ShadeResult &out)
{
    out.shadowRay = false;
    out.alive = false;
    if (hitDrawInstance == MISS_INDEX)
    {
        if (flagIBL(PC)) addBounce(PC, color, throughput * sampleSky(sv, rayD), bounce); // main.rgen:247-255
        return;
    }
    const VisibleSurface surface = evaluateSurface(sv, rayD, geom, taps);
    if (PC.drawType != PRT_DRAW_DEFAULT && PC.drawType != PRT_DRAW_MESHLET_ID)
    {
        color = debugColor(sv, PC.drawType, hitDrawInstance, hitPrimitive, surface); // main.rgen:259-264
        return;
    }
    // evaluateDirectLighting, main.rgen:195-223
    if (surface.material.alpha != 0)
    {
        const uint32_t lightCount = 1 + sv.pointLightCount + sv.spotLightCount;
        uint32_t lightIndex = (uint32_t)(rng.rnd01() * (float)lightCount);
        if (lightIndex > lightCount - 1) lightIndex = lightCount - 1;
        V3 l, irradiance;
        float d;
        sampleLight(sv, surface, lightIndex, l, d, irradiance);
        if (dot(l, surface.normalWS) > 0)
        {
            irradiance = irradiance * (float)lightCount;
            V3 c = throughput * irradiance * evalBRDFTimesNoL(l, surface);
            if (bounce > 0 && flagClampIndirect(PC)) c = clamp3(c, 0.f, 2.f);
            out.shadowRay = true;
            out.shadowO = surface.positionWS;
            out.shadowD = l;
            out.shadowTMax = d;
            out.shadowSeed = pcg(rng.state.x ^ rng.state.y); // main.rgen:55
            out.shadowContribution = c;
        }
    }
    V3 rd;
    importanceSampleBounce(rng, surface, throughput, rd);
    if (bounce > PC.rouletteStartBounce) // main.rgen:271-276 (no 1/p compensation)
    {
        if (rng.rnd01() < fmaxf_(.05f, 1.f - max3(throughput))) return;
    }
    if (bounce + 1 >= MAX_BOUNCES || bounce + 1 >= PC.maxBounces) return;
    out.alive = true;
    out.next.o = offsetRay(surface.positionWS, surface.normalWS);
    out.next.d = rd;
    out.next.tMin = 0;
    out.next.tMax = INFINITY;
}

} // namespace prt
