// oracle/oracle.cpp -- TEST INFRASTRUCTURE (CPU oracle, liboracle.so). See oracle.h.
//
// CPU restatement of prosper's RtReference path. The reference has no CPU implementation of
// this path (GLSL + Vulkan driver BVH), so this file restates res/shader/rt/reference/main.rgen
// and what it includes function by function (citations on each function), on top of the
// oracle's own BVH (o_bvh.h). PARITY UNPINNED by the reference's own tests (it has none for
// this path); the pins are tests/golden/kat.json + analytic properties +
// brute-force-vs-BVH agreement (tests/test_oracle.py).
//
// Build: oracle/Makefile (g++ -O2 -ffp-contract=off). Never linked into the product.
#include "o_bvh.h"
#include "o_scene.h"
#include "o_shading.h"
#include "oracle.h"

#include <atomic>
#include <chrono>
#include <thread>

namespace orc {

constexpr uint32_t MISS_INDEX = 0xFFFFFFFFu; // main.rgen:47
constexpr uint32_t MAX_BOUNCES = 6;          // RtRef.hpp:22, RtRef.cpp:410

enum DrawType : uint32_t // src/scene/DrawType.hpp:8-10
{
    DrawType_Default = 0, DrawType_PrimitiveID, DrawType_MeshletID, DrawType_MeshID, DrawType_MaterialID,
    DrawType_Position, DrawType_ShadingNormal, DrawType_TexCoord0, DrawType_Albedo, DrawType_Roughness,
    DrawType_Metallic
};

// ---- one BLAS per model (World.cpp:585-747) -----------------------------------------------
struct BlasTri
{
    vec3 v0, v1, v2;
    uint32_t geometry; // sub-model index inside the model = gl_GeometryIndexEXT
    uint32_t primitive; // gl_PrimitiveID
    uint32_t opaque;
};
struct Blas
{
    std::vector<BlasTri> tris; // in BVH item order
    Bvh bvh;
};

struct InstanceRec
{
    float w2o[12];
    uint32_t customIndex;
    uint32_t mask;
    uint32_t model;
    uint32_t index; // position in the TLAS instance array (tie-break key)
};

struct Hit
{
    uint32_t drawInstanceIndex = MISS_INDEX;
    uint32_t primitiveID = MISS_INDEX;
    vec2 bary{0.f, 0.f};
    float t = 0.f;
    uint32_t instance = MISS_INDEX; // tie-break key, not part of the payload
};

} // namespace orc

using namespace orc;

struct orc_ctx
{
    Scene scene;
    bool bruteForce = false;
    std::vector<Blas> blases;
    std::vector<InstanceRec> instances;
    Bvh tlas;
    double buildSeconds = 0.0;
};

namespace orc {

// ---- texture fetch: what the Vulkan sampler does for texture() in a stage without derivatives:
// LOD 0, so the magnification filter applies (materials.glsl:44; WorldData.cpp:686-720).
static int wrapCoord(int i, int n, uint32_t mode)
{
    if (mode == 0) // repeat
    {
        i %= n;
        return i < 0 ? i + n : i;
    }
    if (mode == 1) // mirrored repeat
    {
        const int period = 2 * n;
        i %= period;
        if (i < 0) i += period;
        return i < n ? i : period - 1 - i;
    }
    return std::min(std::max(i, 0), n - 1); // clamp to edge
}

static vec4 texel(const std::vector<uint8_t> &img, uint32_t w, int x, int y)
{
    const uint8_t *p = &img[((size_t)y * w + x) * 4];
    return {p[0] / 255.f, p[1] / 255.f, p[2] / 255.f, p[3] / 255.f};
}

static vec4 sampleTexture(const Scene &sc, uint32_t tex, uint32_t smp, vec2 uv)
{
    if (tex >= sc.s.textureCount || sc.decoded[tex].empty()) return {1.f, 1.f, 1.f, 1.f};
    const orc_texture &tx = sc.s.textures[tex];
    SamplerDesc sd{1, 1, 0, 0};
    if (smp < sc.s.samplerCount) sd = ((const SamplerDesc *)sc.s.samplers)[smp];
    const int w = (int)tx.width, h = (int)tx.height;
    const std::vector<uint8_t> &img = sc.decoded[tex];
    if (sd.magFilter == 0) // nearest
    {
        const int x = wrapCoord((int)std::floor(uv.x * w), w, sd.addressU);
        const int y = wrapCoord((int)std::floor(uv.y * h), h, sd.addressV);
        return texel(img, tx.width, x, y);
    }
    const float fx = uv.x * w - 0.5f, fy = uv.y * h - 0.5f;
    const float x0f = std::floor(fx), y0f = std::floor(fy);
    const float ax = fx - x0f, ay = fy - y0f;
    const int x0 = wrapCoord((int)x0f, w, sd.addressU), x1 = wrapCoord((int)x0f + 1, w, sd.addressU);
    const int y0 = wrapCoord((int)y0f, h, sd.addressV), y1 = wrapCoord((int)y0f + 1, h, sd.addressV);
    const vec4 t00 = texel(img, tx.width, x0, y0), t10 = texel(img, tx.width, x1, y0);
    const vec4 t01 = texel(img, tx.width, x0, y1), t11 = texel(img, tx.width, x1, y1);
    return (t00 * (1.f - ax) + t10 * ax) * (1.f - ay) + (t01 * (1.f - ax) + t11 * ax) * ay;
}

// textureLod(skybox, d, 0).rgb (main.rgen:251, skybox.glsl:4): cube face selection of the Vulkan
// spec (major axis table), bilinear inside the face, clamped at the face border.
static vec3 sampleSky(const Scene &sc, vec3 d)
{
    const uint32_t n = sc.s.skyFaceSize;
    if (n == 0 || sc.s.skyRgba16f == nullptr) return {0.f, 0.f, 0.f};
    // A NaN direction (orthonormalBasis() of a normal with n.z == 0, sampling.glsl:41-42) is an
    // undefined cube lookup in the reference; contract: it fetches black.
    if (!std::isfinite(d.x) || !std::isfinite(d.y) || !std::isfinite(d.z)) return {0.f, 0.f, 0.f};
    const float ax = std::fabs(d.x), ay = std::fabs(d.y), az = std::fabs(d.z);
    uint32_t face;
    float sc_, tc, ma;
    if (ax >= ay && ax >= az)
    {
        ma = ax;
        if (d.x >= 0) { face = 0; sc_ = -d.z; tc = -d.y; }
        else { face = 1; sc_ = d.z; tc = -d.y; }
    }
    else if (ay >= az)
    {
        ma = ay;
        if (d.y >= 0) { face = 2; sc_ = d.x; tc = d.z; }
        else { face = 3; sc_ = d.x; tc = -d.z; }
    }
    else
    {
        ma = az;
        if (d.z >= 0) { face = 4; sc_ = d.x; tc = -d.y; }
        else { face = 5; sc_ = -d.x; tc = -d.y; }
    }
    const float s = 0.5f * (sc_ / ma + 1.f), t = 0.5f * (tc / ma + 1.f);
    const uint16_t *px = (const uint16_t *)sc.s.skyRgba16f + (size_t)face * n * n * 4;
    const float fx = s * n - 0.5f, fy = t * n - 0.5f;
    const float x0f = std::floor(fx), y0f = std::floor(fy);
    const float wx = fx - x0f, wy = fy - y0f;
    auto cl = [&](int i) { return std::min(std::max(i, 0), (int)n - 1); };
    const int x0 = cl((int)x0f), x1 = cl((int)x0f + 1), y0 = cl((int)y0f), y1 = cl((int)y0f + 1);
    auto ld = [&](int x, int y) {
        const uint16_t *p = px + ((size_t)y * n + x) * 4;
        return vec3{halfToFloat(p[0]), halfToFloat(p[1]), halfToFloat(p[2])};
    };
    return (ld(x0, y0) * (1.f - wx) + ld(x1, y0) * wx) * (1.f - wy) + (ld(x0, y1) * (1.f - wx) + ld(x1, y1) * wx) * wy;
}

// ---- scene/materials.glsl -------------------------------------------------------------------
static float sRGBtoLinear(float x) // materials.glsl:26-29
{
    return x <= 0.04045f ? x / 12.92f : std::pow((x + 0.055f) / 1.055f, 2.4f);
}

static Material sampleMaterial(const Scene &sc, uint32_t index, vec2 uv) // materials.glsl:47-119
{
    const MaterialData &data = sc.material(index);
    Material ret{};
    vec4 linearBaseColor;
    const uint32_t baseColorTex = data.baseColorTextureSampler & 0xFFFFFFu;
    const uint32_t baseColorSampler = data.baseColorTextureSampler >> 24;
    if (baseColorTex > 0)
    {
        const vec4 s = sampleTexture(sc, baseColorTex, baseColorSampler, uv);
        linearBaseColor = {sRGBtoLinear(s.x), sRGBtoLinear(s.y), sRGBtoLinear(s.z), s.w};
    }
    else
        linearBaseColor = {1.f, 1.f, 1.f, 1.f};
    linearBaseColor = linearBaseColor * vec4{data.baseColorFactor[0], data.baseColorFactor[1], data.baseColorFactor[2],
                                             data.baseColorFactor[3]};
    if (data.alphaMode == AlphaModeBlend)
        ret.alpha = linearBaseColor.w;
    else
    {
        if (data.alphaMode == AlphaModeMask)
        {
            if (linearBaseColor.w < data.alphaCutoff)
            {
                ret.alpha = 0;
                return ret;
            }
        }
        ret.alpha = -1;
    }
    ret.albedo = xyz(linearBaseColor);
    const uint32_t mrTex = data.metallicRoughnessTextureSampler & 0xFFFFFFu;
    const uint32_t mrSampler = data.metallicRoughnessTextureSampler >> 24;
    if (mrTex > 0)
    {
        const vec4 mr = sampleTexture(sc, mrTex, mrSampler, uv);
        ret.roughness = mr.y * data.roughnessFactor;
        ret.metallic = mr.z * data.metallicFactor;
    }
    else
    {
        ret.roughness = data.roughnessFactor;
        ret.metallic = data.metallicFactor;
    }
    ret.roughness = std::max(ret.roughness, 0.05f);
    const uint32_t nTex = data.normalTextureSampler & 0xFFFFFFu;
    const uint32_t nSampler = data.normalTextureSampler >> 24;
    if (nTex > 0)
    {
        const vec4 tn = sampleTexture(sc, nTex, nSampler, uv);
        ret.normal = xyz(tn) * 2.f - vec3{1.f, 1.f, 1.f};
    }
    else
        ret.normal = {-2.f, -2.f, -2.f};
    return ret;
}

// materials.glsl:121-147. NOTE the float overload of sRGBtoLinear IS applied to alpha here
// (unlike sampleMaterial). `margin` reports |alpha - threshold| of the decision for tie masks.
static float sampleAlpha(const Scene &sc, uint32_t index, vec2 uv, float *linearAlphaOut)
{
    const MaterialData &data = sc.material(index);
    float linearAlpha = 1;
    const uint32_t baseColorTex = data.baseColorTextureSampler & 0xFFFFFFu;
    const uint32_t baseColorSampler = data.baseColorTextureSampler >> 24;
    if (baseColorTex > 0) linearAlpha = sRGBtoLinear(sampleTexture(sc, baseColorTex, baseColorSampler, uv).w);
    linearAlpha *= data.baseColorFactor[3];
    if (linearAlphaOut) *linearAlphaOut = linearAlpha;
    if (data.alphaMode == AlphaModeBlend) return linearAlpha;
    if (data.alphaMode == AlphaModeMask)
    {
        if (linearAlpha < data.alphaCutoff) return 0;
    }
    return -1;
}

// ---- rt/scene.rahit:18-39 ---------------------------------------------------------------------
// Returns true when the candidate is accepted. randomSeed is the payload's (one u per RAY).
static bool anyHit(const Scene &sc, uint32_t drawInstanceIndex, uint32_t primitiveID, vec2 bary, uint32_t randomSeed,
                   float *margin)
{
    const DrawInstance &instance = sc.drawInstance(drawInstanceIndex);
    const vec2 uv0 = loadUV(sc, instance.meshIndex, primitiveID * 3 + 0);
    const vec2 uv1 = loadUV(sc, instance.meshIndex, primitiveID * 3 + 1);
    const vec2 uv2 = loadUV(sc, instance.meshIndex, primitiveID * 3 + 2);
    const vec2 uv = interpolate(uv0, uv1, uv2, bary);
    float linearAlpha = 0.f;
    const float alpha = sampleAlpha(sc, instance.materialIndex, uv, &linearAlpha);
    const MaterialData &md = sc.material(instance.materialIndex);
    if (margin)
    {
        float m = 1e30f;
        if (md.alphaMode == AlphaModeMask) m = std::fabs(linearAlpha - md.alphaCutoff);
        else if (md.alphaMode == AlphaModeBlend)
            m = std::min(std::fabs(linearAlpha - (float)pcg(randomSeed) / 4294967296.0f), std::fabs(linearAlpha));
        *margin = std::min(*margin, m);
    }
    if (alpha == 0) return false;
    if (alpha > 0)
    {
        const float u = (float)pcg(randomSeed) / 4294967296.0f; // float(0xFFFFFFFFu) == 2^32
        if (u > alpha) return false;
    }
    return true;
}

// ---- traversal: traceRayEXT (main.rgen:57, 73) --------------------------------------------------
struct TraceArgs
{
    vec3 o, d;
    float tMin, tMax;
    uint32_t seed;
    bool terminateOnFirstHit;
    float *margin;
};

static inline bool better(float t, uint32_t inst, uint32_t geom, uint32_t prim, const Hit &h, uint32_t hGeom)
{
    if (h.drawInstanceIndex == MISS_INDEX) return true;
    if (t != h.t) return t < h.t;
    if (inst != h.instance) return inst < h.instance;
    if (geom != hGeom) return geom < hGeom;
    return prim < h.primitiveID;
}

static void testTriangle(const orc_ctx &ctx, const InstanceRec &inst, const BlasTri &tri, vec3 oo, vec3 od,
                         const TraceArgs &a, Hit &best, uint32_t &bestGeom, bool &done)
{
    float t, u, v;
    const float tMax = a.tMax;
    if (!intersectTriangle(oo, od, a.tMin, tMax, tri.v0, tri.v1, tri.v2, t, u, v)) return;
    // closest-so-far filter with the deterministic tie rule
    if (!better(t, inst.index, tri.geometry, tri.primitive, best, bestGeom)) return;
    const uint32_t drawInstanceIndex = inst.customIndex + tri.geometry;
    if (!tri.opaque && !anyHit(ctx.scene, drawInstanceIndex, tri.primitive, {u, v}, a.seed, a.margin)) return;
    best.drawInstanceIndex = drawInstanceIndex;
    best.primitiveID = tri.primitive;
    best.bary = {u, v};
    best.t = t;
    best.instance = inst.index;
    bestGeom = tri.geometry;
    if (a.terminateOnFirstHit) done = true;
}

static void traverseBlas(const orc_ctx &ctx, const InstanceRec &inst, const TraceArgs &a, Hit &best, uint32_t &bestGeom,
                         bool &done)
{
    const Blas &blas = ctx.blases[inst.model];
    if (blas.tris.empty()) return;
    const vec3 oo = transformPoint(inst.w2o, a.o);
    const vec3 od = transformVector(inst.w2o, a.d);
    if (ctx.bruteForce)
    {
        for (const BlasTri &tri : blas.tris)
        {
            testTriangle(ctx, inst, tri, oo, od, a, best, bestGeom, done);
            if (done) return;
        }
        return;
    }
    const vec3 invD = {1.f / od.x, 1.f / od.y, 1.f / od.z};
    uint32_t stack[128];
    int sp = 0;
    stack[sp++] = 0;
    while (sp > 0)
    {
        const BvhNode &n = blas.bvh.nodes[stack[--sp]];
        float tn;
        const float tFar = best.drawInstanceIndex == MISS_INDEX ? a.tMax : std::min(a.tMax, best.t);
        if (!slab(n.box, oo, invD, a.tMin, tFar, tn)) continue;
        if (n.count)
        {
            for (uint32_t i = 0; i < n.count; ++i)
            {
                testTriangle(ctx, inst, blas.tris[n.left + i], oo, od, a, best, bestGeom, done);
                if (done) return;
            }
        }
        else
        {
            stack[sp++] = n.left;
            stack[sp++] = n.left + 1;
        }
    }
}

static Hit trace(const orc_ctx &ctx, const TraceArgs &a)
{
    Hit best;
    uint32_t bestGeom = 0;
    bool done = false;
    if (ctx.bruteForce)
    {
        for (const InstanceRec &inst : ctx.instances)
        {
            if ((inst.mask & 0xFFu) == 0) continue;
            traverseBlas(ctx, inst, a, best, bestGeom, done);
            if (done) break;
        }
        return best;
    }
    if (ctx.tlas.nodes.empty()) return best;
    const vec3 invD = {1.f / a.d.x, 1.f / a.d.y, 1.f / a.d.z};
    uint32_t stack[128];
    int sp = 0;
    stack[sp++] = 0;
    while (sp > 0 && !done)
    {
        const BvhNode &n = ctx.tlas.nodes[stack[--sp]];
        float tn;
        const float tFar = best.drawInstanceIndex == MISS_INDEX ? a.tMax : std::min(a.tMax, best.t);
        if (!slab(n.box, a.o, invD, a.tMin, tFar, tn)) continue;
        if (n.count)
        {
            for (uint32_t i = 0; i < n.count && !done; ++i)
            {
                const InstanceRec &inst = ctx.instances[ctx.tlas.items[n.left + i]];
                if ((inst.mask & 0xFFu) == 0) continue;
                traverseBlas(ctx, inst, a, best, bestGeom, done);
            }
        }
        else
        {
            stack[sp++] = n.left;
            stack[sp++] = n.left + 1;
        }
    }
    return best;
}

// ---- main.rgen ------------------------------------------------------------------------------
struct PathCtx
{
    const orc_ctx &ctx;
    const ReferencePC &PC;
    Rng rng;
    uint64_t closestRays = 0, shadowRays = 0, shadedHits = 0;
};

static bool flagSkipHistory(const ReferencePC &pc) { return (pc.flags >> 0) & 1u; }
static bool flagAccumulate(const ReferencePC &pc) { return (pc.flags >> 1) & 1u; }
static bool flagIBL(const ReferencePC &pc) { return (pc.flags >> 2) & 1u; }
static bool flagDepthOfField(const ReferencePC &pc) { return (pc.flags >> 3) & 1u; }
static bool flagClampIndirect(const ReferencePC &pc) { return (pc.flags >> 4) & 1u; }

static vec3 mappedNormal(vec3 tangentSpaceNormal, vec3 normal, vec3 tangent, float sgn) // main.rgen:37-45
{
    const vec3 vB = cross(normal, tangent) * sgn;
    return normalize(tangent * tangentSpaceNormal.x + vB * tangentSpaceNormal.y + normal * tangentSpaceNormal.z);
}

static float shadow(PathCtx &pc, vec3 p, vec3 l, float tMin, float lDist) // main.rgen:49-60
{
    TraceArgs a{p, l, tMin, lDist, pcg(pc.rng.state.x ^ pc.rng.state.y), true, nullptr};
    pc.shadowRays++;
    const Hit h = trace(pc.ctx, a);
    return h.drawInstanceIndex == MISS_INDEX ? 1.f : 0.f;
}

static Hit traceClosest(PathCtx &pc, const Ray &ray, float *margin) // main.rgen:62-81
{
    TraceArgs a{ray.o, ray.d, ray.tMin, ray.tMax, pcg(pc.rng.state.x ^ pc.rng.state.z), false, margin};
    pc.closestRays++;
    return trace(pc.ctx, a);
}

static void addBounce(const ReferencePC &PC, vec3 &accColor, vec3 color, uint32_t bounce) // main.rgen:83-88
{
    if (bounce > 0 && flagClampIndirect(PC)) color = clamp(color, 0.f, 2.f);
    accColor += color;
}

static void importanceSampleBounce(PathCtx &pc, const VisibleSurface &surface, vec3 &throughput, vec3 &rd)
{ // main.rgen:90-144
    const bool specularOnly = surface.material.metallic > 0.999f;
    const float specularWeight = specularOnly ? 1.f : 0.5f;
    const float diffuseWeight = 1.f - specularWeight;
    const Basis normalBasis = orthonormalBasis(surface.normalWS);
    const vec3 vInBasis = toBasis(normalBasis, surface.invViewRayWS);
    const float alpha = surface.material.roughness * surface.material.roughness;
    const bool pickDiffuse = pc.rng.rnd01() < diffuseWeight;
    vec3 brdf;
    float NoL, pdf;
    if (pickDiffuse)
    {
        rd = cosineSampleHemisphere(surface.normalWS, pc.rng.rnd2d01());
        NoL = saturate(dot(surface.normalWS, rd));
        brdf = lambertBRFD(surface.material.albedo);
        pdf = cosineHemispherePdf(NoL);
        pdf *= diffuseWeight;
    }
    else
    {
        rd = sampleVisibleTrowbridgeReitz(vInBasis, alpha, pc.rng.rnd2d01());
        rd = fromBasis(normalBasis, rd);
        NoL = saturate(dot(surface.normalWS, rd));
        const vec3 h = normalize(surface.invViewRayWS + rd);
        const float NoH = saturate(dot(surface.normalWS, h));
        const float VoH = saturate(dot(surface.invViewRayWS, h));
        const vec3 f0 = fresnelZero(surface);
        brdf = cookTorranceBRDF(NoL, surface.NoV, NoH, VoH, f0, surface.material.roughness);
        pdf = visibleTrowbridgeReitzPdf(vInBasis, toBasis(normalBasis, rd), alpha);
        pdf *= specularWeight;
    }
    // GLSL max(x, 0) with NaN x is undefined; both the oracle and the CUDA path use
    // "NaN -> 0" (fmaxf semantics), DESIGN.md section 4.
    const vec3 w = brdf * NoL / pdf;
    throughput *= vec3{std::fmax(w.x, 0.f), std::fmax(w.y, 0.f), std::fmax(w.z, 0.f)};
}

static VisibleSurface evaluateSurface(PathCtx &pc, const Ray &ray, const Hit &hit) // main.rgen:146-179
{
    const Scene &sc = pc.ctx.scene;
    const DrawInstance &instance = sc.drawInstance(hit.drawInstanceIndex);
    const ModelInstanceTransforms &trfn = sc.transforms(instance.modelInstanceIndex);
    const GeometryMetadata &metadata = sc.metadata(instance.meshIndex);
    const Vertex v0 = loadVertexThroughIndexBuffer(sc, metadata, hit.primitiveID * 3 + 0);
    const Vertex v1 = loadVertexThroughIndexBuffer(sc, metadata, hit.primitiveID * 3 + 1);
    const Vertex v2 = loadVertexThroughIndexBuffer(sc, metadata, hit.primitiveID * 3 + 2);
    const Vertex vi = interpolate(v0, v1, v2, hit.bary);
    const Vertex v = transform(vi, trfn);
    VisibleSurface surface{};
    surface.positionWS = v.Position;
    surface.invViewRayWS = -ray.d;
    surface.uv = v.TexCoord0;
    surface.material = sampleMaterial(sc, instance.materialIndex, v.TexCoord0);
    if (surface.material.normal.x != -2 && v.Tangent.w != 0)
        surface.normalWS = mappedNormal(surface.material.normal, v.Normal, xyz(v.Tangent), v.Tangent.w);
    else
        surface.normalWS = normalize(v.Normal);
    surface.NoV = saturate(dot(surface.normalWS, surface.invViewRayWS));
    surface.linearDepth = length(surface.positionWS - ray.o);
    pc.shadedHits++;
    return surface;
}

static vec3 uintToColor(uint32_t x) // random.glsl:30-40
{
    const uint32_t xr = pcg(x);
    return vec3{(float)((xr >> 20) & 0x3FF), (float)((xr >> 10) & 0x3FF), (float)(xr & 0x3FF)} / (float)0x3FF;
}

static vec3 debugColor(PathCtx &pc, const Hit &hit, const VisibleSurface &surface) // main.rgen:181-193; debug.glsl:17-38
{
    const DrawInstance &instance = pc.ctx.scene.drawInstance(hit.drawInstanceIndex);
    const Material &m = surface.material;
    switch (pc.PC.drawType)
    {
    case DrawType_PrimitiveID: return uintToColor(hit.primitiveID);
    case DrawType_MeshID: return uintToColor(instance.meshIndex);
    case DrawType_MaterialID: return uintToColor(instance.materialIndex);
    case DrawType_Position: return surface.positionWS;
    case DrawType_TexCoord0: return {surface.uv.x, surface.uv.y, 0.f};
    case DrawType_Albedo: return m.albedo;
    case DrawType_ShadingNormal: return surface.normalWS * 0.5f + vec3{0.5f, 0.5f, 0.5f};
    case DrawType_Roughness: return {m.roughness, m.roughness, m.roughness};
    case DrawType_Metallic: return {m.metallic, m.metallic, m.metallic};
    default: return {1.f, 0.f, 1.f};
    }
}

// scene/lighting.glsl:15-89
static void sampleLight(const Scene &sc, const VisibleSurface &surface, uint32_t lightIndex, vec3 &l, float &d,
                        vec3 &irradiance)
{
    if (lightIndex == 0)
    {
        const DirectionalLight &dl = *(const DirectionalLight *)sc.s.directionalLight;
        l = -normalize(vec3{dl.direction[0], dl.direction[1], dl.direction[2]});
        d = 100.f;
        irradiance = {dl.irradiance[0], dl.irradiance[1], dl.irradiance[2]};
        return;
    }
    lightIndex -= 1;
    if (lightIndex < sc.s.pointLightCount)
    {
        const PointLight &light = ((const PointLight *)sc.s.pointLights)[lightIndex];
        const vec3 pos = {light.position[0], light.position[1], light.position[2]};
        const vec3 radiance = {light.radianceAndRadius[0], light.radianceAndRadius[1], light.radianceAndRadius[2]};
        const float radius = light.radianceAndRadius[3];
        const vec3 toLight = pos - surface.positionWS;
        const float d2 = dot(toLight, toLight);
        d = std::sqrt(d2);
        l = toLight / d;
        const float dPerR = d / radius;
        const float dPerR2 = dPerR * dPerR;
        const float dPerR4 = dPerR2 * dPerR2;
        const float radialAttenuation = std::max(std::min(1.0f - dPerR4, 1.f), 0.f);
        irradiance = radiance * radialAttenuation / d2;
        return;
    }
    lightIndex -= sc.s.pointLightCount;
    if (lightIndex < sc.s.spotLightCount)
    {
        const SpotLight &light = ((const SpotLight *)sc.s.spotLights)[lightIndex];
        const vec3 toLight =
            vec3{light.positionAndAngleOffset[0], light.positionAndAngleOffset[1], light.positionAndAngleOffset[2]} -
            surface.positionWS;
        const float d2 = dot(toLight, toLight);
        d = std::sqrt(d2);
        l = toLight / d;
        const float cd = dot(-vec3{light.direction[0], light.direction[1], light.direction[2]}, l);
        float angularAttenuation = saturate(cd * light.radianceAndAngleScale[3] + light.positionAndAngleOffset[3]);
        angularAttenuation *= angularAttenuation;
        irradiance =
            vec3{light.radianceAndAngleScale[0], light.radianceAndAngleScale[1], light.radianceAndAngleScale[2]} *
            angularAttenuation / d2;
        return;
    }
    l = {0.f, 1.f, 0.f};
    d = 1.f;
    irradiance = {0.f, 0.f, 0.f};
}

static vec3 evaluateDirectLighting(PathCtx &pc, const VisibleSurface &surface, vec3 throughput) // main.rgen:195-223
{
    if (surface.material.alpha == 0) return {0.f, 0.f, 0.f};
    const Scene &sc = pc.ctx.scene;
    const uint32_t lightCount = 1 + sc.s.pointLightCount + sc.s.spotLightCount;
    const uint32_t lightIndex = std::min((uint32_t)(pc.rng.rnd01() * (float)lightCount), lightCount - 1);
    vec3 l;
    float d;
    vec3 irradiance;
    sampleLight(sc, surface, lightIndex, l, d, irradiance);
    if (dot(l, surface.normalWS) <= 0) return {0.f, 0.f, 0.f};
    irradiance = irradiance * shadow(pc, surface.positionWS, l, 0.1f, d);
    irradiance = irradiance * (float)lightCount;
    return throughput * irradiance * evalBRDFTimesNoL(l, surface);
}

static vec3 pathSample(const orc_ctx &ctx, const ReferencePC &PC, uint32_t px, uint32_t py, uint32_t width,
                       uint32_t height, uint64_t *counters) // main.rgen:225-283
{
    PathCtx pc{ctx, PC, Rng{{px, py, PC.frameIndex}}};
    const vec2 r = pc.rng.rnd2d01();
    const vec2 jitteredUV = {((float)px + r.x) / (float)width, ((float)py + r.y) / (float)height};
    vec3 color = {0.f, 0.f, 0.f};
    vec3 throughput = {1.f, 1.f, 1.f};
    uint32_t bounce = 0;
    const CameraUniforms &cam = ctx.scene.camera();
    Ray ray;
    if (flagDepthOfField(PC))
    {
        const vec2 lens = pc.rng.rnd2d01();
        ray = thinLensCameraRay(cam, jitteredUV, lens, PC.apertureDiameter, PC.focusDistance, PC.focalLength);
    }
    else
        ray = pinholeCameraRay(cam, jitteredUV);
    while (bounce < MAX_BOUNCES)
    {
        if (bounce >= PC.maxBounces) break;
        const Hit hit = traceClosest(pc, ray, nullptr);
        if (hit.drawInstanceIndex == MISS_INDEX)
        {
            if (flagIBL(PC))
            {
                const vec3 skyColor = sampleSky(ctx.scene, ray.d);
                addBounce(PC, color, throughput * skyColor, bounce);
            }
            break;
        }
        const VisibleSurface surface = evaluateSurface(pc, ray, hit);
        if (PC.drawType != DrawType_Default && PC.drawType != DrawType_MeshletID)
        {
            color = debugColor(pc, hit, surface);
            break;
        }
        addBounce(PC, color, evaluateDirectLighting(pc, surface, throughput), bounce);
        vec3 rd;
        importanceSampleBounce(pc, surface, throughput, rd);
        if (bounce > PC.rouletteStartBounce)
        {
            if (pc.rng.rnd01() < std::max(.05f, 1.f - max3(throughput))) break;
        }
        ray.o = offsetRay(surface.positionWS, surface.normalWS);
        ray.d = rd;
        ray.tMin = 0;
        ray.tMax = INFINITY;
        bounce++;
    }
    if (counters)
    {
        counters[0] += pc.closestRays;
        counters[1] += pc.shadowRays;
        counters[2] += pc.shadedHits;
        counters[3] += 1;
    }
    return color;
}

static Ray primaryRay(const orc_ctx &ctx, const ReferencePC &PC, uint32_t px, uint32_t py, uint32_t width,
                      uint32_t height, Rng &rng)
{
    rng = Rng{{px, py, PC.frameIndex}};
    const vec2 r = rng.rnd2d01();
    const vec2 jitteredUV = {((float)px + r.x) / (float)width, ((float)py + r.y) / (float)height};
    const CameraUniforms &cam = ctx.scene.camera();
    if (flagDepthOfField(PC))
    {
        const vec2 lens = rng.rnd2d01();
        return thinLensCameraRay(cam, jitteredUV, lens, PC.apertureDiameter, PC.focusDistance, PC.focalLength);
    }
    return pinholeCameraRay(cam, jitteredUV);
}

template <typename F> static void parallelRows(uint32_t rows, int threads, F &&fn)
{
    int n = threads > 0 ? threads : (int)std::thread::hardware_concurrency();
    n = std::max(1, std::min(n, (int)rows));
    std::atomic<uint32_t> next{0};
    auto worker = [&](int tid) {
        for (;;)
        {
            const uint32_t y = next.fetch_add(1);
            if (y >= rows) break;
            fn(y, tid);
        }
    };
    if (n == 1) { worker(0); return; }
    std::vector<std::thread> pool;
    for (int i = 0; i < n; ++i) pool.emplace_back(worker, i);
    for (auto &t : pool) t.join();
}

// ---- ReSTIR DI trace pass: rt/direct_illumination/main.rgen (+ restir_di/*.glsl, initial_reservoirs.comp) ---------------------
struct GBuffer
{
    const uint32_t *albedoRoughness; // R8G8B8A8_UNORM (GBufferRenderer.cpp:31)
    const uint32_t *normalMetallic;  // A2B10G10R10_UNORM_PACK32 (GBufferRenderer.cpp:32)
    const float *nonLinearDepth;
    uint32_t width, height;
};
static vec4 unpackUnorm8x4(uint32_t p)
{
    return {(float)(p & 255u) / 255.f, (float)((p >> 8) & 255u) / 255.f, (float)((p >> 16) & 255u) / 255.f, (float)(p >> 24) / 255.f};
}
static vec4 unpackA2B10G10R10(uint32_t p)
{
    return {(float)(p & 1023u) / 1023.f, (float)((p >> 10) & 1023u) / 1023.f, (float)((p >> 20) & 1023u) / 1023.f, (float)(p >> 30) / 3.f};
}
static uint32_t unormCode(float v, float top) { return (uint32_t)(std::fmin(std::fmax(v, 0.f), 1.f) * top + 0.5f); }

static vec3 signedOctDecode(vec3 n) // scene/material.glsl:20-31
{
    vec3 OutN;
    OutN.x = (n.x - n.y);
    OutN.y = (n.x + n.y) - 1.0f;
    OutN.z = n.z * 2.0f - 1.0f;
    OutN.z = OutN.z * (1.0f - std::fabs(OutN.x) - std::fabs(OutN.y));
    return normalize(OutN);
}
static vec3 signedOctEncode(vec3 n) // gbuffer.frag:41-55
{
    vec3 OutN;
    n = n / (std::fabs(n.x) + std::fabs(n.y) + std::fabs(n.z));
    OutN.y = n.y * 0.5f + 0.5f;
    OutN.x = n.x * 0.5f + OutN.y;
    OutN.y = n.x * -0.5f + OutN.y;
    const float fltMax = 3.40282e+38f;
    OutN.z = saturate(n.z * fltMax);
    return OutN;
}
static Material loadFromGbuffer(const GBuffer &gb, uint32_t px, uint32_t py) // scene/material.glsl:33-48
{
    const size_t i = (size_t)py * gb.width + px;
    const vec4 albedoRoughness = unpackUnorm8x4(gb.albedoRoughness[i]);
    const vec4 normalMetallic = unpackA2B10G10R10(gb.normalMetallic[i]);
    Material m;
    m.albedo = {albedoRoughness.x, albedoRoughness.y, albedoRoughness.z};
    m.roughness = albedoRoughness.w;
    m.normal = signedOctDecode(vec3{normalMetallic.x, normalMetallic.y, normalMetallic.w});
    m.metallic = normalMetallic.z;
    m.alpha = -1.0f;
    return m;
}
static vec4 mulMat4(const float m[4][4], vec4 v)
{
    return {m[0][0] * v.x + m[1][0] * v.y + m[2][0] * v.z + m[3][0] * v.w, m[0][1] * v.x + m[1][1] * v.y + m[2][1] * v.z + m[3][1] * v.w,
            m[0][2] * v.x + m[1][2] * v.y + m[2][2] * v.z + m[3][2] * v.w, m[0][3] * v.x + m[1][3] * v.y + m[2][3] * v.z + m[3][3] * v.w};
}
static vec3 worldPos(const CameraUniforms &c, vec2 uv, float nonLinearDepth) // scene/camera.glsl:24-30
{
    const vec4 surfaceVS = mulMat4(c.clipToWorld, vec4{uv.x * 2 - 1, uv.y * 2 - 1, nonLinearDepth, 1});
    return vec3{surfaceVS.x, surfaceVS.y, surfaceVS.z} / surfaceVS.w;
}
static float linearizeDepth(const CameraUniforms &c, float nonLinearDepth) // scene/camera.glsl:11-22
{
    return -c.cameraToClip[3][2] / (nonLinearDepth + c.cameraToClip[2][2]);
}
static VisibleSurface gbufferSurface(const Scene &sc, const GBuffer &gb, uint32_t px, uint32_t py) // main.rgen:117-131
{
    const CameraUniforms &camera = *(const CameraUniforms *)sc.s.camera;
    const vec2 uv = {(float)px / (float)gb.width, (float)py / (float)gb.height};
    const float nonLinearDepth = gb.nonLinearDepth[(size_t)py * gb.width + px];
    VisibleSurface surface{};
    surface.positionWS = worldPos(camera, uv, nonLinearDepth);
    surface.invViewRayWS = normalize(vec3{camera.eye[0], camera.eye[1], camera.eye[2]} - surface.positionWS);
    surface.material = loadFromGbuffer(gb, px, py);
    surface.normalWS = surface.material.normal;
    surface.uv = {0, 0};
    surface.NoV = saturate(dot(surface.normalWS, surface.invViewRayWS));
    surface.linearDepth = linearizeDepth(camera, nonLinearDepth);
    return surface;
}

struct LightReservoir // restir_di/reservoir.glsl:4-9
{
    int32_t lightIndex;
    float unbiasedContributionWeight;
};
static float luminance(vec3 c) { return dot(vec3{0.299f, 0.587f, 0.114f}, c); } // common/math.glsl:15
static float pHatLight(const Scene &sc, const VisibleSurface &surface, uint32_t lightIndex) // resampling_phat.glsl:4-19
{
    vec3 l;
    float d;
    vec3 irradiance;
    sampleLight(sc, surface, lightIndex, l, d, irradiance);
    const vec3 brdf = evalBRDFTimesNoL(l, surface);
    return luminance(irradiance * brdf);
}
static LightReservoir initialLightCandidate(const Scene &sc, const VisibleSurface &surface, Rng &rng) // initial_reservoirs.comp:31-60
{
    const int lightCount = 1 + (int)(sc.s.pointLightCount + sc.s.spotLightCount);
    LightReservoir reservoir = {-1, 0.f};
    float sumResamplingWeights = 0;
    const uint32_t sampleCount = 5;
    for (uint32_t i = 0; i < sampleCount; ++i)
    {
        const int lightIndex = std::min((int)(rng.rnd01() * (float)lightCount), lightCount - 1);
        const float misWeight = 1.f / sampleCount;
        const float unbiasedContributionWeight = (float)lightCount;
        const float resamplingWeight = misWeight * pHatLight(sc, surface, (uint32_t)lightIndex) * unbiasedContributionWeight;
        sumResamplingWeights += resamplingWeight;
        if (rng.rnd01() < resamplingWeight / sumResamplingWeights) reservoir.lightIndex = lightIndex;
    }
    if (reservoir.lightIndex >= 0)
        reservoir.unbiasedContributionWeight = sumResamplingWeights / pHatLight(sc, surface, (uint32_t)reservoir.lightIndex);
    return reservoir;
}
static vec3 evaluateDirectLightingReSTIR(PathCtx &pc, const VisibleSurface &surface, LightReservoir reservoir) // main.rgen:89-110
{
    if (surface.material.alpha == 0 || reservoir.lightIndex < 0) return {0, 0, 0};
    vec3 l;
    float d;
    vec3 irradiance;
    sampleLight(pc.ctx.scene, surface, (uint32_t)reservoir.lightIndex, l, d, irradiance);
    if (dot(l, surface.normalWS) <= 0) return {0, 0, 0};
    irradiance = irradiance * shadow(pc, surface.positionWS, l, 0.1f, d);
    return irradiance * evalBRDFTimesNoL(l, surface) * reservoir.unbiasedContributionWeight;
}

extern "C" {
void orc_rtdi_trace(orc_ctx *ctx, const uint32_t tracePC[3], uint32_t width, uint32_t height, const uint32_t *albedoRoughness,
                    const uint32_t *normalMetallic, const float *nonLinearDepth, const float *reservoirs, float *rgba, int threads,
                    uint64_t *shadowRays)
{ // rt/direct_illumination/main.rgen:112-164
    const uint32_t drawType = tracePC[0], frameIndex = tracePC[1], flags = tracePC[2];
    const GBuffer gb{albedoRoughness, normalMetallic, nonLinearDepth, width, height};
    const int nThreads = threads > 0 ? threads : (int)std::thread::hardware_concurrency();
    std::vector<uint64_t> perThread((size_t)std::max(nThreads, 1), 0);
    ReferencePC dummy{};
    parallelRows(height, nThreads, [&](uint32_t y, int tid) {
        for (uint32_t x = 0; x < width; ++x)
        {
            PathCtx pc{*ctx, dummy, Rng{{x, y, frameIndex}}};
            const VisibleSurface surface = gbufferSurface(ctx->scene, gb, x, y);
            float *out = rgba + ((size_t)y * width + x) * 4;
            if (drawType != 0) // DrawType_Default
            {
                const vec3 c = drawType == 5 /* DrawType_Position */ ? surface.positionWS : surface.material.albedo;
                out[0] = c.x; out[1] = c.y; out[2] = c.z; out[3] = 1.f;
                continue;
            }
            const size_t i = (size_t)y * width + x;
            LightReservoir reservoir;
            std::memcpy(&reservoir.lightIndex, &reservoirs[i * 2], 4); // floatBitsToInt
            reservoir.unbiasedContributionWeight = reservoirs[i * 2 + 1];
            const vec3 color = evaluateDirectLightingReSTIR(pc, surface, reservoir);
            perThread[(size_t)tid] += pc.shadowRays;
            if ((flags & 1u) || !((flags >> 1) & 1u))
            {
                out[0] = color.x; out[1] = color.y; out[2] = color.z; out[3] = 1.f;
            }
            else
            {
                const vec3 historyColor = {out[0], out[1], out[2]};
                const float historyCount = out[3] + 1;
                const vec3 newColor = historyColor + (color - historyColor) / historyCount;
                out[0] = newColor.x; out[1] = newColor.y; out[2] = newColor.z; out[3] = historyCount;
            }
        }
    });
    if (shadowRays)
        for (uint64_t v : perThread) *shadowRays += v;
}

void orc_rtdi_initial_reservoirs(orc_ctx *ctx, uint32_t frameIndex, uint32_t width, uint32_t height, const uint32_t *albedoRoughness,
                                 const uint32_t *normalMetallic, const float *nonLinearDepth, float *reservoirsOut, int threads)
{ // restir_di/initial_reservoirs.comp:62-89
    const GBuffer gb{albedoRoughness, normalMetallic, nonLinearDepth, width, height};
    parallelRows(height, threads, [&](uint32_t y, int) {
        for (uint32_t x = 0; x < width; ++x)
        {
            Rng rng{{x, y, frameIndex}};
            const VisibleSurface surface = gbufferSurface(ctx->scene, gb, x, y);
            const LightReservoir r = initialLightCandidate(ctx->scene, surface, rng);
            const size_t i = (size_t)y * width + x;
            std::memcpy(&reservoirsOut[i * 2], &r.lightIndex, 4); // intBitsToFloat
            reservoirsOut[i * 2 + 1] = r.unbiasedContributionWeight;
        }
    });
}

// The G-buffer texel gbuffer.frag:57-110 writes, for the primary hit through each pixel CENTRE (the raster sample point).
void orc_rtdi_gbuffer(orc_ctx *ctx, uint32_t frameIndex, uint32_t width, uint32_t height, uint32_t *albedoRoughness, uint32_t *normalMetallic,
                      float *nonLinearDepth, int threads)
{
    const CameraUniforms &camera = *(const CameraUniforms *)ctx->scene.s.camera;
    ReferencePC dummy{};
    parallelRows(height, threads, [&](uint32_t y, int) {
        for (uint32_t x = 0; x < width; ++x)
        {
            const size_t i = (size_t)y * width + x;
            albedoRoughness[i] = 0; normalMetallic[i] = 0; nonLinearDepth[i] = 0.f;
            PathCtx pc{*ctx, dummy, Rng{{x, y, frameIndex}}};
            const Ray ray = pinholeCameraRay(camera, vec2{((float)x + 0.5f) / (float)width, ((float)y + 0.5f) / (float)height});
            TraceArgs a{ray.o, ray.d, ray.tMin, ray.tMax, pcg(x ^ frameIndex), false, nullptr};
            const Hit hit = trace(*ctx, a);
            if (hit.drawInstanceIndex == MISS_INDEX) continue;
            const VisibleSurface surface = evaluateSurface(pc, ray, hit);
            if (surface.material.alpha == 0) continue; // gbuffer.frag:62-63
            const Material &m = surface.material;
            albedoRoughness[i] = unormCode(m.albedo.x, 255.f) | (unormCode(m.albedo.y, 255.f) << 8) | (unormCode(m.albedo.z, 255.f) << 16) |
                                 (unormCode(m.roughness, 255.f) << 24);
            const vec3 e = signedOctEncode(surface.normalWS);
            normalMetallic[i] = unormCode(e.x, 1023.f) | (unormCode(e.y, 1023.f) << 10) | (unormCode(m.metallic, 1023.f) << 20) | (unormCode(e.z, 3.f) << 30);
            const vec4 view = mulMat4(camera.worldToCamera, vec4{surface.positionWS.x, surface.positionWS.y, surface.positionWS.z, 1.f});
            const vec4 clip = mulMat4(camera.cameraToClip, view);
            nonLinearDepth[i] = clip.z / clip.w;
        }
    });
}
} // extern "C"

static void buildStructures(orc_ctx &ctx)
{
    const auto t0 = std::chrono::steady_clock::now();
    const Scene &sc = ctx.scene;
    const orc_scene &s = sc.s;
    ctx.blases.resize(s.modelCount);
    for (uint32_t m = 0; m < s.modelCount; ++m)
    {
        Blas &blas = ctx.blases[m];
        std::vector<BlasTri> tris;
        for (uint32_t g = s.modelOffsets[m]; g < s.modelOffsets[m + 1]; ++g)
        {
            const SubModel &sm = sc.subModel(g);
            const GeometryMetadata &md = sc.metadata(sm.meshIndex);
            const MeshInfo &info = sc.meshInfo(sm.meshIndex);
            // opaque flag from the MESH's material (World.cpp:646-651)
            const uint32_t opaque = sc.material(info.materialIndex).alphaMode == AlphaModeOpaque ? 1u : 0u;
            const uint32_t primCount = info.indexCount / 3;
            for (uint32_t p = 0; p < primCount; ++p)
            {
                BlasTri t;
                const uint32_t i0 = loadIndex(sc, md.bufferIndex, md.indicesOffset, p * 3 + 0, md.usesShortIndices);
                const uint32_t i1 = loadIndex(sc, md.bufferIndex, md.indicesOffset, p * 3 + 1, md.usesShortIndices);
                const uint32_t i2 = loadIndex(sc, md.bufferIndex, md.indicesOffset, p * 3 + 2, md.usesShortIndices);
                t.v0 = loadPosition(sc, md, i0);
                t.v1 = loadPosition(sc, md, i1);
                t.v2 = loadPosition(sc, md, i2);
                t.geometry = g - s.modelOffsets[m];
                t.primitive = p;
                t.opaque = opaque;
                tris.push_back(t);
            }
        }
        if (tris.empty()) continue;
        Aabb all;
        for (const BlasTri &t : tris) { all.grow(t.v0); all.grow(t.v1); all.grow(t.v2); }
        const float maxAbs = std::max(std::max(std::max(std::fabs(all.lo.x), std::fabs(all.hi.x)),
                                               std::max(std::fabs(all.lo.y), std::fabs(all.hi.y))),
                                      std::max(std::fabs(all.lo.z), std::fabs(all.hi.z)));
        const float pad = std::max(maxAbs * (1.0f / 16384.0f), 1e-6f);
        std::vector<Aabb> boxes(tris.size());
        for (size_t i = 0; i < tris.size(); ++i)
        {
            Aabb b;
            b.grow(tris[i].v0); b.grow(tris[i].v1); b.grow(tris[i].v2);
            b.lo = b.lo - vec3{pad, pad, pad};
            b.hi = b.hi + vec3{pad, pad, pad};
            boxes[i] = b;
        }
        if (ctx.bruteForce)
        {
            blas.tris = std::move(tris);
            blas.bvh.nodes.emplace_back();
            Aabb rb;
            for (const Aabb &b : boxes) rb.grow(b);
            blas.bvh.nodes[0].box = rb;
            continue;
        }
        blas.bvh = buildBvh(boxes, 4);
        blas.tris.resize(tris.size());
        for (size_t i = 0; i < tris.size(); ++i) blas.tris[i] = tris[blas.bvh.items[i]];
    }
    // TLAS (World.cpp:878-928, 749-802)
    std::vector<Aabb> instBoxes;
    for (uint32_t i = 0; i < s.instanceCount; ++i)
    {
        const TlasInstance &ti = sc.tlasInstance(i);
        if (ti.blasReference == 0) continue; // inactive instance (World.cpp:898-905)
        const uint32_t model = (uint32_t)(ti.blasReference - 1);
        if (model >= s.modelCount || ctx.blases[model].tris.empty()) continue;
        InstanceRec rec;
        invertTransform(&ti.transform[0][0], rec.w2o);
        rec.customIndex = ti.customIndexAndMask & 0xFFFFFFu;
        rec.mask = ti.customIndexAndMask >> 24;
        rec.model = model;
        rec.index = i;
        ctx.instances.push_back(rec);
        const Aabb ob = ctx.blases[model].bvh.bounds();
        Aabb wb;
        const float *m = &ti.transform[0][0];
        for (int c = 0; c < 8; ++c)
        {
            const vec3 p = {(c & 1) ? ob.hi.x : ob.lo.x, (c & 2) ? ob.hi.y : ob.lo.y, (c & 4) ? ob.hi.z : ob.lo.z};
            wb.grow(transformPoint(m, p));
        }
        const vec3 e = wb.hi - wb.lo;
        const float pad = std::max(std::max(e.x, std::max(e.y, e.z)) * (1.0f / 16384.0f), 1e-6f);
        wb.lo = wb.lo - vec3{pad, pad, pad};
        wb.hi = wb.hi + vec3{pad, pad, pad};
        instBoxes.push_back(wb);
    }
    if (!ctx.bruteForce) ctx.tlas = buildBvh(instBoxes, 1);
    ctx.buildSeconds = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
}

} // namespace orc

// =================================================================================================
extern "C" {

orc_ctx *orc_create(const orc_scene *scene, int bruteForce)
{
    orc_ctx *ctx = new orc_ctx;
    ctx->scene.s = *scene;
    ctx->bruteForce = bruteForce != 0;
    ctx->scene.decodeTextures();
    buildStructures(*ctx);
    return ctx;
}
void orc_destroy(orc_ctx *ctx) { delete ctx; }
double orc_bvh_build_seconds(const orc_ctx *ctx) { return ctx->buildSeconds; }

void orc_render(orc_ctx *ctx, const void *pcIn, uint32_t width, uint32_t height, uint32_t sampleCount, float *rgba,
                int threads, uint64_t *counters)
{
    const ReferencePC pc0 = *(const ReferencePC *)pcIn;
    const int nThreads = threads > 0 ? threads : (int)std::thread::hardware_concurrency();
    std::vector<uint64_t> perThread((size_t)std::max(nThreads, 1) * 4, 0);
    parallelRows(height, nThreads, [&](uint32_t y, int tid) {
        for (uint32_t x = 0; x < width; ++x)
        {
            float *out = rgba + ((size_t)y * width + x) * 4;
            for (uint32_t k = 0; k < sampleCount; ++k)
            {
                ReferencePC PC = pc0;
                PC.frameIndex = (pc0.frameIndex + k) % 4096u; // RtRef.cpp:170
                if (k > 0) PC.flags &= ~1u;                   // skipHistory only on the first frame
                const vec3 color = pathSample(*ctx, PC, x, y, width, height, &perThread[(size_t)tid * 4]);
                if (flagSkipHistory(PC) || !flagAccumulate(PC)) // main.rgen:285-298
                {
                    out[0] = color.x; out[1] = color.y; out[2] = color.z; out[3] = 1.f;
                }
                else
                {
                    const vec3 historyColor = {out[0], out[1], out[2]};
                    const float historyCount = out[3] + 1;
                    const vec3 newColor = historyColor + (color - historyColor) / historyCount; // main.rgen:293
                    out[0] = newColor.x; out[1] = newColor.y; out[2] = newColor.z; out[3] = historyCount;
                }
            }
        }
    });
    if (counters)
        for (int t = 0; t < std::max(nThreads, 1); ++t)
            for (int c = 0; c < 4; ++c) counters[c] += perThread[(size_t)t * 4 + c];
}

void orc_trace_primary(orc_ctx *ctx, const void *pcIn, uint32_t width, uint32_t height, void *hitsOut,
                       float *alphaMargin, int threads)
{
    const ReferencePC PC = *(const ReferencePC *)pcIn;
    struct Rec { uint32_t di, prim; float bx, by; };
    Rec *hits = (Rec *)hitsOut;
    parallelRows(height, threads, [&](uint32_t y, int) {
        for (uint32_t x = 0; x < width; ++x)
        {
            Rng rng;
            const Ray ray = primaryRay(*ctx, PC, x, y, width, height, rng);
            float margin = 1e30f;
            TraceArgs a{ray.o, ray.d, ray.tMin, ray.tMax, pcg(rng.state.x ^ rng.state.z), false, &margin};
            const Hit h = trace(*ctx, a);
            hits[(size_t)y * width + x] = {h.drawInstanceIndex, h.primitiveID, h.bary.x, h.bary.y};
            if (alphaMargin) alphaMargin[(size_t)y * width + x] = margin;
        }
    });
}

void orc_trace_rays(orc_ctx *ctx, const float *originTmin, const float *dirTmax, const uint32_t *seeds, uint32_t count,
                    int shadow, void *hitsOut, float *alphaMargin, int threads)
{
    struct Rec { uint32_t di, prim; float bx, by; };
    Rec *hits = (Rec *)hitsOut;
    const uint32_t chunk = 1024;
    const uint32_t rows = (count + chunk - 1) / chunk;
    parallelRows(rows, threads, [&](uint32_t r, int) {
        for (uint32_t i = r * chunk; i < std::min(count, (r + 1) * chunk); ++i)
        {
            float margin = 1e30f;
            TraceArgs a{{originTmin[i * 4], originTmin[i * 4 + 1], originTmin[i * 4 + 2]},
                        {dirTmax[i * 4], dirTmax[i * 4 + 1], dirTmax[i * 4 + 2]},
                        originTmin[i * 4 + 3], dirTmax[i * 4 + 3], seeds ? seeds[i] : 0u, shadow != 0, &margin};
            const Hit h = trace(*ctx, a);
            hits[i] = {h.drawInstanceIndex, h.primitiveID, h.bary.x, h.bary.y};
            if (alphaMargin) alphaMargin[i] = margin;
        }
    });
}

// ---- unit entry points ------------------------------------------------------------------------
uint32_t orc_pcg(uint32_t v) { return pcg(v); }
void orc_pcg3d(const uint32_t in[3], uint32_t out[3])
{
    const uvec3 r = pcg3d({in[0], in[1], in[2]});
    out[0] = r.x; out[1] = r.y; out[2] = r.z;
}
float orc_rng_to_01(uint32_t u) { return rngTo01(u); }
void orc_cosine_sample_hemisphere(const float n[3], const float u[2], float out[3])
{
    const vec3 r = cosineSampleHemisphere({n[0], n[1], n[2]}, {u[0], u[1]});
    out[0] = r.x; out[1] = r.y; out[2] = r.z;
}
void orc_sample_vndf(const float ve[3], float alpha, const float u[2], float out[3])
{
    const vec3 r = sampleVisibleTrowbridgeReitz({ve[0], ve[1], ve[2]}, alpha, {u[0], u[1]});
    out[0] = r.x; out[1] = r.y; out[2] = r.z;
}
float orc_vndf_pdf(const float ve[3], const float le[3], float alpha)
{
    return visibleTrowbridgeReitzPdf({ve[0], ve[1], ve[2]}, {le[0], le[1], le[2]}, alpha);
}
void orc_eval_brdf_times_nol(const float l[3], const float v[3], const float n[3], const float albedo[3],
                             float roughness, float metallic, float out[3])
{
    VisibleSurface s{};
    s.normalWS = {n[0], n[1], n[2]};
    s.invViewRayWS = {v[0], v[1], v[2]};
    s.NoV = saturate(dot(s.normalWS, s.invViewRayWS));
    s.material.albedo = {albedo[0], albedo[1], albedo[2]};
    s.material.roughness = roughness;
    s.material.metallic = metallic;
    const vec3 r = evalBRDFTimesNoL({l[0], l[1], l[2]}, s);
    out[0] = r.x; out[1] = r.y; out[2] = r.z;
}
void orc_offset_ray(const float p[3], const float n[3], float out[3])
{
    const vec3 r = offsetRay({p[0], p[1], p[2]}, {n[0], n[1], n[2]});
    out[0] = r.x; out[1] = r.y; out[2] = r.z;
}
void orc_pinhole_ray(const void *camera, const float uv[2], float o[3], float d[3])
{
    const Ray r = pinholeCameraRay(*(const CameraUniforms *)camera, {uv[0], uv[1]});
    o[0] = r.o.x; o[1] = r.o.y; o[2] = r.o.z;
    d[0] = r.d.x; d[1] = r.d.y; d[2] = r.d.z;
}
void orc_thin_lens_ray(const void *camera, const float uv[2], const float lens[2], float aperture, float focusDistance,
                       float focalLength, float o[3], float d[3])
{
    const Ray r = thinLensCameraRay(*(const CameraUniforms *)camera, {uv[0], uv[1]}, {lens[0], lens[1]}, aperture,
                                    focusDistance, focalLength);
    o[0] = r.o.x; o[1] = r.o.y; o[2] = r.o.z;
    d[0] = r.d.x; d[1] = r.d.y; d[2] = r.d.z;
}
void orc_unpack_snorm10(uint32_t packed, float out[3])
{
    const vec3 r = unpackSnormR10G10B10(packed);
    out[0] = r.x; out[1] = r.y; out[2] = r.z;
}
float orc_half_to_float(uint16_t h) { return halfToFloat(h); }
uint16_t orc_float_to_half(float f) { return floatToHalf(f); }
float orc_srgb_to_linear(float x) { return sRGBtoLinear(x); }
int orc_bc7_decode_block(const uint8_t block[16], uint8_t out[64]) { return bc7DecodeBlock(block, out); }
void orc_sample_texture(orc_ctx *ctx, uint32_t tex, uint32_t smp, float u, float v, float out[4])
{
    const vec4 r = sampleTexture(ctx->scene, tex, smp, {u, v});
    out[0] = r.x; out[1] = r.y; out[2] = r.z; out[3] = r.w;
}
void orc_sample_sky(orc_ctx *ctx, const float d[3], float out[3])
{
    const vec3 r = sampleSky(ctx->scene, {d[0], d[1], d[2]});
    out[0] = r.x; out[1] = r.y; out[2] = r.z;
}
int orc_intersect_triangle(const float o[3], const float d[3], float tmin, float tmax, const float v0[3],
                           const float v1[3], const float v2[3], float tuv[3])
{
    float t, u, v;
    if (!intersectTriangle({o[0], o[1], o[2]}, {d[0], d[1], d[2]}, tmin, tmax, {v0[0], v0[1], v0[2]},
                           {v1[0], v1[1], v1[2]}, {v2[0], v2[1], v2[2]}, t, u, v))
        return 0;
    tuv[0] = t; tuv[1] = u; tuv[2] = v;
    return 1;
}
void orc_invert_transform(const float m[12], float out[12]) { invertTransform(m, out); }

} // extern "C"

// ---- display transform (tone_map.comp) ----
namespace {
// GLSL mod
float modGlsl(float a, float b) { return a - b * std::floor(a / b); }
// DXGI R9G9B9E5_SHAREDEXP: value = mantissa / 512 * 2^(exponent - 15)
void rgb9e5(uint32_t t, float rgb[3])
{
    const float scale = std::ldexp(1.0f, (int)(t >> 27) - 15 - 9);
    for (int ch = 0; ch < 3; ++ch) rgb[ch] = (float)((t >> (9 * ch)) & 511u) * scale;
}
// linear filter, clamp to edge, LOD 0 (Vulkan spec "Texel filtering": i0 = floor(u - 0.5), weight = frac(u - 0.5))
void lutFetch(const uint32_t *lut, int dim, const float uvw[3], float rgb[3])
{
    int lo[3], hi[3];
    float fr[3];
    for (int a = 0; a < 3; ++a)
    {
        const float u = uvw[a] * (float)dim - 0.5f;
        const float fl = std::floor(u);
        fr[a] = u - fl;
        if (!(fr[a] >= 0.f)) fr[a] = 0.f;
        long i = std::isnan(fl) ? 0 : (fl < -1.f ? -1 : (fl > (float)dim ? dim : (long)fl));
        lo[a] = (int)std::min<long>(std::max<long>(i, 0), dim - 1);
        hi[a] = (int)std::min<long>(std::max<long>(i + 1, 0), dim - 1);
    }
    rgb[0] = rgb[1] = rgb[2] = 0.f;
    for (int corner = 0; corner < 8; ++corner)
    {
        const int xi = (corner & 1) ? hi[0] : lo[0];
        const int yi = (corner & 2) ? hi[1] : lo[1];
        const int zi = (corner & 4) ? hi[2] : lo[2];
        const float wx = (corner & 1) ? fr[0] : 1.f - fr[0];
        const float wy = (corner & 2) ? fr[1] : 1.f - fr[1];
        const float wz = (corner & 4) ? fr[2] : 1.f - fr[2];
        float texel[3];
        rgb9e5(lut[((size_t)zi * dim + yi) * dim + xi], texel);
        const float w = wx * wy * wz;
        for (int ch = 0; ch < 3; ++ch) rgb[ch] += w * texel[ch];
    }
}
uint8_t unorm8(float v)
{
    if (std::isnan(v) || v <= 0.f) return 0;
    if (v >= 1.f) return 255;
    return (uint8_t)std::nearbyint(v * 255.f); // round to nearest even
}
} // namespace

extern "C" void orc_decode_rgb9e5(uint32_t texel, float out[3]) { rgb9e5(texel, out); }

extern "C" void orc_tone_map(const uint16_t *img, uint32_t width, uint32_t height, float exposure, float contrast,
                             const uint32_t *lut, uint32_t dim, uint8_t *out)
{
    for (size_t p = 0; p < (size_t)width * height; ++p)
    {
        // tone_map.comp:39-43
        float r = halfToFloat(img[4 * p + 0]) * exposure;
        float g = halfToFloat(img[4 * p + 1]) * exposure;
        float b = halfToFloat(img[4 * p + 2]) * exposure;
        // math.glsl:18-42
        const float value = std::max(std::max(r, g), b);
        const float low = std::min(std::min(r, g), b);
        const float chroma = value - low;
        float hue = 0.f;
        if (chroma != 0.f)
        {
            if (value == r) hue = modGlsl((g - b) / chroma, 6.f);
            else if (value == g) hue = (b - r) / chroma + 2.f;
            else hue = (r - g) / chroma + 4.f;
        }
        const float sat = value == 0.f ? 0.f : chroma / value;
        // tone_map.comp:50
        const float v2 = std::pow(value, contrast);
        // math.glsl:45-77
        const float c2 = v2 * sat;
        const float x = c2 * (1.f - std::fabs(modGlsl(hue, 2.f) - 1.f));
        const int sector = hue < 1.f ? 0 : hue < 2.f ? 1 : hue < 3.f ? 2 : hue < 4.f ? 3 : hue < 5.f ? 4 : 5;
        static const int pick[6][3] = {{0, 1, 2}, {1, 0, 2}, {2, 0, 1}, {2, 1, 0}, {1, 2, 0}, {0, 2, 1}}; // 0 chroma, 1 x, 2 zero
        const float parts[3] = {c2, x, 0.f};
        const float m = v2 - c2;
        const float rgb[3] = {parts[pick[sector][0]] + m, parts[pick[sector][1]] + m, parts[pick[sector][2]] + m};
        // tone_map.comp:17-27
        const float dims = (float)dim;
        float uvw[3], mapped[3];
        for (int ch = 0; ch < 3; ++ch) uvw[ch] = rgb[ch] / (rgb[ch] + 1.f) * ((dims - 1.f) / dims) + 0.5f / dims;
        lutFetch(lut, (int)dim, uvw, mapped);
        // tone_map.comp:57-61
        for (int ch = 0; ch < 3; ++ch) out[4 * p + ch] = unorm8(std::pow(mapped[ch], 1.f / 2.2f));
        out[4 * p + 3] = 255;
    }
}
