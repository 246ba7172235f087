// core_scene.h -- device view of the scene buffers and the fetch functions of
// res/shader/scene/{geometry,instances,materials}.glsl, one function per GLSL function.
#pragma once
#include "hd.h"
#include "core_bc7.h"
#include "../../include/prt_types.h"

#ifdef __CUDACC__
#include <cuda_runtime.h>
#endif

namespace prt {

// One bindless slot = (texture, sampler) pair (materials.glsl:21-24, 37-38).
//  * `payload` is level 0 of the texture as the texture cache stores it (BC7 blocks or RGBA8 texels) in
//    HBM: the software bilinear of the any-hit stage decodes texels straight from it, so the
//    accept/ignore decisions of scene.rahit are bit-identical with the oracle's and the traversal
//    kernels contain no texture instruction at all;
//  * `linear` is the hardware-filtered CUDA texture object over the same data (BC7 CUDA array), used for
//    the material colour taps in the shade stage;
//  * `texels` is a point-sampled view of that array, used only by the test probe that pins what the
//    texture units decode against the software decode.
struct TexSlot
{
    const uint8_t *payload;
#ifdef __CUDACC__
    cudaTextureObject_t linear;
    cudaTextureObject_t texels;
#else
    uint64_t linear, texels;
#endif
    uint32_t format; // PRT_TEX_*
    uint32_t width, height;
    uint32_t magFilter; // 0 nearest, 1 linear
    uint32_t addressU, addressV;
    uint32_t valid;
    uint32_t alphaOpaque;  // every level-0 texel has alpha 255: the alpha of any tap is the filter of ones
    const uint8_t *alpha;  // level-0 alpha channel, one byte per texel exactly as the BC7/RGBA8 payload decodes
                           // (built at upload, prt_set_texture); nullptr: decode the payload per tap
};

// Everything scene.rahit's sampleAlpha needs for one draw instance, flattened into one 32-byte record (two
// 128-bit loads) so the any-hit stage does not chase draw instance -> material -> texture slot. Derived data,
// rebuilt by k_alpha_mats whenever draw instances, materials or textures change.
struct alignas(16) AlphaMat
{
    const uint8_t *alpha; // level-0 alpha plane of the base colour texture (nullptr unless kind == PLANE)
    uint32_t width, height;
    uint32_t bits;        // see ALPHAMAT_*
    float factor;         // baseColorFactor.a
    float cutoff;         // alphaCutoff
    uint32_t pad;
};
static_assert(sizeof(AlphaMat) == 32, "AlphaMat");
constexpr uint32_t ALPHAMAT_TEXTURED = 1u;      // a valid base colour texture is bound
constexpr uint32_t ALPHAMAT_TEX_OPAQUE = 2u;    // ... whose texels all have alpha 255
constexpr uint32_t ALPHAMAT_NEAREST = 4u;       // mag filter
constexpr uint32_t ALPHAMAT_ADDRU_SHIFT = 4, ALPHAMAT_ADDRV_SHIFT = 6, ALPHAMAT_MODE_SHIFT = 8;
constexpr uint32_t ALPHAMAT_USABLE = 1u << 31;  // 0: this draw instance must take the generic any-hit path

struct SceneView
{
    const uint32_t *const *geometryBuffers;
    const prt_geometry_metadata *metadatas;
    const prt_material_data *materials;
    const prt_draw_instance *drawInstances;
    const prt_model_instance_transforms *transforms;
    const TexSlot *texSlots; // [textureCount * samplerCount]
    uint32_t textureCount, samplerCount;
    const prt_directional_light *directional;
    const prt_point_light *pointLights;
    const prt_spot_light *spotLights;
    uint32_t pointLightCount, spotLightCount;
    const prt_camera_uniforms *camera;
    const uint16_t *sky; // RGBA16F, 6 faces
    uint32_t skyFaceSize;
    uint32_t exactMaterialFilter; // 1: software bilinear for material taps too (parity tests)
    const AlphaMat *alphaMats;    // [draw instance]; nullptr: generic any-hit path only
};

// ---- scene/geometry.glsl ---------------------------------------------------------------------------
HD uint32_t loadIndex(const SceneView &sv, uint32_t bufferIndex, uint32_t bufferOffset, uint32_t index, uint32_t usesShort)
{ // geometry.glsl:51-59
    if (usesShort == 1) return (uint32_t)((const uint16_t *)sv.geometryBuffers[bufferIndex])[bufferOffset + index];
    return sv.geometryBuffers[bufferIndex][bufferOffset + index];
}
HD V2 unpackHalf2x16(uint32_t p) { return {halfToFloat((uint16_t)(p & 0xFFFFu)), halfToFloat((uint16_t)(p >> 16))}; }
HD V2 loadR16G16(const SceneView &sv, uint32_t bufferIndex, uint32_t bufferOffset, uint32_t index)
{ // geometry.glsl:71-80
    if (bufferOffset == 0xFFFFFFFFu) return {0.f, 0.f};
    return unpackHalf2x16(sv.geometryBuffers[bufferIndex][bufferOffset + index]);
}
HD V3 loadR16G16B16A16(const SceneView &sv, uint32_t bufferIndex, uint32_t bufferOffset, uint32_t index)
{ // geometry.glsl:82-93
    if (bufferOffset == 0xFFFFFFFFu) return {0.f, 0.f, 0.f};
    const uint32_t *b = sv.geometryBuffers[bufferIndex] + bufferOffset + index * 2;
    const V2 a = unpackHalf2x16(b[0]), c = unpackHalf2x16(b[1]);
    return {a.x, a.y, c.x};
}
HD V3 unpackSnormR10G10B10(uint32_t packed)
{ // geometry.glsl:95-103
    const int32_t x = (int32_t)(packed << 22) >> 22;
    const int32_t y = (int32_t)(packed << 12) >> 22;
    const int32_t z = (int32_t)(packed << 2) >> 22;
    const V3 v = vmax(V3{(float)x, (float)y, (float)z} / 511.f, V3{-1.f, -1.f, -1.f});
    return normalize(v);
}
HD V3 loadR10G10B10Snorm(const SceneView &sv, uint32_t bufferIndex, uint32_t bufferOffset, uint32_t index)
{ // geometry.glsl:105-114
    if (bufferOffset == 0xFFFFFFFFu) return {0.f, 0.f, 0.f};
    return unpackSnormR10G10B10(sv.geometryBuffers[bufferIndex][bufferOffset + index]);
}
HD V4 loadTangentWithSign(const SceneView &sv, uint32_t bufferIndex, uint32_t bufferOffset, uint32_t index)
{ // geometry.glsl:116-127
    if (bufferOffset == 0xFFFFFFFFu) return {0.f, 0.f, 0.f, 0.f};
    const uint32_t packed = sv.geometryBuffers[bufferIndex][bufferOffset + index];
    const V3 t = unpackSnormR10G10B10(packed);
    return {t.x, t.y, t.z, (float)((int32_t)packed >> 30)};
}

struct Vertex // vertex.glsl:4-10
{
    V3 Position;
    V3 Normal;
    V4 Tangent;
    V2 TexCoord0;
};

HD Vertex loadVertex(const SceneView &sv, const prt_geometry_metadata &m, uint32_t index)
{ // geometry.glsl:220-234
    Vertex r;
    r.Position = loadR16G16B16A16(sv, m.bufferIndex, m.positionsOffset, index);
    r.Normal = loadR10G10B10Snorm(sv, m.bufferIndex, m.normalsOffset, index);
    r.Tangent = loadTangentWithSign(sv, m.bufferIndex, m.tangentsOffset, index);
    r.TexCoord0 = loadR16G16(sv, m.bufferIndex, m.texCoord0sOffset, index);
    return r;
}
HD Vertex loadVertexThroughIndexBuffer(const SceneView &sv, const prt_geometry_metadata &m, uint32_t index)
{ // geometry.glsl:236-244
    return loadVertex(sv, m, loadIndex(sv, m.bufferIndex, m.indicesOffset, index, m.usesShortIndices));
}
HD V2 loadUV(const SceneView &sv, uint32_t meshIndex, uint32_t index)
{ // geometry.glsl:246-256
    const prt_geometry_metadata &m = sv.metadatas[meshIndex];
    const uint32_t vi = loadIndex(sv, m.bufferIndex, m.indicesOffset, index, m.usesShortIndices);
    return loadR16G16(sv, m.bufferIndex, m.texCoord0sOffset, vi);
}

HD V2 bary2(V2 v0, V2 v1, V2 v2, float a, float b, float c)
{
    return {v0.x * a + v1.x * b + v2.x * c, v0.y * a + v1.y * b + v2.y * c};
}
HD V3 bary3(V3 v0, V3 v1, V3 v2, float a, float b, float c) { return v0 * a + v1 * b + v2 * c; }

// The any-hit uv has to be bit-identical with the oracle's: single roundings, left to right.
HD V2 interpolateUvExact(V2 v0, V2 v1, V2 v2, float bx, float by)
{ // geometry.glsl:288-295
    const float a = xsub(xsub(1.f, bx), by), b = bx, c = by;
    return {xadd(xadd(xmul(v0.x, a), xmul(v1.x, b)), xmul(v2.x, c)),
            xadd(xadd(xmul(v0.y, a), xmul(v1.y, b)), xmul(v2.y, c))};
}

HD Vertex interpolateVertex(const Vertex &v0, const Vertex &v1, const Vertex &v2, float bx, float by)
{ // geometry.glsl:271-286
    const float a = 1 - bx - by, b = bx, c = by;
    Vertex r;
    r.Position = bary3(v0.Position, v1.Position, v2.Position, a, b, c);
    r.Normal = bary3(v0.Normal, v1.Normal, v2.Normal, a, b, c);
    r.Tangent = {v0.Tangent.x * a + v1.Tangent.x * b + v2.Tangent.x * c, v0.Tangent.y * a + v1.Tangent.y * b + v2.Tangent.y * c,
                 v0.Tangent.z * a + v1.Tangent.z * b + v2.Tangent.z * c, v0.Tangent.w * a + v1.Tangent.w * b + v2.Tangent.w * c};
    r.TexCoord0 = bary2(v0.TexCoord0, v1.TexCoord0, v2.TexCoord0, a, b, c);
    return r;
}

// scene/instances.glsl:36-53. GLSL `vec * mat`: result[i] = dot(vec, column i).
HD Vertex transformVertex(const Vertex &v, const prt_model_instance_transforms &t)
{
    Vertex r;
    const float(*m)[4] = t.modelToWorld;
    const float(*nm)[4] = t.normalToWorld;
    const V3 p = v.Position;
    r.Position = {p.x * m[0][0] + p.y * m[0][1] + p.z * m[0][2] + m[0][3], p.x * m[1][0] + p.y * m[1][1] + p.z * m[1][2] + m[1][3],
                  p.x * m[2][0] + p.y * m[2][1] + p.z * m[2][2] + m[2][3]};
    const V3 n = v.Normal;
    r.Normal = normalize(V3{n.x * nm[0][0] + n.y * nm[0][1] + n.z * nm[0][2], n.x * nm[1][0] + n.y * nm[1][1] + n.z * nm[1][2],
                            n.x * nm[2][0] + n.y * nm[2][1] + n.z * nm[2][2]});
    if (v.Tangent.w != 0)
    {
        const V3 g = {v.Tangent.x, v.Tangent.y, v.Tangent.z};
        const V3 tw = normalize(V3{g.x * m[0][0] + g.y * m[0][1] + g.z * m[0][2], g.x * m[1][0] + g.y * m[1][1] + g.z * m[1][2],
                                   g.x * m[2][0] + g.y * m[2][1] + g.z * m[2][2]});
        r.Tangent = {tw.x, tw.y, tw.z, v.Tangent.w};
    }
    else
        r.Tangent = v.Tangent;
    r.TexCoord0 = v.TexCoord0;
    return r;
}

// ---- texture fetch: texture() in a stage without derivatives = LOD 0, magnification filter
// (materials.glsl:44; WorldData.cpp:686-720) ----------------------------------------------------------
HD int wrapCoord(int i, int n, uint32_t mode)
{
    if (mode == PRT_ADDRESS_REPEAT)
    {
        if ((n & (n - 1)) == 0) return i & (n - 1); // power of two: the same non-negative remainder
        i %= n;
        return i < 0 ? i + n : i;
    }
    if (mode == PRT_ADDRESS_MIRRORED_REPEAT)
    {
        const int period = 2 * n;
        i %= period;
        if (i < 0) i += period;
        return i < n ? i : period - 1 - i;
    }
    return i < 0 ? 0 : (i > n - 1 ? n - 1 : i);
}

// texel (x, y) of level 0, decoded in software from the cached payload: unorm8 / 255 like the oracle
HD V4 fetchTexel(const TexSlot &s, int x, int y)
{
    uint32_t px;
    if (s.format == PRT_TEX_BC7_UNORM)
    {
        const uint32_t bw = (s.width + 3u) >> 2;
        const uint8_t *blk = s.payload + ((size_t)((uint32_t)y >> 2) * bw + ((uint32_t)x >> 2)) * 16;
#if defined(__CUDA_ARCH__)
        const uint4 b = __ldg(reinterpret_cast<const uint4 *>(blk));
        const uint64_t lo = (uint64_t)b.x | ((uint64_t)b.y << 32), hi = (uint64_t)b.z | ((uint64_t)b.w << 32);
#else
        uint64_t lo, hi;
        memcpy(&lo, blk, 8);
        memcpy(&hi, blk + 8, 8);
#endif
        px = bc7DecodeTexel(lo, hi, (((uint32_t)y & 3u) << 2) | ((uint32_t)x & 3u));
    }
    else
    {
#if defined(__CUDA_ARCH__)
        px = __ldg(reinterpret_cast<const uint32_t *>(s.payload) + (size_t)y * s.width + x);
#else
        memcpy(&px, s.payload + ((size_t)y * s.width + x) * 4, 4);
#endif
    }
    return {xdiv((float)(px & 255u), 255.f), xdiv((float)((px >> 8) & 255u), 255.f), xdiv((float)((px >> 16) & 255u), 255.f),
            xdiv((float)(px >> 24), 255.f)};
}

// Software bilinear with the oracle's exact operation order (oracle.cpp sampleTexture).
HD V4 sampleTextureExact(const TexSlot &s, V2 uv)
{
    const int w = (int)s.width, h = (int)s.height;
    if (s.magFilter == PRT_FILTER_NEAREST)
    {
        const int x = wrapCoord((int)floorf(xmul(uv.x, (float)w)), w, s.addressU);
        const int y = wrapCoord((int)floorf(xmul(uv.y, (float)h)), h, s.addressV);
        return fetchTexel(s, x, y);
    }
    const float fx = xsub(xmul(uv.x, (float)w), 0.5f), fy = xsub(xmul(uv.y, (float)h), 0.5f);
    const float x0f = floorf(fx), y0f = floorf(fy);
    const float ax = xsub(fx, x0f), ay = xsub(fy, y0f);
    const int x0 = wrapCoord((int)x0f, w, s.addressU), x1 = wrapCoord((int)x0f + 1, w, s.addressU);
    const int y0 = wrapCoord((int)y0f, h, s.addressV), y1 = wrapCoord((int)y0f + 1, h, s.addressV);
    const V4 t00 = fetchTexel(s, x0, y0), t10 = fetchTexel(s, x1, y0);
    const V4 t01 = fetchTexel(s, x0, y1), t11 = fetchTexel(s, x1, y1);
    const float bx = xsub(1.f, ax), by = xsub(1.f, ay);
    auto lerp2 = [&](float c00, float c10, float c01, float c11) {
        const float top = xadd(xmul(c00, bx), xmul(c10, ax));
        const float bot = xadd(xmul(c01, bx), xmul(c11, ax));
        return xadd(xmul(top, by), xmul(bot, ay));
    };
    return {lerp2(t00.x, t10.x, t01.x, t11.x), lerp2(t00.y, t10.y, t01.y, t11.y), lerp2(t00.z, t10.z, t01.z, t11.z),
            lerp2(t00.w, t10.w, t01.w, t11.w)};
}

// .w of sampleTextureExact, bit for bit: the same coordinates, texel values and filter arithmetic, but the four
// texels come from the decoded alpha plane (or are the constant 1 for a texture without transparency) instead
// of four software BC7 block decodes. This is what scene.rahit's sampleAlpha costs per candidate.
// `fallback` (may be nullptr) decodes the payload when no plane was built.
HD float texelAlphaAt(const uint8_t *plane, const TexSlot *fallback, uint32_t width, int x, int y)
{
    if (plane == nullptr) return fallback ? fetchTexel(*fallback, x, y).w : 1.f;
#if defined(__CUDA_ARCH__)
    const uint32_t a = __ldg(plane + (size_t)y * width + x);
#else
    const uint32_t a = plane[(size_t)y * width + x];
#endif
    // 0 / 255 and 255 / 255 without the division (its zero-dividend case is a slow path on the device)
    return a == 0u ? 0.f : (a == 255u ? 1.f : xdiv((float)a, 255.f));
}

HD float alphaFilterExact(const uint8_t *plane, bool allOpaque, const TexSlot *fallback, uint32_t width, uint32_t height, bool nearest,
                          uint32_t addressU, uint32_t addressV, V2 uv)
{
    const int w = (int)width, h = (int)height;
    if (nearest)
    {
        if (allOpaque) return 1.f; // xdiv(255, 255)
        const int x = wrapCoord((int)floorf(xmul(uv.x, (float)w)), w, addressU);
        const int y = wrapCoord((int)floorf(xmul(uv.y, (float)h)), h, addressV);
        return texelAlphaAt(plane, fallback, width, x, y);
    }
    const float fx = xsub(xmul(uv.x, (float)w), 0.5f), fy = xsub(xmul(uv.y, (float)h), 0.5f);
    const float x0f = floorf(fx), y0f = floorf(fy);
    const float ax = xsub(fx, x0f), ay = xsub(fy, y0f);
    float c00 = 1.f, c10 = 1.f, c01 = 1.f, c11 = 1.f;
    if (!allOpaque)
    {
        const int x0 = wrapCoord((int)x0f, w, addressU), x1 = wrapCoord((int)x0f + 1, w, addressU);
        const int y0 = wrapCoord((int)y0f, h, addressV), y1 = wrapCoord((int)y0f + 1, h, addressV);
        c00 = texelAlphaAt(plane, fallback, width, x0, y0); c10 = texelAlphaAt(plane, fallback, width, x1, y0);
        c01 = texelAlphaAt(plane, fallback, width, x0, y1); c11 = texelAlphaAt(plane, fallback, width, x1, y1);
    }
    const float bx = xsub(1.f, ax), by = xsub(1.f, ay);
    const float top = xadd(xmul(c00, bx), xmul(c10, ax));
    const float bot = xadd(xmul(c01, bx), xmul(c11, ax));
    return xadd(xmul(top, by), xmul(bot, ay));
}

HD float sampleTextureAlphaExact(const TexSlot &s, V2 uv)
{
    return alphaFilterExact(s.alpha, s.alphaOpaque != 0, &s, s.width, s.height, s.magFilter == PRT_FILTER_NEAREST, s.addressU, s.addressV, uv);
}

HD const TexSlot *texSlot(const SceneView &sv, uint32_t textureSampler)
{
    const uint32_t tex = textureSampler & 0xFFFFFFu, smp = textureSampler >> 24;
    if (tex >= sv.textureCount) return nullptr;
    const TexSlot *s = &sv.texSlots[(size_t)tex * sv.samplerCount + (smp < sv.samplerCount ? smp : 0)];
    return s->valid ? s : nullptr;
}

// texture() through the software filter
HD V4 sampleTexture(const SceneView &sv, uint32_t textureSampler, V2 uv)
{
    const TexSlot *s = texSlot(sv, textureSampler);
    if (s == nullptr) return {1.f, 1.f, 1.f, 1.f};
    return sampleTextureExact(*s, uv);
}

// ---- scene/materials.glsl ----------------------------------------------------------------------------
struct Material // scene/material.glsl:4-16
{
    V3 albedo;
    V3 normal;
    float roughness;
    float metallic;
    float alpha;
};

HD float sRGBtoLinear(float x) // materials.glsl:26-29
{
    return x <= 0.04045f ? x / 12.92f : powf((x + 0.055f) / 1.055f, 2.4f);
}
// The same function for the any-hit stage with its two most common arguments answered without the IEEE division's
// zero-dividend slow path and without powf: 0 / 12.92 == 0 and powf(1.055 / 1.055, 2.4) == powf(1, 2.4) == 1 exactly.
HD float sRGBtoLinearAlpha(float x)
{
    if (x == 0.f) return x;
    if (x == 1.f) return 1.f;
    return sRGBtoLinear(x);
}

// The three texture() results sampleMaterial needs (base colour, metallic-roughness, normal); a tap of
// a material without that texture is never read.
struct MaterialTaps
{
    V4 base, metallicRoughness, normal;
};

HD MaterialTaps softwareMaterialTaps(const SceneView &sv, uint32_t index, V2 uv)
{
    const prt_material_data &data = sv.materials[index];
    MaterialTaps t;
    t.base = t.metallicRoughness = t.normal = V4{1.f, 1.f, 1.f, 1.f};
    if ((data.baseColorTextureSampler & 0xFFFFFFu) > 0) t.base = sampleTexture(sv, data.baseColorTextureSampler, uv);
    if ((data.metallicRoughnessTextureSampler & 0xFFFFFFu) > 0) t.metallicRoughness = sampleTexture(sv, data.metallicRoughnessTextureSampler, uv);
    if ((data.normalTextureSampler & 0xFFFFFFu) > 0) t.normal = sampleTexture(sv, data.normalTextureSampler, uv);
    return t;
}

HD Material sampleMaterial(const SceneView &sv, uint32_t index, const MaterialTaps &taps)
{ // materials.glsl:47-119
    const prt_material_data &data = sv.materials[index];
    Material ret;
    ret.albedo = {0.f, 0.f, 0.f}; ret.normal = {0.f, 0.f, 0.f}; ret.roughness = 0.f; ret.metallic = 0.f; ret.alpha = 0.f;
    V4 c = {1.f, 1.f, 1.f, 1.f};
    if ((data.baseColorTextureSampler & 0xFFFFFFu) > 0)
    {
        const V4 s = taps.base;
        c = {sRGBtoLinear(s.x), sRGBtoLinear(s.y), sRGBtoLinear(s.z), s.w};
    }
    c = {c.x * data.baseColorFactor[0], c.y * data.baseColorFactor[1], c.z * data.baseColorFactor[2], c.w * data.baseColorFactor[3]};
    if (data.alphaMode == PRT_ALPHA_BLEND)
        ret.alpha = c.w;
    else
    {
        if (data.alphaMode == PRT_ALPHA_MASK && c.w < data.alphaCutoff)
        {
            ret.alpha = 0;
            return ret;
        }
        ret.alpha = -1;
    }
    ret.albedo = {c.x, c.y, c.z};
    if ((data.metallicRoughnessTextureSampler & 0xFFFFFFu) > 0)
    {
        ret.roughness = taps.metallicRoughness.y * data.roughnessFactor;
        ret.metallic = taps.metallicRoughness.z * data.metallicFactor;
    }
    else
    {
        ret.roughness = data.roughnessFactor;
        ret.metallic = data.metallicFactor;
    }
    ret.roughness = fmaxf_(ret.roughness, 0.05f);
    if ((data.normalTextureSampler & 0xFFFFFFu) > 0)
        ret.normal = V3{taps.normal.x, taps.normal.y, taps.normal.z} * 2.f - V3{1.f, 1.f, 1.f};
    else
        ret.normal = {-2.f, -2.f, -2.f};
    return ret;
}

// materials.glsl:121-147: the float overload of sRGBtoLinear IS applied to alpha here. Always the
// exact software filter: this feeds the accept/ignore decision of scene.rahit.
HD float sampleAlpha(const SceneView &sv, uint32_t index, V2 uv)
{
    const prt_material_data &data = sv.materials[index];
    float linearAlpha = 1;
    if ((data.baseColorTextureSampler & 0xFFFFFFu) > 0)
    {
        const TexSlot *s = texSlot(sv, data.baseColorTextureSampler);
        linearAlpha = sRGBtoLinearAlpha(s ? sampleTextureAlphaExact(*s, uv) : 1.f);
    }
    linearAlpha = xmul(linearAlpha, data.baseColorFactor[3]);
    if (data.alphaMode == PRT_ALPHA_BLEND) return linearAlpha;
    if (data.alphaMode == PRT_ALPHA_MASK && linearAlpha < data.alphaCutoff) return 0;
    return -1;
}

HD uint32_t pcg(uint32_t v) // common/random.glsl:7-12
{
    const uint32_t state = v * 747796405u + 2891336453u;
    const uint32_t word = ((state >> ((state >> 28u) + 4u)) ^ state) * 277803737u;
    return (word >> 22u) ^ word;
}

// rt/scene.rahit:18-39. One u per RAY (payload.randomSeed is not advanced between candidates).
HD bool anyHitAccept(const SceneView &sv, uint32_t drawInstanceIndex, uint32_t primitiveID, float bx, float by, uint32_t randomSeed)
{
    const prt_draw_instance &instance = sv.drawInstances[drawInstanceIndex];
    const V2 uv0 = loadUV(sv, instance.meshIndex, primitiveID * 3 + 0);
    const V2 uv1 = loadUV(sv, instance.meshIndex, primitiveID * 3 + 1);
    const V2 uv2 = loadUV(sv, instance.meshIndex, primitiveID * 3 + 2);
    const V2 uv = interpolateUvExact(uv0, uv1, uv2, bx, by);
    const float alpha = sampleAlpha(sv, instance.materialIndex, uv);
    if (alpha == 0) return false;
    if (alpha > 0)
    {
        const float u = (float)pcg(randomSeed) / 4294967296.0f;
        if (u > alpha) return false;
    }
    return true;
}

// The AlphaMat of one draw instance (k_alpha_mats; the CPU simulator calls it too).
HD AlphaMat makeAlphaMat(const SceneView &sv, uint32_t drawInstanceIndex)
{
    const prt_draw_instance &instance = sv.drawInstances[drawInstanceIndex];
    const prt_material_data &data = sv.materials[instance.materialIndex];
    AlphaMat m;
    m.alpha = nullptr; m.width = 1; m.height = 1; m.pad = 0;
    m.factor = data.baseColorFactor[3];
    m.cutoff = data.alphaCutoff;
    m.bits = ALPHAMAT_USABLE | (data.alphaMode << ALPHAMAT_MODE_SHIFT);
    if ((data.baseColorTextureSampler & 0xFFFFFFu) > 0)
    {
        const TexSlot *s = texSlot(sv, data.baseColorTextureSampler);
        if (s != nullptr) // an unbound slot samples as 1 and sRGBtoLinear(1) == 1: the same as no texture
        {
            m.bits |= ALPHAMAT_TEXTURED | (s->magFilter == PRT_FILTER_NEAREST ? ALPHAMAT_NEAREST : 0u) |
                      (s->addressU << ALPHAMAT_ADDRU_SHIFT) | (s->addressV << ALPHAMAT_ADDRV_SHIFT);
            m.width = s->width; m.height = s->height;
            if (s->alphaOpaque) m.bits |= ALPHAMAT_TEX_OPAQUE;
            else if (s->alpha) m.alpha = s->alpha;
            else m.bits &= ~ALPHAMAT_USABLE; // no alpha plane built: generic path decodes the payload
        }
    }
    return m;
}

// anyHitAccept with the flattened records: bit-identical decisions (same uv interpolation, same filter, same
// sRGB->linear, same comparisons), two 128-bit loads + four alpha bytes instead of the scene-buffer chase.
HD bool anyHitAcceptFast(const AlphaMat &m, const AlphaUv &t, float bx, float by, uint32_t randomSeed)
{
    const V2 uv = interpolateUvExact(V2{t.uv[0], t.uv[1]}, V2{t.uv[2], t.uv[3]}, V2{t.uv[4], t.uv[5]}, bx, by);
    float linearAlpha = 1;
    if (m.bits & ALPHAMAT_TEXTURED)
        linearAlpha = sRGBtoLinearAlpha(alphaFilterExact(m.alpha, (m.bits & ALPHAMAT_TEX_OPAQUE) != 0, nullptr, m.width, m.height,
                                                    (m.bits & ALPHAMAT_NEAREST) != 0, (m.bits >> ALPHAMAT_ADDRU_SHIFT) & 3u,
                                                    (m.bits >> ALPHAMAT_ADDRV_SHIFT) & 3u, uv));
    linearAlpha = xmul(linearAlpha, m.factor);
    const uint32_t mode = (m.bits >> ALPHAMAT_MODE_SHIFT) & 3u;
    if (mode == PRT_ALPHA_BLEND)
    {
        if (linearAlpha == 0) return false;
        if (linearAlpha > 0)
        {
            const float u = (float)pcg(randomSeed) / 4294967296.0f;
            if (u > linearAlpha) return false;
        }
        return true;
    }
    if (mode == PRT_ALPHA_MASK && linearAlpha < m.cutoff) return false;
    return true;
}

} // namespace prt
